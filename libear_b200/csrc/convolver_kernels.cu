// K4: bank of uniformly-partitioned overlap-add FFT convolvers.
//
// Replaces BlockConvolver::process (ref src/dsp/block_convolver_impl.cpp:143-237) and
// the r2c / c2r transforms of the FFT plug-in it calls (ref src/fft_kiss.cpp:61-88):
// per block of B samples, zero-pad to 2B, r2c, multiply-accumulate the last
// `num_blocks` input spectra with the filter partitions (two spectra per block while
// a crossfade is in flight), c2r, add the previous block's tail, scale by 1/(2B).
//
// libear processes one block of one channel per call and carries spectra queues and
// `last_tail` between calls.  Here one CTA processes a *run* of R consecutive blocks
// of one channel; everything a run needs from before its first block (the num_blocks
// previous input blocks and the previous tail) is recomputed from time-domain input
// ("halo"), so all runs of all channels are independent and one launch covers the
// whole call.  State carried between calls: the last num_blocks input blocks and the
// last tail, per channel.
//
// The 2B-point real transforms are done as B-point complex Stockham FFTs in shared
// memory (radix 4, plus one radix-2 pass when log2 B is odd) with the usual even/odd
// split; twiddles come from a table computed in double precision on the host.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>

#include "kernels.hpp"

namespace earb {
  namespace {
    constexpr int kConvThreads = CONV_THREADS;

    __device__ __forceinline__ float2 cmul(float2 a, float2 b) {
      return make_float2(fmaf(a.x, b.x, -a.y * b.y), fmaf(a.x, b.y, a.y * b.x));
    }
    __device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
    __device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
    __device__ __forceinline__ float2 conjf2(float2 a) { return make_float2(a.x, -a.y); }

    // tw2 holds exp(-2 pi i k / (2B)) for k in [0, 2B).  Twiddle exp(sign 2 pi i t / B).
    template <bool INVERSE>
    __device__ __forceinline__ float2 twiddle(const float2 *__restrict__ tw2, int t) {
      float2 w = __ldg(tw2 + 2 * t);
      if (INVERSE) w.y = -w.y;
      return w;
    }

    // B-point complex FFT, Stockham autosort: natural order in and out.  Data ping-pongs
    // between `a` and `b`; returns the buffer holding the result.  All threads of the CTA
    // take part; contains __syncthreads().
    template <bool INVERSE>
    __device__ float2 *fft_stockham(float2 *a, float2 *b, int B, const float2 *__restrict__ tw2) {
      float2 *src = a, *dst = b;
      int Ns = 1;
      // radix-4 passes
      while (Ns * 4 <= B) {
        const int quarter = B >> 2;
        for (int j = threadIdx.x; j < quarter; j += kConvThreads) {
          const int k = j & (Ns - 1);
          const int tstep = k * (B / (Ns * 4));  // twiddle index step: k * B / (Ns*R)
          float2 v0 = src[j];
          float2 v1 = cmul(src[j + quarter], twiddle<INVERSE>(tw2, tstep));
          float2 v2 = cmul(src[j + 2 * quarter], twiddle<INVERSE>(tw2, 2 * tstep));
          float2 v3 = cmul(src[j + 3 * quarter], twiddle<INVERSE>(tw2, 3 * tstep));
          float2 s02 = cadd(v0, v2), d02 = csub(v0, v2);
          float2 s13 = cadd(v1, v3), d13 = csub(v1, v3);
          // multiply d13 by -i (forward) or +i (inverse)
          float2 r13 = INVERSE ? make_float2(-d13.y, d13.x) : make_float2(d13.y, -d13.x);
          const int base = ((j - k) << 2) + k;  // (j / Ns) * Ns * 4 + k
          dst[base] = cadd(s02, s13);
          dst[base + Ns] = cadd(d02, r13);
          dst[base + 2 * Ns] = csub(s02, s13);
          dst[base + 3 * Ns] = csub(d02, r13);
        }
        __syncthreads();
        float2 *t = src;
        src = dst;
        dst = t;
        Ns <<= 2;
      }
      if (Ns < B) {  // one radix-2 pass
        const int half = B >> 1;
        for (int j = threadIdx.x; j < half; j += kConvThreads) {
          const int k = j & (Ns - 1);
          float2 v0 = src[j];
          float2 v1 = cmul(src[j + half], twiddle<INVERSE>(tw2, k * (B / (Ns * 2))));
          const int base = ((j - k) << 1) + k;
          dst[base] = cadd(v0, v1);
          dst[base + Ns] = csub(v0, v1);
        }
        __syncthreads();
        float2 *t = src;
        src = dst;
        dst = t;
      }
      return src;
    }

    // spectrum X[0..B] (B+1 bins) of the zero-padded 2B-point real signal whose first B
    // samples are `x` (scaled per sample by fade(n)); za/zb: B-complex scratch.
    // FADE: 0 none, 1 down (1 - n/B), 2 up (n/B)   (ref block_convolver_impl.cpp:127-141)
    // n_real: how many of the 2B real samples are present (B for the convolver's zero-padded
    // blocks, 2B for the plug-in transform)
    __device__ void r2c_block(const float *__restrict__ x, int fade, float2 *za, float2 *zb, int B,
                              const float2 *__restrict__ tw2, float2 *X, int n_real) {
      const float i_scale = 1.0f / (float)B;
      for (int n = threadIdx.x; n < B; n += kConvThreads) {
        float2 v = make_float2(0.0f, 0.0f);
        if (2 * n < n_real) {
          float e = x ? x[2 * n] : 0.0f, o = x ? x[2 * n + 1] : 0.0f;
          if (fade) {
            float ae = (float)(2 * n) * i_scale, ao = (float)(2 * n + 1) * i_scale;
            if (fade == 1) {
              ae = 1.0f - ae;
              ao = 1.0f - ao;
            }
            e = ae * e;
            o = ao * o;
          }
          v = make_float2(e, o);
        }
        za[n] = v;
      }
      __syncthreads();
      const float2 *Z = fft_stockham<false>(za, zb, B, tw2);
      for (int k = threadIdx.x; k <= B; k += kConvThreads) {
        const float2 zk = Z[k & (B - 1)];
        const float2 zc = conjf2(Z[(B - k) & (B - 1)]);
        const float2 E = make_float2(0.5f * (zk.x + zc.x), 0.5f * (zk.y + zc.y));
        const float2 D = csub(zk, zc);                          // 2i * O
        const float2 O = make_float2(0.5f * D.y, -0.5f * D.x);  // D / (2i)
        X[k] = cadd(E, cmul(__ldg(tw2 + k), O));
      }
      __syncthreads();
    }

    struct ConvArgs {
      int B, nb, C, R, Kg;
      long K;
      const float *in;
      long in_stride;
      float *out;
      long out_stride;
      const float *hist;   // [C][nb][B] the nb input blocks before the call (oldest first)
      const float *tail;   // [C][B] last tail of the previous call (read by run 0)
      float *tail_new;     // [C][B] last tail of this call (written by the final run)
      int vec_out;         // out rows are 8-byte aligned
      const float2 *filt;  // [n_filters][nb][B+1]
      // filter ids for the first Kg blocks of the call (while crossfades / filter
      // changes are still travelling through the queue): [C][Kg][nb][2] = (old, new)
      const short *ftab;
      const int *uniform;            // [C] filter id from block Kg on (-1: none)
      const unsigned char *faded;    // [C][nb + Kg] input block was split into down/up
      const float2 *tw2;
    };

    __global__ void __launch_bounds__(kConvThreads) convolver_kernel(ConvArgs a) {
      extern __shared__ float2 smem2[];
      const int B = a.B, nb = a.nb;
      const int c = blockIdx.y;
      const long k0 = (long)blockIdx.x * a.R;
      const long k1 = min(a.K, k0 + a.R);
      float2 *za = smem2;               // B
      float2 *zb = za + B;              // B
      float2 *Y = zb + B;               // B+1 (+ pad)
      float2 *ring = Y + (B + 2);       // [nb][2][B+2]  (old, new) spectra, slot = block mod nb
      float *tail = reinterpret_cast<float *>(ring + (size_t)nb * 2 * (B + 2));  // B
      __shared__ unsigned char ring_faded[64];

      const float norm = 1.0f / (float)(2 * B);
      const float *hist = a.hist + (size_t)c * nb * B;

      if (k0 == 0) {
        for (int n = threadIdx.x; n < B; n += kConvThreads) tail[n] = a.tail[(size_t)c * B + n];
      }
      __syncthreads();

      for (long j = k0 - nb; j < k1; j++) {
        // ---- input spectra of block j
        const float *x;
        if (j < 0)
          x = hist + (size_t)(j + nb) * B;
        else
          x = a.in ? a.in + (size_t)c * a.in_stride + (size_t)j * B : nullptr;
        const bool is_faded = (j < a.Kg) ? a.faded[(size_t)c * (nb + a.Kg) + (j + nb)] != 0 : false;
        const int slot = (int)((j + nb) % nb);
        float2 *Xold = ring + (size_t)(slot * 2) * (B + 2);
        float2 *Xnew = Xold + (B + 2);
        if (is_faded) {
          r2c_block(x, 1, za, zb, B, a.tw2, Xold, B);
          r2c_block(x, 2, za, zb, B, a.tw2, Xnew, B);
        } else {
          r2c_block(x, 0, za, zb, B, a.tw2, Xnew, B);
        }
        if (threadIdx.x == 0) ring_faded[slot] = is_faded;
        __syncthreads();
        if (j < k0 - 1) continue;  // halo blocks before k0-1 only feed the ring

        // ---- Y_j = sum_i Fold(j,i)[i] * Xold_{j-i} + Fnew(j,i)[i] * Xnew_{j-i}
        if (j < 0) {
          // block before the call: its tail is carried in a.tail (loaded above)
          continue;
        }
        for (int k = threadIdx.x; k <= B; k += kConvThreads) {
          float2 acc = make_float2(0.0f, 0.0f);
          for (int i = 0; i < nb; i++) {
            const long src = j - i;
            if (src < -nb) break;
            int f_old, f_new;
            if (j < a.Kg) {
              const short *ft = a.ftab + (((size_t)c * a.Kg + j) * nb + i) * 2;
              f_old = ft[0];
              f_new = ft[1];
            } else {
              f_old = f_new = a.uniform[c];
            }
            const int s = (int)((src + nb) % nb);
            const float2 *Xo = ring + (size_t)(s * 2) * (B + 2);
            const float2 *Xn = Xo + (B + 2);
            if (f_old >= 0 && ring_faded[s])
              acc = cadd(acc, cmul(__ldg(a.filt + ((size_t)f_old * nb + i) * (B + 1) + k), Xo[k]));
            if (f_new >= 0)
              acc = cadd(acc, cmul(__ldg(a.filt + ((size_t)f_new * nb + i) * (B + 1) + k), Xn[k]));
          }
          Y[k] = acc;
        }
        __syncthreads();

        // ---- c2r: Z[k] = (Y[k] + conj Y[B-k]) + i (Y[k] - conj Y[B-k]) e^{+2 pi i k / 2B}
        for (int k = threadIdx.x; k < B; k += kConvThreads) {
          const float2 yk = Y[k];
          const float2 yc = conjf2(Y[B - k]);
          const float2 E = cadd(yk, yc);
          float2 w = __ldg(a.tw2 + k);
          w.y = -w.y;
          const float2 O = cmul(csub(yk, yc), w);
          za[k] = make_float2(E.x - O.y, E.y + O.x);  // E + i O
        }
        __syncthreads();
        const float2 *z = fft_stockham<true>(za, zb, B, a.tw2);
        // z[n] = (y[2n], y[2n+1]); first half of y -> output (+ tail), second half -> tail
        if (j >= k0) {
          float *o = a.out + (size_t)c * a.out_stride + (size_t)j * B;
          for (int n = threadIdx.x; n < B / 2; n += kConvThreads) {
            const float2 v = z[n];
            const float2 r = make_float2((v.x + tail[2 * n]) * norm, (v.y + tail[2 * n + 1]) * norm);
            if (a.vec_out) {
              *reinterpret_cast<float2 *>(o + 2 * n) = r;
            } else {
              o[2 * n] = r.x;
              o[2 * n + 1] = r.y;
            }
          }
        }
        __syncthreads();
        for (int n = threadIdx.x; n < B / 2; n += kConvThreads) {
          const float2 v = z[n + B / 2];
          tail[2 * n] = v.x;
          tail[2 * n + 1] = v.y;
        }
        __syncthreads();
      }
      if (k1 == a.K) {
        for (int n = threadIdx.x; n < B; n += kConvThreads) a.tail_new[(size_t)c * B + n] = tail[n];
      }
    }

    // ---- the FFT plug-in (ref include/ear/fft.hpp:27-62, src/fft_kiss.cpp:61-88) as a batched
    // transform: one CTA per 2B-point real transform, same butterflies as the convolver
    __global__ void __launch_bounds__(kConvThreads)
        fft_r2c_kernel(int B, long batch, const float *__restrict__ in, float2 *__restrict__ out,
                       const float2 *__restrict__ tw2) {
      extern __shared__ float2 smem2[];
      float2 *za = smem2, *zb = za + B, *X = zb + B;
      for (long t = blockIdx.x; t < batch; t += gridDim.x) {
        r2c_block(in + (size_t)t * 2 * B, 0, za, zb, B, tw2, X, 2 * B);
        for (int k = threadIdx.x; k <= B; k += kConvThreads) out[(size_t)t * (B + 1) + k] = X[k];
        __syncthreads();
      }
    }

    // unnormalised inverse: B+1 bins -> 2B reals (the imaginary parts of bins 0 and B are ignored,
    // as a c2r transform does)
    __global__ void __launch_bounds__(kConvThreads)
        fft_c2r_kernel(int B, long batch, const float2 *__restrict__ in, float *__restrict__ out,
                       const float2 *__restrict__ tw2) {
      extern __shared__ float2 smem2[];
      float2 *za = smem2, *zb = za + B, *Y = zb + B;
      for (long t = blockIdx.x; t < batch; t += gridDim.x) {
        for (int k = threadIdx.x; k <= B; k += kConvThreads) {
          float2 v = in[(size_t)t * (B + 1) + k];
          if (k == 0 || k == B) v.y = 0.0f;
          Y[k] = v;
        }
        __syncthreads();
        for (int k = threadIdx.x; k < B; k += kConvThreads) {
          const float2 yk = Y[k];
          const float2 yc = conjf2(Y[B - k]);
          const float2 E = cadd(yk, yc);
          float2 w = __ldg(tw2 + k);
          w.y = -w.y;
          const float2 O = cmul(csub(yk, yc), w);
          za[k] = make_float2(E.x - O.y, E.y + O.x);
        }
        __syncthreads();
        const float2 *z = fft_stockham<true>(za, zb, B, tw2);
        float2 *o = reinterpret_cast<float2 *>(out + (size_t)t * 2 * B);
        for (int n = threadIdx.x; n < B; n += kConvThreads) o[n] = z[n];
        __syncthreads();
      }
    }

    // new history = the last nb input blocks seen (old history shifted when K < nb)
    __global__ void conv_history_kernel(int B, int nb, long K, const float *in, long in_stride,
                                        const float *hist_old, float *hist_new) {
      const int c = blockIdx.y;
      const long total = (long)nb * B;
      for (long idx = (long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
           idx += (long)gridDim.x * blockDim.x) {
        const long slot = idx / B, n = idx % B;
        const long j = K - nb + slot;  // block index relative to this call
        float v;
        if (j >= 0)
          v = in ? in[(size_t)c * in_stride + (size_t)j * B + n] : 0.0f;
        else
          v = hist_old[(size_t)c * total + (size_t)(j + nb) * B + n];
        hist_new[(size_t)c * total + idx] = v;
      }
    }
    // libear's last_tail after the last block of a call, from the time-domain state: with a constant
    // filter h (nb partitions of B taps) and the last nb input blocks x_{K-1-i},
    //   tail[n] = sum_i (x_{K-1-i} * h_i)[B + n] = sum_i sum_{m > n} x_{K-1-i}[m] h[i B + B + n - m]
    // (the second halves of the 2B-point products of block K-1; ref block_convolver_impl.cpp:214-236),
    // unnormalised like the generic kernel's (x 2B).  One thread per (channel, n), double accumulator.
    __global__ void conv_generic_tail_kernel(int B, int nb, const float *__restrict__ hist_new,
                                             const float *__restrict__ taps, const int *__restrict__ filter_of,
                                             float *__restrict__ tail_new) {
      const int c = blockIdx.y;
      const int n = blockIdx.x * blockDim.x + threadIdx.x;
      if (n >= B) return;
      const int f = filter_of[c];
      double acc = 0.0;
      if (f >= 0) {
        const float *h = taps + (size_t)f * nb * B;
        const float *x = hist_new + (size_t)c * nb * B;
        for (int i = 0; i < nb; i++) {
          const float *xb = x + (size_t)(nb - 1 - i) * B;  // block K-1-i (history is oldest first)
          const float *hb = h + (size_t)i * B + B + n;     // hb[-m] = h[i B + B + n - m]
          for (int m = n + 1; m < B; m++) acc += (double)xb[m] * (double)hb[-m];
        }
      }
      tail_new[(size_t)c * B + n] = (float)(acc * (double)(2 * B));
    }
  }  // namespace

  void launch_conv_generic_tail(const ConvLaunch &L, const int *d_filter_of, cudaStream_t stream) {
    const int threads = std::min(L.B, 128);
    dim3 grid((unsigned)((L.B + threads - 1) / threads), (unsigned)L.C);
    conv_generic_tail_kernel<<<grid, threads, 0, stream>>>(L.B, L.nb, L.hist_new, L.taps, d_filter_of, L.tail_new);
  }

  size_t conv_smem_bytes(int B, int nb) {
    return sizeof(float2) * ((size_t)2 * B + (B + 2) + (size_t)nb * 2 * (B + 2)) + sizeof(float) * B;
  }

  void launch_fft_r2c(int B, long batch, const float *in, float *out, const float *tw2, int n_sms,
                      cudaStream_t stream) {
    if (batch <= 0) return;
    const size_t smem = sizeof(float2) * ((size_t)3 * B + 2);
    if (smem > 48 * 1024)
      cudaFuncSetAttribute(fft_r2c_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const unsigned grid = (unsigned)std::min<long>(batch, (long)n_sms * 8);
    fft_r2c_kernel<<<grid, kConvThreads, smem, stream>>>(B, batch, in, reinterpret_cast<float2 *>(out),
                                                        reinterpret_cast<const float2 *>(tw2));
  }

  void launch_fft_c2r(int B, long batch, const float *in, float *out, const float *tw2, int n_sms,
                      cudaStream_t stream) {
    if (batch <= 0) return;
    const size_t smem = sizeof(float2) * ((size_t)3 * B + 2);
    if (smem > 48 * 1024)
      cudaFuncSetAttribute(fft_c2r_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const unsigned grid = (unsigned)std::min<long>(batch, (long)n_sms * 8);
    fft_c2r_kernel<<<grid, kConvThreads, smem, stream>>>(B, batch, reinterpret_cast<const float2 *>(in), out,
                                                        reinterpret_cast<const float2 *>(tw2));
  }

  void launch_convolver(const ConvLaunch &L, cudaStream_t stream) {
    if (L.K <= 0) return;
    ConvArgs a;
    a.B = L.B;
    a.nb = L.nb;
    a.C = L.C;
    a.R = L.R;
    a.Kg = L.Kg;
    a.K = L.K;
    a.in = L.in;
    a.in_stride = L.in_stride;
    a.out = L.out;
    a.out_stride = L.out_stride;
    a.hist = L.hist;
    a.tail = L.tail;
    a.tail_new = L.tail_new;
    a.vec_out = ((reinterpret_cast<uintptr_t>(L.out) & 7) == 0) && (L.out_stride % 2 == 0);
    a.filt = reinterpret_cast<const float2 *>(L.filt);
    a.ftab = L.ftab;
    a.uniform = L.uniform;
    a.faded = L.faded;
    a.tw2 = reinterpret_cast<const float2 *>(L.tw2);
    const size_t smem = conv_smem_bytes(L.B, L.nb);
    if (smem > 48 * 1024)
      cudaFuncSetAttribute(convolver_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    dim3 grid((unsigned)((L.K + L.R - 1) / L.R), (unsigned)L.C);
    convolver_kernel<<<grid, kConvThreads, smem, stream>>>(a);
    launch_conv_history(L, stream);
  }

  void launch_conv_history(const ConvLaunch &L, cudaStream_t stream) {
    dim3 hgrid((unsigned)std::min<long>(((long)L.nb * L.B + 255) / 256, 64), (unsigned)L.C);
    conv_history_kernel<<<hgrid, 256, 0, stream>>>(L.B, L.nb, L.K, L.in, L.in_stride, L.hist, L.hist_new);
  }
}  // namespace earb
