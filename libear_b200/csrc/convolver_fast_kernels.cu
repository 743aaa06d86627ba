// K4 fast path: convolver bank whose filters have at most 512 taps (B * num_blocks <= 512), no filter
// change in flight -- the decorrelator case of the Objects signal flow (BASELINE config 4: B = 512,
// one partition) and every finer partitioning of the same length (256 x 2, 128 x 4, 64 x 8 ...).
// With a constant filter a uniformly partitioned convolver IS the linear convolution with the whole
// filter, whatever the partition size, so those shapes run as ONE 512-tap partition over
// super-blocks of 512 samples; what they carry between calls stays in libear's format (the last
// num_blocks input blocks and last_tail of B samples): the start-of-call overlap is recomputed
// from the input history, the tail for the next call by launch_conv_generic_tail.
//
// Same arithmetic as BlockConvolver::process (ref src/dsp/block_convolver_impl.cpp:143-237):
// zero-pad to 2B, transform, multiply by the filter spectrum, transform back, overlap-add the
// previous tail, scale by 1/(2B).  The differences are in how the transforms are carried out:
//
//  * two consecutive blocks of a channel ride one 1024-point COMPLEX transform as real and
//    imaginary part.  The filter is real, so conv(h, x0 + i x1) = conv(h, x0) + i conv(h, x1):
//    no even/odd split and no Hermitian unpacking is needed, the filter spectrum is simply
//    applied to all 1024 bins.
//  * one warp owns one pair: lane L holds elements L + 32 r (r = 0..31) in registers.  A 1024-point
//    transform is two register-resident 32-point transforms per lane around ONE transposition
//    through shared memory (1024 = 32 x 32).  Three of the four 32-point stages are decimation in
//    time (bit-reversed in, natural out; six-FMA butterflies), one is decimation in frequency
//    (natural in, bit-reversed out), chained so that no reordering pass exists -- every register
//    index is a compile-time constant.  After the inverse, y[n] and y[n + 512] sit in the same
//    lane, so the overlap-add and the tail carried to the next pair never leave the registers.
//  * a warp walks a run of consecutive pairs of its channel; the tail a run needs from before
//    its first pair is recomputed from one halo block.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdint>

#include "kernels.hpp"

namespace earb {
  namespace {
    constexpr int kFastB = 512;
    constexpr int kWarps = 8;
    constexpr int kRow = 34;  // padded row of the 32 x 32 transposition tile (float2 units): 16-byte
                              // aligned rows, float4 row reads and float2 column writes conflict-free

    __host__ __device__ constexpr int br5(int x) {
      return ((x & 1) << 4) | ((x & 2) << 2) | (x & 4) | ((x & 8) >> 2) | ((x & 16) >> 4);
    }

    // cos / sin of 2 pi k / 32; k is a compile-time constant after unrolling, so the switch folds
    __device__ __forceinline__ float cos32(int k) {
      switch (k & 31) {
        case 0: return 1.0f;
        case 1: case 31: return 0.98078528040323043f;
        case 2: case 30: return 0.92387953251128674f;
        case 3: case 29: return 0.83146961230254524f;
        case 4: case 28: return 0.70710678118654752f;
        case 5: case 27: return 0.55557023301960222f;
        case 6: case 26: return 0.38268343236508977f;
        case 7: case 25: return 0.19509032201612827f;
        case 8: case 24: return 0.0f;
        case 9: case 23: return -0.19509032201612827f;
        case 10: case 22: return -0.38268343236508977f;
        case 11: case 21: return -0.55557023301960222f;
        case 12: case 20: return -0.70710678118654752f;
        case 13: case 19: return -0.83146961230254524f;
        case 14: case 18: return -0.92387953251128674f;
        case 15: case 17: return -0.98078528040323043f;
        default: return -1.0f;
      }
    }
    __device__ __forceinline__ float sin32(int k) { return cos32(k - 8); }

    // ---- complex numbers as packed fp32 pairs (sm_100 FADD2 / FMUL2 / FFMA2: one issue slot for
    // the real and the imaginary part; swapping the halves and negating either one are operand
    // modifiers of those instructions, so the "multiply by i" patterns below cost nothing extra).
    // The kernel is bound by instruction issue in the butterflies, not by HBM.
    typedef unsigned long long cx;
    __device__ __forceinline__ cx mk(float re, float im) {
      cx r;
      asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(re), "f"(im));
      return r;
    }
    __device__ __forceinline__ float re(cx v) {
      float a;
      [[maybe_unused]] float b;
      asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v));
      return a;
    }
    __device__ __forceinline__ float im(cx v) {
      [[maybe_unused]] float a;
      float b;
      asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v));
      return b;
    }
    __device__ __forceinline__ cx add2(cx a, cx b) {
      cx r;
      asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
      return r;
    }
    __device__ __forceinline__ cx sub2(cx a, cx b) {
      cx r;
      asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
      return r;
    }
    __device__ __forceinline__ cx mul2(cx a, cx b) {
      cx r;
      asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
      return r;
    }
    __device__ __forceinline__ cx fma2(cx a, cx b, cx c) {
      cx r;
      asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
      return r;
    }
    __device__ __forceinline__ cx bc(float a) { return mk(a, a); }
    __device__ __forceinline__ cx mul_i(cx a) { return mk(-im(a), re(a)); }    // i a
    __device__ __forceinline__ cx mul_mi(cx a) { return mk(im(a), -re(a)); }   // -i a
    __device__ __forceinline__ cx neg(cx a) { return mk(-re(a), -im(a)); }
    // a * (c + i s), c and s scalars: (a.x c - a.y s, a.y c + a.x s)
    __device__ __forceinline__ cx cmul_cs(cx a, float c, float s) { return fma2(a, bc(c), mul2(mul_i(a), bc(s))); }

    // a * exp(-+ 2 pi i k / 32): forward uses the negative exponent
    template <bool INV>
    __device__ __forceinline__ cx mul_w32(cx a, int k) {
      k &= 31;
      if (k == 0) return a;
      if (k == 8) return INV ? mul_i(a) : mul_mi(a);
      if (k == 16) return neg(a);
      if (k == 24) return INV ? mul_mi(a) : mul_i(a);
      const float c = cos32(k), s = INV ? sin32(k) : -sin32(k);  // w = c + i s
      if (k == 4 || k == 12 || k == 20 || k == 28) {
        // |c| == |s|: one addition and one multiplication
        const float m = 0.70710678118654752f;
        const cx t = (c > 0.0f) == (s > 0.0f) ? add2(a, mul_i(a)) : sub2(a, mul_i(a));  // a (1 +- i)
        return mul2(t, bc(c > 0.0f ? m : -m));
      }
      return cmul_cs(a, c, s);
    }

    // 32-point transform in registers, decimation in frequency: natural order in, v[br5(k)] = X[k] out.
    // UPPER_ZERO: v[16..31] are known to be zero on entry (zero-padded block), first stage pruned.
    template <bool INV, bool UPPER_ZERO>
    __device__ __forceinline__ void dif32(cx (&v)[32]) {
#pragma unroll
      for (int i = 0; i < 16; i++) {
        if (UPPER_ZERO) {
          v[i + 16] = mul_w32<INV>(v[i], i);
        } else {
          const cx a = v[i], b = v[i + 16];
          v[i] = add2(a, b);
          v[i + 16] = mul_w32<INV>(sub2(a, b), i);
        }
      }
#pragma unroll
      for (int s = 8; s >= 1; s >>= 1) {
#pragma unroll
        for (int g = 0; g < 32; g += 2 * s) {
#pragma unroll
          for (int i = 0; i < s; i++) {
            const cx a = v[g + i], b = v[g + i + s];
            v[g + i] = add2(a, b);
            v[g + i + s] = mul_w32<INV>(sub2(a, b), i * (16 / s));
          }
        }
      }
    }

    // (a, b) <- (a + w b, a - w b), w = exp(-+ 2 pi i k / 32), in three packed FMAs: the complex
    // product is written as c (b + t i b) with t = s / c (or with the roles of c and s swapped when
    // |s| > |c|), so that the scaling by c folds into the two additions (Linzer-Feig).
    template <bool INV>
    __device__ __forceinline__ void bfly_dit(cx &a, cx &b, int k) {
      k &= 15;
      const cx a0 = a, b0 = b;
      if (k == 0) {
        a = add2(a0, b0);
        b = sub2(a0, b0);
        return;
      }
      if (k == 8) {  // w = -i (forward), +i (inverse)
        const cx wb = INV ? mul_i(b0) : mul_mi(b0);
        a = add2(a0, wb);
        b = sub2(a0, wb);
        return;
      }
      const float c = cos32(k), sn = INV ? sin32(k) : -sin32(k);  // w = c + i sn
      cx uv;
      float scale;
      if (fabsf(c) >= fabsf(sn)) {
        const float t = sn / c;                 // compile-time constant
        uv = fma2(mul_i(b0), bc(t), b0);        // w b / c  = b + t i b
        scale = c;
      } else {
        const float t = c / sn;
        uv = fma2(b0, bc(t), mul_i(b0));        // w b / sn = t b + i b
        scale = sn;
      }
      a = fma2(uv, bc(scale), a0);
      b = fma2(uv, bc(-scale), a0);
    }

    // decimation in time: v[br5(a)] = x[a] in, natural order out.
    // ODD_ZERO: the odd slots (inputs a >= 16) are known to be zero: first stage pruned.
    template <bool INV, bool ODD_ZERO>
    __device__ __forceinline__ void dit32(cx (&v)[32]) {
      if (ODD_ZERO) {
#pragma unroll
        for (int g = 0; g < 32; g += 2) v[g + 1] = v[g];
      }
#pragma unroll
      for (int s = ODD_ZERO ? 2 : 1; s <= 16; s <<= 1) {
#pragma unroll
        for (int g = 0; g < 32; g += 2 * s) {
#pragma unroll
          for (int i = 0; i < s; i++) bfly_dit<INV>(v[g + i], v[g + i + s], i * (16 / s));
        }
      }
    }

    // a * (bx + i by) and a * conj(bx + i by), the second factor given as two scalars
    __device__ __forceinline__ cx cmulf(cx a, float bx, float by) { return fma2(a, bc(bx), mul2(mul_i(a), bc(by))); }
    __device__ __forceinline__ cx cmulcf(cx a, float bx, float by) { return fma2(a, bc(bx), mul2(mul_mi(a), bc(by))); }

    struct FastArgs {
      long K;         // blocks per channel in this call
      long n_pairs;   // ceil(K / 2)
      int run_pairs;  // pairs per warp
      const float *in;
      long in_stride;
      float *out;
      long out_stride;
      const float *hist;      // [C][hist_len] the input samples before the call (hist_len <= 512)
      int hist_len;
      int carried_tail;       // 1 (B = 512): overlap of the first block = tail; 0: recomputed from hist
      const float *tail;      // [C][512] unnormalised tail before the call (carried_tail only)
      float *tail_new;        // [C][512] (carried_tail only)
      const float2 *filt;     // [n_filters][513]: 1024-point spectrum of the whole filter
      const int *filter_of;   // [C] filter id per channel (-1: none)
      const float2 *tw;       // [32][32]: tw[k1][b] = exp(-2 pi i b k1 / 1024), then (wide kernel)
                              // [32][32]: exp(-2 pi i (lane + 32 r) / 2048) at [r][lane]
      const float2 *filt_wide;  // wide kernel: [n_filters][2048], 2048-point spectrum of the whole filter
      int run_sb;               // wide kernel: super-blocks of 512 samples per warp (a multiple of 6)
      const float *add;       // nullable: out += delay(add)  (ConvLaunch::add)
      long add_stride, add_delay;
      const float *add_mem;
    };

    // forward transform of the pair held in v (v[br5(a)] = z[lane + 32 a] for a < 16, odd slots
    // zero), filter multiply, inverse transform.  On return v[n2] = y[lane + 32 n2] (real part:
    // first block of the pair, imaginary part: second block).
    // tw_s / H_s hold two consecutive slots per float4: [slot / 2][lane] = (slot, slot + 1);
    // twi_s pairs the slots the inverse needs together: [n1][lane] = (n1, n1 + 16).  H_s carries the
    // 1 / (2B) scale (a power of two: exact), so v comes back normalised.
    // PRUNED = false: all 32 slots hold input (v[br5(a)] = z[lane + 32 a], a < 32).
    template <bool PRUNED = true>
    __device__ __forceinline__ void convolve_pair(cx (&v)[32], cx *tile, const float4 *tw_s, const float4 *twi_s,
                                                  const float4 *H_s, int lane) {
      dit32<false, PRUNED>(v);  // over a: v[k1] = A_lane[k1]
#pragma unroll
      for (int k1 = 0; k1 < 32; k1 += 2) {
        const float4 t = tw_s[(k1 >> 1) * 32 + lane];
        tile[k1 * kRow + lane] = cmulf(v[k1], t.x, t.y);
        tile[(k1 + 1) * kRow + lane] = cmulf(v[k1 + 1], t.z, t.w);
      }
      __syncwarp();
      const ulonglong2 *row = reinterpret_cast<const ulonglong2 *>(tile + lane * kRow);
#pragma unroll
      for (int b = 0; b < 32; b += 2) {
        const ulonglong2 t = row[b >> 1];
        v[br5(b)] = t.x;
        v[br5(b + 1)] = t.y;
      }
      __syncwarp();
      dit32<false, false>(v);  // over b: v[k2] = Z[lane + 32 k2]
#pragma unroll
      for (int k2 = 0; k2 < 32; k2 += 2) {
        const float4 h = H_s[(k2 >> 1) * 32 + lane];
        v[k2] = cmulf(v[k2], h.x, h.y);
        v[k2 + 1] = cmulf(v[k2 + 1], h.z, h.w);
      }
      dif32<true, false>(v);  // inverse over k2: v[br5(n1)]
#pragma unroll
      for (int q = 0; q < 32; q += 2) {
        // slots q, q + 1 hold n1 = br5(q), br5(q) + 16
        const int n1 = br5(q);
        const float4 t = twi_s[n1 * 32 + lane];
        tile[n1 * kRow + lane] = cmulcf(v[q], t.x, t.y);
        tile[(n1 + 16) * kRow + lane] = cmulcf(v[q + 1], t.z, t.w);
      }
      __syncwarp();
#pragma unroll
      for (int b = 0; b < 32; b += 2) {
        const ulonglong2 t = row[b >> 1];
        v[br5(b)] = t.x;
        v[br5(b + 1)] = t.y;
      }
      __syncwarp();
      dit32<true, false>(v);  // inverse over b: v[n2] = y[lane + 32 n2]
    }

    template <bool ADD>
    __global__ void __launch_bounds__(kWarps * 32, 2) convolver_fast_kernel(FastArgs a) {
      extern __shared__ float4 fsm[];
      float4 *tw_s = fsm;        // [16][32]: (tw[2j][lane], tw[2j + 1][lane])
      float4 *H_s = tw_s + 512;  // [16][32]: (H[lane + 32 (2j)], H[lane + 32 (2j + 1)])
      float4 *twi_s = H_s + 512;  // [16][32]: (tw[n1][lane], tw[n1 + 16][lane])
      cx *tiles = reinterpret_cast<cx *>(twi_s + 512);  // kWarps x 32 x kRow complex numbers
      const int c = blockIdx.y;
      const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
      const int f = a.filter_of[c];
      // the carried tails stay unnormalised in global memory (shared with the generic kernel): scaled
      // by the power of two on the way in and out, which is exact
      const float norm = 1.0f / (float)(2 * kFastB);
      for (int i = threadIdx.x; i < 1024; i += kWarps * 32) {
        const int slot = i >> 5, L = i & 31;
        float2 *tw_dst = reinterpret_cast<float2 *>(tw_s) + ((slot >> 1) * 32 + L) * 2 + (slot & 1);
        float2 *H_dst = reinterpret_cast<float2 *>(H_s) + ((slot >> 1) * 32 + L) * 2 + (slot & 1);
        const float2 w = a.tw[i];
        *tw_dst = w;
        reinterpret_cast<float2 *>(twi_s)[((slot & 15) * 32 + L) * 2 + (slot >> 4)] = w;
        const int k = L + 32 * slot;
        float2 h = make_float2(0.0f, 0.0f);
        if (f >= 0) {
          // full spectrum of the real filter from its B + 1 stored bins
          h = k <= kFastB ? a.filt[(size_t)f * (kFastB + 1) + k] : a.filt[(size_t)f * (kFastB + 1) + (2 * kFastB - k)];
          if (k > kFastB) h.y = -h.y;
        }
        *H_dst = make_float2(h.x * norm, h.y * norm);
      }
      __syncthreads();

      const long p0 = ((long)blockIdx.x * kWarps + warp) * a.run_pairs;
      if (p0 >= a.n_pairs) return;
      const long p1 = min(a.n_pairs, p0 + a.run_pairs);
      cx *tile = tiles + (size_t)warp * 32 * kRow;
      const float *x = a.in + (size_t)c * a.in_stride;
      float *y = a.out + (size_t)c * a.out_stride;

      cx v[32];
      float tail[16];
      if (p0 == 0 && a.carried_tail) {
#pragma unroll
        for (int r = 0; r < 16; r++) tail[r] = a.tail[(size_t)c * kFastB + lane + 32 * r] * norm;
      } else {
        // halo: the 512 samples before the run as the imaginary part of a pair -- input of this call,
        // or (first run, B != 512) the input history, which is all a filter of hist_len taps reaches
        const float *xb = p0 == 0 ? a.hist + (size_t)c * a.hist_len - (kFastB - a.hist_len)
                                  : x + (size_t)(2 * p0 - 1) * kFastB;
        const int first = p0 == 0 ? kFastB - a.hist_len : 0;
#pragma unroll
        for (int r = 0; r < 16; r++) {
          const int t = lane + 32 * r;
          v[br5(r)] = mk(0.0f, t >= first ? __ldg(xb + t) : 0.0f);
        }
        convolve_pair(v, tile, tw_s, twi_s, H_s, lane);
#pragma unroll
        for (int r = 0; r < 16; r++) tail[r] = im(v[r + 16]);
      }
      for (long p = p0; p < p1; p++) {
        const bool has_second = 2 * p + 1 < a.K;
        const float *x0 = x + (size_t)(2 * p) * kFastB;
        // the next pair's 4 KB (32 lines) on their way into L2 while this pair is transformed
        if (p + 1 < p1 && 2 * p + 3 < a.K)
          asm volatile("prefetch.global.L2 [%0];" ::"l"(x0 + 2 * kFastB + lane * 32));
        float xr[16], xi[16];
#pragma unroll
        for (int r = 0; r < 16; r++) xr[r] = __ldcs(x0 + lane + 32 * r);
#pragma unroll
        for (int r = 0; r < 16; r++) xi[r] = has_second ? __ldcs(x0 + kFastB + lane + 32 * r) : 0.0f;
#pragma unroll
        for (int r = 0; r < 16; r++) v[br5(r)] = mk(xr[r], xi[r]);
        convolve_pair(v, tile, tw_s, twi_s, H_s, lane);
        float *y0 = y + (size_t)(2 * p) * kFastB;
        if (ADD) {
          // the delayed direct bus joins here: sample t of [add_mem; add] (the lines of the two
          // blocks are loaded before the stores so that their latency overlaps)
          const long t0 = (long)(2 * p) * kFastB + lane;
          const float *am = a.add_mem + (long)c * a.add_delay, *ad = a.add + (long)c * a.add_stride - a.add_delay;
          float d0[16], d1[16];
#pragma unroll
          for (int r = 0; r < 16; r++) {
            const long t = t0 + 32 * r;
            d0[r] = t < a.add_delay ? am[t] : __ldcs(ad + t);
          }
#pragma unroll
          for (int r = 0; r < 16; r++) {
            const long t = t0 + kFastB + 32 * r;
            d1[r] = !has_second ? 0.0f : (t < a.add_delay ? am[t] : __ldcs(ad + t));
          }
#pragma unroll
          for (int r = 0; r < 16; r++) __stcs(y0 + lane + 32 * r, (re(v[r]) + tail[r]) + d0[r]);
          if (has_second) {
#pragma unroll
            for (int r = 0; r < 16; r++) {
              __stcs(y0 + kFastB + lane + 32 * r, (im(v[r]) + re(v[r + 16])) + d1[r]);
              tail[r] = im(v[r + 16]);
            }
          } else {
#pragma unroll
            for (int r = 0; r < 16; r++) tail[r] = re(v[r + 16]);
          }
        } else {
#pragma unroll
          for (int r = 0; r < 16; r++) __stcs(y0 + lane + 32 * r, re(v[r]) + tail[r]);
          if (has_second) {
#pragma unroll
            for (int r = 0; r < 16; r++) {
              __stcs(y0 + kFastB + lane + 32 * r, im(v[r]) + re(v[r + 16]));
              tail[r] = im(v[r + 16]);
            }
          } else {
#pragma unroll
            for (int r = 0; r < 16; r++) tail[r] = re(v[r + 16]);
          }
        }
      }
      if (p1 == a.n_pairs && a.carried_tail) {
#pragma unroll
        for (int r = 0; r < 16; r++) a.tail_new[(size_t)c * kFastB + lane + 32 * r] = tail[r] * (float)(2 * kFastB);
      }
    }

    // ---------------------------------------------------------------- long calls: 2048-point units
    // With a constant 512-tap filter nothing ties the transform to 1,024 points: a 2,048-point
    // transform carries blocks of 1,536 samples (1,536 + 511 <= 2,048) -- three quarters of the
    // transform are signal instead of half.  A unit is 3,072 consecutive samples of a channel, the
    // first 1,536 as real part and the second 1,536 as imaginary part of z; its transform is split
    // by one radix-2 step (decimation in frequency) into two 1,024-point problems that run through
    // the register transforms above:
    //     even bins: s[n] = z[n] + z[n + 1024]                     * H[2k]     -> e[n]
    //     odd bins:  d[n] = (z[n] - z[n + 1024]) exp(-2 pi i n / 2048) * H[2k + 1] -> o[n]
    //     y[n] = e[n] + exp(+2 pi i n / 2048) o[n],   y[n + 1024] = e[n] - exp(+2 pi i n / 2048) o[n]
    // (z[n + 1024] is zero for n >= 512).  y[n], y[n + 512], y[n + 1024] and y[n + 1536] of both
    // blocks end in the same lane, so the overlap-add and the carried tail stay in registers as
    // before.  Per output sample: 2 x 1,024-point transform pairs per 3,072 samples instead of one
    // per 1,024 -- about 0.72 of the packed operations of the 1024-point kernel (which prunes its
    // zero half) and two thirds of its shared-memory traffic.  The halo of a run and the super-blocks
    // that do not fill a unit at the end of the channel go through the 1024-point pair code, with
    // the even bins as its spectrum (they ARE the 1,024-point spectrum of the filter).
#ifndef EARB_WIDE_WARPS
#define EARB_WIDE_WARPS 12
#endif
    constexpr int kWideWarps = EARB_WIDE_WARPS;       // per CTA
    constexpr int kWideCtas = kWideWarps <= 6 ? 2 : 1;  // per SM (168 registers: 12 warps per SM)
    constexpr int kWideResident = kWideWarps * kWideCtas;
    constexpr int kUnitSb = 6;  // super-blocks of 512 samples per unit
#ifndef EARB_WIDE_UNITS
#define EARB_WIDE_UNITS 16
#endif

    template <bool ADD>
    __global__ void __launch_bounds__(kWideWarps * 32, kWideCtas) convolver_wide_kernel(FastArgs a) {
      extern __shared__ float4 fsm[];
      float4 *tw_s = fsm;           // [16][32]: (tw[2j][lane], tw[2j + 1][lane])
      float4 *twi_s = tw_s + 512;   // [16][32]: (tw[n1][lane], tw[n1 + 16][lane])
      float4 *He_s = twi_s + 512;   // [16][32]: even bins (k = lane + 32 slot -> bin 2k), x 1 / 2048
      float4 *Ho_s = He_s + 512;    // [16][32]: odd bins (bin 2k + 1), x 1 / 2048
      float2 *w2_s = reinterpret_cast<float2 *>(Ho_s + 512);  // [32][32]: exp(-2 pi i (lane + 32 r) / 2048)
      cx *tiles = reinterpret_cast<cx *>(w2_s + 1024);        // kWideWarps x 32 x kRow complex numbers
      const int c = blockIdx.y;
      const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
      const int f = a.filter_of[c];
      const float norm = 1.0f / 2048.0f;                    // of the 2048-point transform (exact)
      const float tail_norm = 1.0f / (float)(2 * kFastB);   // the carried tails are libear's: x 2B
      for (int i = threadIdx.x; i < 1024; i += kWideWarps * 32) {
        const int slot = i >> 5, L = i & 31;
        const float2 w = a.tw[i];
        reinterpret_cast<float2 *>(tw_s)[((slot >> 1) * 32 + L) * 2 + (slot & 1)] = w;
        reinterpret_cast<float2 *>(twi_s)[((slot & 15) * 32 + L) * 2 + (slot >> 4)] = w;
        w2_s[i] = a.tw[1024 + i];
        const int k = L + 32 * slot;
        float2 he = make_float2(0.0f, 0.0f), ho = he;
        if (f >= 0) {
          he = a.filt_wide[(size_t)f * 2048 + 2 * k];
          ho = a.filt_wide[(size_t)f * 2048 + 2 * k + 1];
        }
        reinterpret_cast<float2 *>(He_s)[((slot >> 1) * 32 + L) * 2 + (slot & 1)] = make_float2(he.x * norm, he.y * norm);
        reinterpret_cast<float2 *>(Ho_s)[((slot >> 1) * 32 + L) * 2 + (slot & 1)] = make_float2(ho.x * norm, ho.y * norm);
      }
      __syncthreads();

      const long s0 = ((long)blockIdx.x * kWideWarps + warp) * a.run_sb;
      if (s0 >= a.K) return;
      const long s1 = min(a.K, s0 + (long)a.run_sb);
      cx *tile = tiles + (size_t)warp * 32 * kRow;
      const float *x = a.in + (size_t)c * a.in_stride;
      float *y = a.out + (size_t)c * a.out_stride;
      // ADD: sample t of [add_mem; add], the delayed direct bus
      const float *am = ADD ? a.add_mem + (long)c * a.add_delay : nullptr;
      const float *ad = ADD ? a.add + (long)c * a.add_stride - a.add_delay : nullptr;
      auto delayed = [&](long t) -> float { return t < a.add_delay ? am[t] : __ldcs(ad + t); };

      cx v[32];
      float tail[16];
      if (s0 == 0 && a.carried_tail) {
#pragma unroll
        for (int r = 0; r < 16; r++) tail[r] = a.tail[(size_t)c * kFastB + lane + 32 * r] * tail_norm;
      } else {
        // halo: the 512 samples before the run as the imaginary part of a 1024-point pair (even bins
        // x 1 / 2048: the result comes out halved)
        const float *xb = s0 == 0 ? a.hist + (size_t)c * a.hist_len - (kFastB - a.hist_len)
                                  : x + (size_t)(s0 - 1) * kFastB;
        const int first = s0 == 0 ? kFastB - a.hist_len : 0;
#pragma unroll
        for (int r = 0; r < 16; r++) {
          const int t = lane + 32 * r;
          v[br5(r)] = mk(0.0f, t >= first ? __ldg(xb + t) : 0.0f);
        }
        convolve_pair<true>(v, tile, tw_s, twi_s, He_s, lane);
#pragma unroll
        for (int r = 0; r < 16; r++) tail[r] = 2.0f * im(v[r + 16]);
      }

      long sb = s0;
      for (; sb + kUnitSb <= s1; sb += kUnitSb) {
        const float *xu = x + (size_t)sb * kFastB;
        float *yu = y + (size_t)sb * kFastB;
        // the next unit's 12 KB (96 lines) on their way into L2 while this one is transformed
        if (sb + 2 * kUnitSb <= s1) {
#pragma unroll
          for (int j = 0; j < 3; j++)
            asm volatile("prefetch.global.L2 [%0];" ::"l"(xu + kUnitSb * kFastB + (lane + 32 * j) * 32));
        }
        if (ADD) {
          // ... and this unit's 12 KB of the delayed direct bus, which the combine step at the end of
          // the unit adds: from DRAM they would cost a full memory latency per step
          const long t_unit = (long)sb * kFastB;
          if (t_unit >= a.add_delay) {
#pragma unroll
            for (int j = 0; j < 3; j++) asm volatile("prefetch.global.L2 [%0];" ::"l"(ad + t_unit + (lane + 32 * j) * 32));
          }
        }
        cx so[32];
#pragma unroll
        for (int r = 0; r < 16; r++) {
          const cx za = mk(__ldcs(xu + lane + 32 * r), __ldcs(xu + 3 * kFastB + lane + 32 * r));
          const cx zb = mk(__ldcs(xu + 2 * kFastB + lane + 32 * r), __ldcs(xu + 5 * kFastB + lane + 32 * r));
          const float2 w = w2_s[r * 32 + lane];
          v[br5(r)] = add2(za, zb);
          so[br5(r)] = cmulf(sub2(za, zb), w.x, w.y);
        }
#pragma unroll
        for (int r = 16; r < 32; r++) {
          const cx za = mk(__ldcs(xu + lane + 32 * r), __ldcs(xu + 3 * kFastB + lane + 32 * r));
          const float2 w = w2_s[r * 32 + lane];
          v[br5(r)] = za;
          so[br5(r)] = cmulf(za, w.x, w.y);
        }
        convolve_pair<false>(v, tile, tw_s, twi_s, He_s, lane);  // v[q] = e[lane + 32 q]
#pragma unroll
        for (int i = 0; i < 32; i++) {
          const cx t = v[i];
          v[i] = so[i];
          so[i] = t;
        }
        convolve_pair<false>(v, tile, tw_s, twi_s, Ho_s, lane);  // v[q] = o[lane + 32 q]; so = e
        float *ya = yu, *yb = yu + 3 * kFastB;
        const long ta = (long)sb * kFastB + lane, tb = ta + 3 * kFastB;
        // ADD: the six samples of the delayed direct bus that step q stores on top of are loaded one
        // step ahead, before the stores of the step in between (the compiler may not move a load
        // above a store that could alias it, and a load issued right before its use costs an L2
        // round trip per step with three warps per scheduler)
        const bool near_mem = ADD && (long)sb * kFastB < a.add_delay;  // the unit reaches into add_mem
        auto load_add = [&](int q, float(&d)[6]) {
          if (near_mem) {
            d[0] = delayed(ta + 32 * q), d[1] = delayed(ta + 32 * (q + 16)), d[2] = delayed(ta + 32 * (q + 32));
            d[3] = delayed(tb + 32 * q), d[4] = delayed(tb + 32 * (q + 16)), d[5] = delayed(tb + 32 * (q + 32));
          } else {
            d[0] = __ldcs(ad + ta + 32 * q), d[1] = __ldcs(ad + ta + 32 * (q + 16)), d[2] = __ldcs(ad + ta + 32 * (q + 32));
            d[3] = __ldcs(ad + tb + 32 * q), d[4] = __ldcs(ad + tb + 32 * (q + 16)), d[5] = __ldcs(ad + tb + 32 * (q + 32));
          }
        };
        float dn[6] = {0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f};
        if (ADD) load_add(0, dn);
#pragma unroll
        for (int q = 0; q < 16; q++) {
          const float2 w0 = w2_s[q * 32 + lane], w1 = w2_s[(q + 16) * 32 + lane];
          const cx t0 = cmulcf(v[q], w0.x, w0.y), t1 = cmulcf(v[q + 16], w1.x, w1.y);
          const cx ylo0 = add2(so[q], t0), yhi0 = sub2(so[q], t0);            // y[n], y[n + 1024], n = lane + 32 q
          const cx ylo1 = add2(so[q + 16], t1), yhi1 = sub2(so[q + 16], t1);  // y[n + 512], y[n + 1536]
          float oa0 = re(ylo0) + tail[q], oa1 = re(ylo1), oa2 = re(yhi0);
          float ob0 = im(ylo0) + re(yhi1), ob1 = im(ylo1), ob2 = im(yhi0);
          tail[q] = im(yhi1);
          if (ADD) {
            oa0 += dn[0], oa1 += dn[1], oa2 += dn[2];
            ob0 += dn[3], ob1 += dn[4], ob2 += dn[5];
            if (q + 1 < 16) load_add(q + 1, dn);
          }
          __stcs(ya + lane + 32 * q, oa0);
          __stcs(ya + lane + 32 * (q + 16), oa1);
          __stcs(ya + lane + 32 * (q + 32), oa2);
          __stcs(yb + lane + 32 * q, ob0);
          __stcs(yb + lane + 32 * (q + 16), ob1);
          __stcs(yb + lane + 32 * (q + 32), ob2);
        }
      }
      // what does not fill a unit (end of the channel only): 1024-point pairs, results doubled
      for (; sb < s1; sb += 2) {
        const bool has_second = sb + 1 < a.K;
        const float *x0 = x + (size_t)sb * kFastB;
#pragma unroll
        for (int r = 0; r < 16; r++)
          v[br5(r)] = mk(__ldcs(x0 + lane + 32 * r), has_second ? __ldcs(x0 + kFastB + lane + 32 * r) : 0.0f);
        convolve_pair<true>(v, tile, tw_s, twi_s, He_s, lane);
#pragma unroll
        for (int i = 0; i < 32; i++) v[i] = add2(v[i], v[i]);
        float *y0 = y + (size_t)sb * kFastB;
        const long t0 = (long)sb * kFastB + lane;
#pragma unroll
        for (int r = 0; r < 16; r++) {
          float o = re(v[r]) + tail[r];
          if (ADD) o += delayed(t0 + 32 * r);
          __stcs(y0 + lane + 32 * r, o);
        }
        if (has_second) {
#pragma unroll
          for (int r = 0; r < 16; r++) {
            float o = im(v[r]) + re(v[r + 16]);
            if (ADD) o += delayed(t0 + kFastB + 32 * r);
            __stcs(y0 + kFastB + lane + 32 * r, o);
            tail[r] = im(v[r + 16]);
          }
        } else {
#pragma unroll
          for (int r = 0; r < 16; r++) tail[r] = re(v[r + 16]);
        }
      }
      if (s1 == a.K && a.carried_tail) {
#pragma unroll
        for (int r = 0; r < 16; r++) a.tail_new[(size_t)c * kFastB + lane + 32 * r] = tail[r] * (float)(2 * kFastB);
      }
    }
  }  // namespace

  size_t conv_fast_tw_floats() { return 2 * 2048; }

  void conv_fast_build_tw(float *tw) {
    const double pi = 3.141592653589793238462643383279502884;
    for (int k1 = 0; k1 < 32; k1++)
      for (int b = 0; b < 32; b++) {
        const double ang = -2.0 * pi * (double)(b * k1) / 1024.0;
        tw[2 * (k1 * 32 + b)] = (float)cos(ang);
        tw[2 * (k1 * 32 + b) + 1] = (float)sin(ang);
      }
    // the radix-2 step of the wide kernel: exp(-2 pi i n / 2048), n = lane + 32 r at [r][lane]
    for (int n = 0; n < 1024; n++) {
      const double ang = -2.0 * pi * (double)n / 2048.0;
      tw[2 * (1024 + n)] = (float)cos(ang);
      tw[2 * (1024 + n) + 1] = (float)sin(ang);
    }
  }

  bool conv_fast_supported(int B, int nb) { return B >= 16 && B * nb <= kFastB && kFastB % B == 0; }

  void launch_convolver_fast(const ConvLaunch &L, const int *d_filter_of, const float *d_tw_fast, int n_sms,
                             cudaStream_t stream) {
    if (L.K <= 0) return;
    FastArgs a;
    a.K = L.K * L.B / kFastB;  // super-blocks of 512 samples (the caller keeps K * B a multiple of 512)
    a.n_pairs = (a.K + 1) / 2;
    // run length: amortise the one-block halo, keep a few waves of warps in flight
    int run = 32;
    while (run > 4 && (a.n_pairs + run - 1) / run * L.C < 16L * kWarps * n_sms) run /= 2;
    a.run_pairs = run;
    a.in = L.in;
    a.in_stride = L.in_stride;
    a.out = L.out;
    a.out_stride = L.out_stride;
    a.hist = L.hist;
    a.hist_len = L.B * L.nb;
    a.carried_tail = L.B == kFastB;
    a.tail = L.tail;
    a.tail_new = L.tail_new;
    a.filt = reinterpret_cast<const float2 *>(a.carried_tail ? L.filt : L.filt_fast);
    a.filter_of = d_filter_of;
    a.tw = reinterpret_cast<const float2 *>(d_tw_fast);
    a.add = L.add;
    a.add_stride = L.add_stride;
    a.add_delay = L.add_delay;
    a.add_mem = L.add_mem;
    a.filt_wide = reinterpret_cast<const float2 *>(L.filt_wide);
    a.run_sb = 0;
    // Which kernel, and how long the runs (units of 3,072 samples per warp; every run pays one halo
    // pair, about half a unit).  A CTA is kWideWarps runs of ONE channel and an SM holds one CTA, so
    // the launch proceeds in waves of n_sms CTAs: cost(u) = waves(u) x (u + 1/2) unit times.
    //   * long calls (cfg 4: 24 channels x 10 min): runs of 16 units, eight waves;
    //   * shorter ones (24 channels x 60 s): the run length that minimises the cost --
    //     usually the shortest that fits the whole call into ONE wave -- if that comes within a
    //     quarter of the ideal (all warps of the GPU busy all the time); else the pair kernel, whose
    //     short runs fill the GPU with less work;
    //   * L.wide == 2 (EARB200_CONV_PATH=wide, tests): wide whenever a run of 16 units exists.
    const long total_units = a.K / kUnitSb;
    auto ctas_at = [&](int u) {
      const long warps = (a.K + (long)u * kUnitSb - 1) / ((long)u * kUnitSb);
      return (long)L.C * ((warps + kWideWarps - 1) / kWideWarps);
    };
    int units = EARB_WIDE_UNITS;
    bool wide = false;
    if (a.filt_wide && L.wide == 2 && a.K >= 16 * kUnitSb) {
      while (units > 2 && (a.K + units * kUnitSb - 1) / (units * kUnitSb) * L.C < 6L * kWideResident * n_sms) units /= 2;
      wide = true;
    } else if (a.filt_wide && L.wide == 1 && a.K >= 16 * kUnitSb) {
      if (ctas_at(units) >= 4L * n_sms) {
        wide = true;
      } else {
        double best = 1e30;
        for (int u = 4; u <= 64 && (long)u <= total_units; u++) {
          const double cost = (double)((ctas_at(u) + n_sms - 1) / n_sms) * ((double)u + 0.5);
          if (cost < best) best = cost, units = u;
        }
        const double ideal = (double)total_units * L.C / ((double)kWideResident * n_sms) + 0.5;
        // (with the delayed bus joining, the one-wave form measured level with the pair kernel on the
        // Objects flow -- 0.864 against 0.850 ms -- while without it is a fifth faster at that size:
        // 24 channels x 60 s in 0.181 against 0.228 ms; those calls stay on the pair kernel)
        wide = best <= 1.25 * ideal && !a.add;
      }
    }
    if (wide) {
      a.run_sb = units * kUnitSb;
      const size_t wsmem = sizeof(float2) * (5 * 1024 + (size_t)kWideWarps * 32 * kRow);
      const long wwarps = (a.K + a.run_sb - 1) / a.run_sb;
      dim3 wgrid((unsigned)((wwarps + kWideWarps - 1) / kWideWarps), (unsigned)L.C);
      if (a.add) {
        cudaFuncSetAttribute(convolver_wide_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wsmem);
        convolver_wide_kernel<true><<<wgrid, kWideWarps * 32, wsmem, stream>>>(a);
      } else {
        cudaFuncSetAttribute(convolver_wide_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)wsmem);
        convolver_wide_kernel<false><<<wgrid, kWideWarps * 32, wsmem, stream>>>(a);
      }
      return;
    }
    const size_t smem = sizeof(float2) * (3072 + (size_t)kWarps * 32 * kRow);
    const long warps = (a.n_pairs + run - 1) / run;
    dim3 grid((unsigned)((warps + kWarps - 1) / kWarps), (unsigned)L.C);
    if (a.add) {
      cudaFuncSetAttribute(convolver_fast_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      convolver_fast_kernel<true><<<grid, kWarps * 32, smem, stream>>>(a);
    } else {
      cudaFuncSetAttribute(convolver_fast_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      convolver_fast_kernel<false><<<grid, kWarps * 32, smem, stream>>>(a);
    }
  }
}  // namespace earb
