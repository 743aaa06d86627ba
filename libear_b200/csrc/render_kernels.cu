// K3: interpolated N -> M mixing (the GainInterpolator hot loop).
//
// Replaces LinearInterpMatrix::apply_interp / apply_constant
// (ref include/ear/dsp/gain_interpolator.hpp:250-299; Vector :214-241 and Single
// :186-209 are the N = 1 / N = M = 1 cases) driven by GainInterpolator::process
// (:53-87).  The reference evaluates, per (n, m, t),
//     p = (float)((block_start + i) - start) * (1.0f / (end - start))
//     out[m][t] += in[n][t] * ((1 - p) * s_nm + p * e_nm)
// This kernel uses the algebraically equal two-contraction form
//     out[m][t] = sum_n in[n][t] * s_nm  +  p_t * sum_n in[n][t] * (e_nm - s_nm)
// (2 FMA per (n, m, t) instead of 5 flops), which keeps the kernel on the HBM side of
// the FP32 ridge.  Results differ from the CPU loop by reassociation only (~1e-7
// relative; the stated tolerance is 1e-6 of peak).
//
// One CTA renders one work item: a run of <= 128*TPT samples inside one interpolation
// segment.  The segment's two gain matrices are staged in shared memory (start gains
// and their difference), every thread streams TPT consecutive samples of all N
// channels straight from HBM with 128-bit loads and keeps 2*TPT*MP accumulators in
// registers.
#include <cuda_runtime.h>

#include "kernels.hpp"

namespace earb {
  namespace {
    constexpr int kThreads = RENDER_THREADS;

    template <int TPT>
    struct Vec;
    template <>
    struct Vec<4> { using type = float4; };
    template <>
    struct Vec<2> { using type = float2; };

    template <int TPT>
    __device__ __forceinline__ void load_vec(const float *p, float (&x)[TPT]) {
      using V = typename Vec<TPT>::type;
      V v = __ldcs(reinterpret_cast<const V *>(p));  // streaming: read once
      const float *f = reinterpret_cast<const float *>(&v);
#pragma unroll
      for (int j = 0; j < TPT; j++) x[j] = f[j];
    }

    template <int MP, int TPT, bool RAMP>
    __global__ void __launch_bounds__(kThreads)
        render_kernel(int N, int M, const float *__restrict__ in, long in_stride,
                      float *__restrict__ out, long out_stride,
                      const RenderItem *__restrict__ items,
                      const RenderSegment *__restrict__ segs,
                      const float *__restrict__ gains) {
      extern __shared__ float4 smem4[];
      float *sS = reinterpret_cast<float *>(smem4);  // [N][MP] start gains
      float *sD = sS + (size_t)N * MP;               // [N][MP] end - start

      const RenderItem item = items[blockIdx.x];
      const RenderSegment seg = segs[item.seg];
      {
        const float *ga = gains + (size_t)seg.pa * N * M;
        const float *gb = gains + (size_t)seg.pb * N * M;
        for (int idx = threadIdx.x; idx < N * MP; idx += kThreads) {
          const int n = idx / MP, m = idx - n * MP;
          float s = 0.0f, d = 0.0f;
          if (m < M) {
            s = ga[n * M + m];
            if (RAMP) d = gb[n * M + m] - s;
          }
          sS[idx] = s;
          if (RAMP) sD[idx] = d;
        }
      }
      __syncthreads();

      const int t_local = threadIdx.x * TPT;
      const long t_first = item.t0 + t_local;
      const int valid = min(TPT, item.len - t_local);  // samples this thread owns
      if (valid <= 0) return;
      const float *ip = in + t_first;
      const bool vec_in = (valid == TPT) && ((reinterpret_cast<uintptr_t>(ip) & (sizeof(float) * TPT - 1)) == 0) &&
                          (in_stride % TPT == 0);

      float accS[TPT][MP], accD[TPT][MP];
#pragma unroll
      for (int j = 0; j < TPT; j++)
#pragma unroll
        for (int m = 0; m < MP; m++) {
          accS[j][m] = 0.0f;
          accD[j][m] = 0.0f;
        }

      const float4 *sS4 = reinterpret_cast<const float4 *>(sS);
      const float4 *sD4 = reinterpret_cast<const float4 *>(sD);

      auto mac = [&](int n, const float (&x)[TPT]) {
#pragma unroll
        for (int q = 0; q < MP / 4; q++) {
          const float4 s4 = sS4[n * (MP / 4) + q];
          const float sv[4] = {s4.x, s4.y, s4.z, s4.w};
#pragma unroll
          for (int k = 0; k < 4; k++)
#pragma unroll
            for (int j = 0; j < TPT; j++) accS[j][4 * q + k] = fmaf(x[j], sv[k], accS[j][4 * q + k]);
          if (RAMP) {
            const float4 d4 = sD4[n * (MP / 4) + q];
            const float dv[4] = {d4.x, d4.y, d4.z, d4.w};
#pragma unroll
            for (int k = 0; k < 4; k++)
#pragma unroll
              for (int j = 0; j < TPT; j++) accD[j][4 * q + k] = fmaf(x[j], dv[k], accD[j][4 * q + k]);
          }
        }
      };

      if (vec_in) {
        constexpr int U = RENDER_UNROLL;
        int n = 0;
        for (; n + U <= N; n += U) {
          float x[U][TPT];
#pragma unroll
          for (int u = 0; u < U; u++) load_vec<TPT>(ip + (long)(n + u) * in_stride, x[u]);
#pragma unroll
          for (int u = 0; u < U; u++) mac(n + u, x[u]);
        }
        for (; n < N; n++) {
          float x[TPT];
          load_vec<TPT>(ip + (long)n * in_stride, x);
          mac(n, x);
        }
      } else {
        for (int n = 0; n < N; n++) {
          float x[TPT];
#pragma unroll
          for (int j = 0; j < TPT; j++) x[j] = (j < valid) ? __ldcs(ip + (long)n * in_stride + j) : 0.0f;
          mac(n, x);
        }
      }

      // out = S + p * D
      float p[TPT];
#pragma unroll
      for (int j = 0; j < TPT; j++)
        p[j] = RAMP ? (float)((t_first + j) - seg.ramp_start) * seg.scale : 0.0f;
      float *op = out + t_first;
      const bool vec_out = (valid == TPT) && ((reinterpret_cast<uintptr_t>(op) & (sizeof(float) * TPT - 1)) == 0) &&
                           (out_stride % TPT == 0);
#pragma unroll
      for (int m = 0; m < MP; m++) {
        if (m < M) {
          float r[TPT];
#pragma unroll
          for (int j = 0; j < TPT; j++) r[j] = RAMP ? fmaf(p[j], accD[j][m], accS[j][m]) : accS[j][m];
          float *o = op + (long)m * out_stride;
          if (vec_out) {
            using V = typename Vec<TPT>::type;
            V v;
            float *f = reinterpret_cast<float *>(&v);
#pragma unroll
            for (int j = 0; j < TPT; j++) f[j] = r[j];
            __stcs(reinterpret_cast<V *>(o), v);
          } else {
#pragma unroll
            for (int j = 0; j < TPT; j++)
              if (j < valid) o[j] = r[j];
          }
        }
      }
    }

    template <int MP, int TPT, bool RAMP>
    void launch_one(int N, int M, const float *in, long in_stride, float *out, long out_stride,
                    const RenderItem *items, long n_items, const RenderSegment *segs,
                    const float *gains, cudaStream_t stream) {
      if (n_items <= 0) return;
      const size_t smem = sizeof(float) * (size_t)N * MP * (RAMP ? 2 : 1);
      auto kern = render_kernel<MP, TPT, RAMP>;
      if (smem > 48 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      kern<<<(unsigned)n_items, kThreads, smem, stream>>>(N, M, in, in_stride, out, out_stride, items,
                                                          segs, gains);
    }
  }  // namespace

  int render_tile_samples(int M) { return kThreads * ((M <= 16) ? 4 : 2); }

  size_t render_smem_bytes(int N, int M) {
    const int MP = (M + 7) / 8 * 8;
    return sizeof(float) * (size_t)N * MP * 2;
  }

  void launch_render(int N, int M, const float *in, long in_stride, float *out, long out_stride,
                     const RenderItem *d_items_ramp, long n_ramp, const RenderItem *d_items_const,
                     long n_const, const RenderSegment *d_segs, const float *d_gains,
                     cudaStream_t stream) {
#define EARB_RENDER_CASE(MP, TPT)                                                                  \
  launch_one<MP, TPT, true>(N, M, in, in_stride, out, out_stride, d_items_ramp, n_ramp, d_segs,   \
                            d_gains, stream);                                                      \
  launch_one<MP, TPT, false>(N, M, in, in_stride, out, out_stride, d_items_const, n_const, d_segs, \
                             d_gains, stream);
    if (M <= 8) {
      EARB_RENDER_CASE(8, 4)
    } else if (M <= 16) {
      EARB_RENDER_CASE(16, 4)
    } else if (M <= 24) {
      EARB_RENDER_CASE(24, 2)
    } else {
      EARB_RENDER_CASE(32, 2)
    }
#undef EARB_RENDER_CASE
  }
}  // namespace earb
