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

#include <stdexcept>

#include <cstdint>

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
                      const float *__restrict__ gains, int accumulate) {
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
          if (accumulate) {  // mix into the bus instead of overwriting it
#pragma unroll
            for (int j = 0; j < TPT; j++)
              if (j < valid) r[j] += o[j];
          }
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

    // ---------------------------------------------------------------- TMA-staged fast path
    // Full, 16-byte aligned 512-sample tiles: one elected thread streams the tile's input
    // rows into a shared-memory ring with bulk asynchronous copies (cp.async.bulk, the 1-D
    // TMA path) completing on mbarriers, kPipeStages stages of kPipeCh channels deep, so
    // ~48 KB per CTA are in flight while all threads run the two-contraction FMAs out of
    // shared memory.  Thread (tg, mg): samples {4 tg .. 4 tg + 3} and {256 + 4 tg ..} of the
    // tile, outputs {4 mg .. 4 mg + 3}; a warp shares mg, so gain loads are broadcasts and
    // sample loads are conflict-free 128-bit reads.
    constexpr int kPipeTile = RENDER_PIPE_TILE;
    constexpr int kPipeCh = 4;
    constexpr int kPipeStages = 6;

    __device__ __forceinline__ uint32_t smem_u32(const void *p) {
      return (uint32_t)__cvta_generic_to_shared(p);
    }
    __device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
    }
    __device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                   : "memory");
    }
    __device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
      asm volatile(
          "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
              smem_u32(dst)),
          "l"(src), "r"(bytes), "r"(smem_u32(bar))
          : "memory");
    }
    __device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
      uint32_t done;
      do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
      } while (!done);
    }

    template <int MP, bool RAMP>
    __global__ void __launch_bounds__(64 * (MP / 4))
        render_pipe_kernel(int N, int M, const float *__restrict__ in, long in_stride,
                           float *__restrict__ out, long out_stride,
                           const RenderItem *__restrict__ items,
                           const RenderSegment *__restrict__ segs,
                           const float *__restrict__ gains, int accumulate) {
      constexpr int NT = 64 * (MP / 4);
      extern __shared__ __align__(128) unsigned char smem_raw[];
      float *sx = reinterpret_cast<float *>(smem_raw);     // [stages][kPipeCh][tile]
      float *sS = sx + kPipeStages * kPipeCh * kPipeTile;  // [N][MP]
      float *sD = sS + (size_t)N * MP;                     // [N][MP]
      __shared__ __align__(8) uint64_t full[kPipeStages];

      const RenderItem item = items[blockIdx.x];
      const RenderSegment seg = segs[item.seg];
      const int tid = threadIdx.x, tg = tid & 63, mg = tid >> 6;
      if (tid == 0) {
        for (int s = 0; s < kPipeStages; s++) mbar_init(&full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      }
      {
        const float *ga = gains + (size_t)seg.pa * N * M;
        const float *gb = gains + (size_t)seg.pb * N * M;
        for (int idx = tid; idx < N * MP; idx += NT) {
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

      const float *base = in + item.t0;
      const int n_iters = (N + kPipeCh - 1) / kPipeCh;
      auto issue = [&](int it) {
        const int slot = it % kPipeStages;
        const int nch = min(kPipeCh, N - it * kPipeCh);
        mbar_expect_tx(&full[slot], (uint32_t)(nch * kPipeTile * sizeof(float)));
        for (int c = 0; c < nch; c++)
          bulk_g2s(sx + (size_t)(slot * kPipeCh + c) * kPipeTile,
                   base + (long)(it * kPipeCh + c) * in_stride, kPipeTile * sizeof(float), &full[slot]);
      };
      if (tid == 0)
        for (int it = 0; it < kPipeStages - 1 && it < n_iters; it++) issue(it);

      float aS[2][4][4], aD[2][4][4];  // [half][sample][m]
#pragma unroll
      for (int h = 0; h < 2; h++)
#pragma unroll
        for (int j = 0; j < 4; j++)
#pragma unroll
          for (int k = 0; k < 4; k++) {
            aS[h][j][k] = 0.0f;
            aD[h][j][k] = 0.0f;
          }
      const float4 *sS4 = reinterpret_cast<const float4 *>(sS);
      const float4 *sD4 = reinterpret_cast<const float4 *>(sD);

      for (int it = 0; it < n_iters; it++) {
        // the slot refilled here was drained in iteration it - 1 (barrier at its end)
        if (tid == 0 && it + kPipeStages - 1 < n_iters) issue(it + kPipeStages - 1);
        const int slot = it % kPipeStages;
        mbar_wait(&full[slot], (uint32_t)((it / kPipeStages) & 1));
        const int nch = min(kPipeCh, N - it * kPipeCh);
#pragma unroll
        for (int c = 0; c < kPipeCh; c++) {
          if (c < nch) {
            const int n = it * kPipeCh + c;
            const float *row = sx + (size_t)(slot * kPipeCh + c) * kPipeTile;
            const float4 xa = *reinterpret_cast<const float4 *>(row + 4 * tg);
            const float4 xb = *reinterpret_cast<const float4 *>(row + 256 + 4 * tg);
            const float x[2][4] = {{xa.x, xa.y, xa.z, xa.w}, {xb.x, xb.y, xb.z, xb.w}};
            const float4 s4 = sS4[n * (MP / 4) + mg];
            const float sv[4] = {s4.x, s4.y, s4.z, s4.w};
#pragma unroll
            for (int h = 0; h < 2; h++)
#pragma unroll
              for (int j = 0; j < 4; j++)
#pragma unroll
                for (int k = 0; k < 4; k++) aS[h][j][k] = fmaf(x[h][j], sv[k], aS[h][j][k]);
            if (RAMP) {
              const float4 d4 = sD4[n * (MP / 4) + mg];
              const float dv[4] = {d4.x, d4.y, d4.z, d4.w};
#pragma unroll
              for (int h = 0; h < 2; h++)
#pragma unroll
                for (int j = 0; j < 4; j++)
#pragma unroll
                  for (int k = 0; k < 4; k++) aD[h][j][k] = fmaf(x[h][j], dv[k], aD[h][j][k]);
            }
          }
        }
        __syncthreads();
      }

#pragma unroll
      for (int h = 0; h < 2; h++) {
        const long t = item.t0 + 256 * h + 4 * tg;
        float p[4];
#pragma unroll
        for (int j = 0; j < 4; j++) p[j] = RAMP ? (float)((t + j) - seg.ramp_start) * seg.scale : 0.0f;
#pragma unroll
        for (int k = 0; k < 4; k++) {
          const int m = 4 * mg + k;
          if (m < M) {
            float4 v;
            v.x = RAMP ? fmaf(p[0], aD[h][0][k], aS[h][0][k]) : aS[h][0][k];
            v.y = RAMP ? fmaf(p[1], aD[h][1][k], aS[h][1][k]) : aS[h][1][k];
            v.z = RAMP ? fmaf(p[2], aD[h][2][k], aS[h][2][k]) : aS[h][2][k];
            v.w = RAMP ? fmaf(p[3], aD[h][3][k], aS[h][3][k]) : aS[h][3][k];
            if (accumulate) {
              const float4 o = __ldcs(reinterpret_cast<const float4 *>(out + (long)m * out_stride + t));
              v.x += o.x;
              v.y += o.y;
              v.z += o.z;
              v.w += o.w;
            }
            __stcs(reinterpret_cast<float4 *>(out + (long)m * out_stride + t), v);
          }
        }
      }
    }

    template <int MP, bool RAMP>
    void launch_pipe(int N, int M, const float *in, long in_stride, float *out, long out_stride,
                     const RenderItem *items, long n_items, const RenderSegment *segs,
                     const float *gains, int accumulate, cudaStream_t stream) {
      if (n_items <= 0) return;
      const size_t smem = sizeof(float) * ((size_t)kPipeStages * kPipeCh * kPipeTile + (size_t)N * MP * 2);
      auto kern = render_pipe_kernel<MP, RAMP>;
      cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      kern<<<(unsigned)n_items, 64 * (MP / 4), smem, stream>>>(N, M, in, in_stride, out, out_stride, items,
                                                              segs, gains, accumulate);
    }

    template <int MP, int TPT, bool RAMP>
    void launch_one(int N, int M, const float *in, long in_stride, float *out, long out_stride,
                    const RenderItem *items, long n_items, const RenderSegment *segs,
                    const float *gains, int accumulate, cudaStream_t stream) {
      if (n_items <= 0) return;
      const size_t smem = sizeof(float) * (size_t)N * MP * (RAMP ? 2 : 1);
      auto kern = render_kernel<MP, TPT, RAMP>;
      if (smem > 48 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      kern<<<(unsigned)n_items, kThreads, smem, stream>>>(N, M, in, in_stride, out, out_stride, items,
                                                          segs, gains, accumulate);
    }
  }  // namespace

  int render_tile_samples(int M) { return kThreads * ((M <= 16) ? 4 : 2); }

  size_t render_smem_bytes(int N, int M) {
    const int MP = (M + 7) / 8 * 8;
    return sizeof(float) * ((size_t)N * MP * 2 + (size_t)kPipeStages * kPipeCh * kPipeTile);
  }

  void launch_render(int N, int M, const float *in, long in_stride, float *out, long out_stride,
                     const RenderPlan &plan, const RenderSegment *d_segs, const float *d_gains,
                     cudaStream_t stream) {
    const RenderItem *it = plan.d_items;
    const RenderItem *fast_ramp = it, *fast_const = fast_ramp + plan.n_fast_ramp;
    const RenderItem *slow_ramp = fast_const + plan.n_fast_const, *slow_const = slow_ramp + plan.n_slow_ramp;
    long n_fast_ramp = plan.n_fast_ramp, n_fast_const = plan.n_fast_const;
    if (plan.use_tc) {
      // the ramp and constant tiles are adjacent in the item list: one persistent launch
      if (!launch_render_tc(N, M, in, in_stride, plan.n_samples, out, out_stride, fast_ramp,
                            n_fast_ramp + n_fast_const, d_segs, plan.d_images, plan.n_points, plan.n_sms,
                            plan.accumulate, plan.bus, stream))
        throw std::runtime_error("cuTensorMapEncodeTiled failed for the render input");
      n_fast_ramp = n_fast_const = 0;
    }
#define EARB_RENDER_CASE(MP, TPT)                                                                      \
  launch_pipe<MP, true>(N, M, in, in_stride, out, out_stride, fast_ramp, n_fast_ramp, d_segs,          \
                        d_gains, plan.accumulate, stream);                                                              \
  launch_pipe<MP, false>(N, M, in, in_stride, out, out_stride, fast_const, n_fast_const, d_segs,       \
                         d_gains, plan.accumulate, stream);                                                             \
  launch_one<MP, TPT, true>(N, M, in, in_stride, out, out_stride, slow_ramp, plan.n_slow_ramp, d_segs, \
                            d_gains, plan.accumulate, stream);                                                          \
  launch_one<MP, TPT, false>(N, M, in, in_stride, out, out_stride, slow_const, plan.n_slow_const,      \
                             d_segs, d_gains, plan.accumulate, stream);
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
  namespace {
    // flags[i] = gain matrix i equals matrix i - 1 element for element (constant_interp,
    // ref include/ear/dsp/gain_interpolator.hpp:139-141); one block per point
    __global__ void render_points_same_kernel(long mat, const float *__restrict__ gains, char *__restrict__ flags) {
      const long i = (long)blockIdx.x + 1;
      const float *a = gains + (i - 1) * mat, *b = a + mat;
      int same = 1;
      for (long k = threadIdx.x; k < mat; k += blockDim.x) same &= (a[k] == b[k]);
      same = __syncthreads_and(same);
      if (threadIdx.x == 0) flags[i] = (char)same;
    }
  }  // namespace
  void launch_render_points_same(long mat, long n_points, const float *d_gains, char *d_flags, cudaStream_t stream) {
    if (n_points < 2) return;
    render_points_same_kernel<<<(unsigned)(n_points - 1), 256, 0, stream>>>(mat, d_gains, d_flags);
  }
}  // namespace earb
