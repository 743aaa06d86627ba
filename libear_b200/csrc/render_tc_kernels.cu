// K3 on the 5th-generation tensor cores (tcgen05 + TMEM), sm_100a only.
//
// Same contract as render_kernels.cu (LinearInterpMatrix::apply_interp/apply_constant,
// ref include/ear/dsp/gain_interpolator.hpp:250-299) for full, 16-byte aligned
// 256-sample tiles with M <= 16 outputs and N <= 256 inputs:
//
//     [S | D][t][m] = sum_n x[n][t] * [s_nm | e_nm - s_nm],   out[m][t] = S + p_t * D
//
// is a (256 x N) x (N x 32) GEMM per tile.  fp32 accuracy is kept with a 3xTF32 split
// (x = x_hi + x_lo, g = g_hi + g_lo; x_hi g_hi accumulates in one TMEM accumulator, the
// two 2^-11-sized cross terms in a second one, summed in fp32 registers at the end).
//
// Data flow per CTA (128 threads, one 256-sample tile = two M=128 UMMA tiles):
//   * one thread streams the tile's input rows into a shared-memory ring with bulk
//     asynchronous copies (cp.async.bulk, the 1-D TMA path) completing on mbarriers;
//   * every thread owns TMEM lane `tid`: it reads its samples of the 8 channels of a
//     stage from shared memory, splits them and writes them to TMEM as the A operand
//     (tcgen05.st, K-major: one 32-bit column per input channel);
//   * one thread issues tcgen05.mma kind::tf32 with A from TMEM and the gain matrix
//     (K-major, no swizzle, built in shared memory once per tile) as the B descriptor,
//     and commits to an mbarrier that releases the A buffer;
//   * the epilogue reads both accumulators back (tcgen05.ld), applies the ramp and
//     writes the M output rows with coalesced stores.
#include <cuda_runtime.h>

#include <cstdint>

#include "kernels.hpp"

namespace earb {
  namespace {
    constexpr int kTile = RENDER_TC_TILE;  // samples per CTA (2 UMMA M-tiles)
    constexpr int kMT = kTile / 128;       // M-tiles
    constexpr int kCh = 16;                // channels per stage = two K = 8 UMMA steps
    constexpr int kStages = 4;
    constexpr int kThreads = 128;        // converter threads (one per TMEM lane)
    constexpr int kThreadsTotal = 192;   // + TMA producer warp + MMA issuer warp
    constexpr int kNP = 32;                // UMMA N: 16 columns for S, 16 for D
    // TMEM columns: [main acc: kMT x 32][corr acc: kMT x 32][A: 2 buffers x kMT x (hi 16 + lo 16)]
    constexpr int kColMain = 0;
    constexpr int kColCorr = kMT * kNP;
    constexpr int kColA = 2 * kMT * kNP;
    constexpr int kTmemCols = 256;
    static_assert(kColA + 2 * kMT * 2 * kCh <= kTmemCols, "TMEM budget");

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
    // spins on the barrier phase; a protocol error traps after ~2 s instead of hanging the GPU
    __device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
      uint32_t done;
      const long long t0 = clock64();
      do {
        if (clock64() - t0 > 4000000000LL) __trap();
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
      } while (!done);
    }
    __device__ __forceinline__ float to_tf32(float x) {
      uint32_t r;
      asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
      return __uint_as_float(r);
    }
    __device__ __forceinline__ void tmem_st16(uint32_t taddr, const float (&v)[16]) {
      asm volatile(
          "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
          "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
          "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])),
          "r"(__float_as_uint(v[3])), "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])),
          "r"(__float_as_uint(v[6])), "r"(__float_as_uint(v[7])), "r"(__float_as_uint(v[8])),
          "r"(__float_as_uint(v[9])), "r"(__float_as_uint(v[10])), "r"(__float_as_uint(v[11])),
          "r"(__float_as_uint(v[12])), "r"(__float_as_uint(v[13])), "r"(__float_as_uint(v[14])),
          "r"(__float_as_uint(v[15]))
          : "memory");
    }
    __device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&v)[32]) {
      uint32_t r[32];
      asm volatile(
          "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
          "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
          "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
          : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
            "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]),
            "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]),
            "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]),
            "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
          : "r"(taddr)
          : "memory");
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
      for (int i = 0; i < 32; i++) v[i] = __uint_as_float(r[i]);
    }
    // D[tmem] (+)= A[tmem] * B[smem desc], kind::tf32, issued by one thread
    __device__ __forceinline__ void umma_tf32_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc,
                                                 uint32_t idesc, uint32_t accumulate) {
      asm volatile(
          "{\n\t.reg .pred p;\n\t"
          "setp.ne.b32 p, %4, 0;\n\t"
          "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
          ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
          : "memory");
    }
    __device__ __forceinline__ void umma_commit(uint64_t *bar) {
      asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                       smem_u32(bar))
                   : "memory");
    }

    // K-major, non-swizzled canonical layout of an [32 rows] x [K] tf32 operand: core matrices
    // of 8 rows x 16 bytes; the 32 rows of one 4-wide K chunk are contiguous (512 bytes)
    __device__ __forceinline__ int b_offset_floats(int row, int k) {
      return (k >> 2) * (kNP * 4) + row * 4 + (k & 3);
    }
    __device__ __forceinline__ uint64_t b_desc(uint32_t smem_addr) {
      const uint64_t lbo = (kNP * 16) >> 4;  // next 4-wide K chunk
      const uint64_t sbo = 128 >> 4;         // next 8-row group
      return (uint64_t)((smem_addr >> 4) & 0x3FFF) | (lbo << 16) | (sbo << 32) | (1ull << 46);
    }

    __device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
      asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
    }

    // Warp roles: warps 0..3 convert (thread = TMEM lane), warp 4 streams input (TMA),
    // warp 5 owns TMEM allocation and issues the UMMAs.  All hand-offs are mbarriers:
    //   full[s]   TMA       -> converters   stage s of the input ring has landed
    //   empty[s]  converters-> TMA          stage s has been read out of shared memory
    //   a_full[b] converters-> MMA          A buffer b is complete in TMEM
    //   a_free[b] MMA commit-> converters   the UMMAs reading A buffer b have finished
    //   b_ready   converters-> MMA          the gain matrix B is in shared memory
    template <bool RAMP>
    __global__ void __launch_bounds__(kThreadsTotal)
        render_tc_kernel(int N, int M, const float *__restrict__ in, long in_stride,
                         float *__restrict__ out, long out_stride,
                         const RenderItem *__restrict__ items,
                         const RenderSegment *__restrict__ segs, const float *__restrict__ gains) {
      extern __shared__ __align__(128) unsigned char smem_raw[];
      const int Kp = (N + kCh - 1) / kCh * kCh;  // K padded to whole stages
      float *sx = reinterpret_cast<float *>(smem_raw);  // [stages][kCh][kTile]
      float *sBhi = sx + kStages * kCh * kTile;         // [Kp/4][32][4]
      float *sBlo = sBhi + (size_t)Kp * kNP;
      __shared__ __align__(8) uint64_t full[kStages], empty[kStages];
      __shared__ __align__(8) uint64_t a_full[2], a_free[2], b_ready;
      __shared__ uint32_t tmem_base_s;

      const RenderItem item = items[blockIdx.x];
      const RenderSegment seg = segs[item.seg];
      const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
      const int n_iters = Kp / kCh;

      if (tid == 0) {
        for (int s = 0; s < kStages; s++) {
          mbar_init(&full[s], 1);
          mbar_init(&empty[s], 4);
        }
        for (int b = 0; b < 2; b++) {
          mbar_init(&a_full[b], 4);
          mbar_init(&a_free[b], 1);
        }
        mbar_init(&b_ready, 4);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      }
      if (warp == 5) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                         smem_u32(&tmem_base_s)),
                     "r"(kTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
      }
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncthreads();
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint32_t tmem_base = tmem_base_s;

      if (warp == 4) {
        // ------------------------------------------------------------ TMA producer
        if (lane == 0) {
          const float *base = in + item.t0;
          for (int it = 0; it < n_iters; it++) {
            const int slot = it % kStages;
            if (it >= kStages) mbar_wait(&empty[slot], (uint32_t)(((it / kStages) - 1) & 1));
            const int nch = min(kCh, N - it * kCh);
            mbar_expect_tx(&full[slot], (uint32_t)(nch * kTile * sizeof(float)));
            for (int c = 0; c < nch; c++)
              bulk_g2s(sx + (size_t)(slot * kCh + c) * kTile, base + (long)(it * kCh + c) * in_stride,
                       kTile * sizeof(float), &full[slot]);
          }
        }
      } else if (warp == 5) {
        // ------------------------------------------------------------ MMA issuer
        if (lane == 0) {
          // instruction descriptor: D f32, A/B tf32, both K-major, N = 32, M = 128
          const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(kNP >> 3) << 17) |
                                 ((uint32_t)(128 >> 4) << 24);
          const uint32_t bhi_addr = smem_u32(sBhi), blo_addr = smem_u32(sBlo);
          mbar_wait(&b_ready, 0);
          for (int it = 0; it < n_iters; it++) {
            const int abuf = it & 1;
            mbar_wait(&a_full[abuf], (uint32_t)((it >> 1) & 1));
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
            for (int ks = 0; ks < kCh / 8; ks++) {
              // one K = 8 step covers two 4-wide chunks of B (2 x 512 bytes)
              const int kstep = it * (kCh / 8) + ks;
              const uint64_t dhi = b_desc(bhi_addr + kstep * 2 * (kNP * 16));
              const uint64_t dlo = b_desc(blo_addr + kstep * 2 * (kNP * 16));
#pragma unroll
              for (int j = 0; j < kMT; j++) {
                const uint32_t a_hi = tmem_base + kColA + (abuf * kMT + j) * 2 * kCh + ks * 8;
                const uint32_t a_lo = a_hi + kCh;
                umma_tf32_ts(tmem_base + kColMain + j * kNP, a_hi, dhi, idesc, kstep > 0);
                umma_tf32_ts(tmem_base + kColCorr + j * kNP, a_lo, dhi, idesc, kstep > 0);
                umma_tf32_ts(tmem_base + kColCorr + j * kNP, a_hi, dlo, idesc, 1);
              }
            }
            umma_commit(&a_free[abuf]);  // arrives when every UMMA issued so far has finished
          }
        }
      } else {
        // ------------------------------------------------------------ converters (+ epilogue)
        const uint32_t lane_base = tmem_base + ((uint32_t)(warp * 32) << 16);
        // gain matrix B: rows 0..15 = start gains s, rows 16..31 = e - s, split hi/lo.  Zero
        // everything first (padding rows m >= M and channels k >= N must be finite), then
        // scatter the N x M raw gains, read coalesced with several loads in flight.
        {
          float4 *z = reinterpret_cast<float4 *>(sBhi);
          const float4 zero = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
          for (int i = tid; i < 2 * Kp * kNP / 4; i += kThreads) z[i] = zero;
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");
        {
          const float *ga = gains + (size_t)seg.pa * N * M;
          const float *gb = gains + (size_t)seg.pb * N * M;
          const int total = N * M;
#pragma unroll 4
          for (int r = tid; r < total; r += kThreads) {
            const float sv = __ldg(ga + r);
            const float dv = RAMP ? __ldg(gb + r) - sv : 0.0f;
            const int k = r / M, m = r - k * M;
            const float shi = to_tf32(sv), dhi = to_tf32(dv);
            const int o = b_offset_floats(m, k);
            sBhi[o] = shi;
            sBlo[o] = to_tf32(sv - shi);
            sBhi[o + 16 * 4] = dhi;  // row 16 + m
            sBlo[o + 16 * 4] = to_tf32(dv - dhi);
          }
        }
        // make the generic-proxy writes of B visible to the tensor core (async proxy)
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) mbar_arrive(&b_ready);

        for (int it = 0; it < n_iters; it++) {
          const int slot = it % kStages;
          const int abuf = it & 1;
          mbar_wait(&full[slot], (uint32_t)((it / kStages) & 1));
          // the UMMAs of iteration it - 2 must have finished reading this A buffer
          if (it >= 2) {
            mbar_wait(&a_free[abuf], (uint32_t)(((it >> 1) - 1) & 1));
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          }
          const int nch = min(kCh, N - it * kCh);
#pragma unroll
          for (int j = 0; j < kMT; j++) {
            float hi[kCh], lo[kCh];
#pragma unroll
            for (int c = 0; c < kCh; c++) {
              const float x = (c < nch) ? sx[(size_t)(slot * kCh + c) * kTile + j * 128 + tid] : 0.0f;
              hi[c] = to_tf32(x);
              lo[c] = to_tf32(x - hi[c]);
            }
            const uint32_t a_col = kColA + (abuf * kMT + j) * 2 * kCh;
            tmem_st16(lane_base + a_col, hi);
            tmem_st16(lane_base + a_col + kCh, lo);
          }
          asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
          asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
          __syncwarp();
          if (lane == 0) {
            mbar_arrive(&a_full[abuf]);
            mbar_arrive(&empty[slot]);
          }
        }
        // the last commit covers every UMMA of the tile
        {
          const int last = n_iters - 1;
          mbar_wait(&a_free[last & 1], (uint32_t)((last >> 1) & 1));
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        }
#pragma unroll
        for (int j = 0; j < kMT; j++) {
          float a[32], b[32];
          tmem_ld32(lane_base + kColMain + j * kNP, a);
          tmem_ld32(lane_base + kColCorr + j * kNP, b);
          const long t = item.t0 + j * 128 + tid;
          const float p = RAMP ? (float)(t - seg.ramp_start) * seg.scale : 0.0f;
#pragma unroll
          for (int m = 0; m < 16; m++) {
            if (m < M) {
              const float sv = a[m] + b[m];
              const float dv = a[16 + m] + b[16 + m];
              __stcs(out + (long)m * out_stride + t, RAMP ? fmaf(p, dv, sv) : sv);
            }
          }
        }
      }
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncthreads();
      if (warp == 5) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols)
                     : "memory");
      }
    }
  }  // namespace

  size_t render_tc_smem_bytes(int N) {
    const int Kp = (N + kCh - 1) / kCh * kCh;
    return sizeof(float) * ((size_t)kStages * kCh * kTile + (size_t)2 * Kp * kNP);
  }

  bool render_tc_supported(int N, int M) { return M <= 16 && N <= 256; }

  void launch_render_tc(int N, int M, const float *in, long in_stride, float *out, long out_stride,
                        const RenderItem *d_items_ramp, long n_ramp, const RenderItem *d_items_const,
                        long n_const, const RenderSegment *d_segs, const float *d_gains,
                        cudaStream_t stream) {
    const size_t smem = render_tc_smem_bytes(N);
    if (n_ramp > 0) {
      cudaFuncSetAttribute(render_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      render_tc_kernel<true><<<(unsigned)n_ramp, kThreadsTotal, smem, stream>>>(
          N, M, in, in_stride, out, out_stride, d_items_ramp, d_segs, d_gains);
    }
    if (n_const > 0) {
      cudaFuncSetAttribute(render_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      render_tc_kernel<false><<<(unsigned)n_const, kThreadsTotal, smem, stream>>>(
          N, M, in, in_stride, out, out_stride, d_items_const, d_segs, d_gains);
    }
  }
}  // namespace earb
