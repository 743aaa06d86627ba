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
// Data flow (one persistent CTA of 448 threads per SM, looping over 256-sample tiles = two
// M=128 UMMA tiles each):
//   * a producer warp streams the tiles' input rows into a shared-memory ring with bulk
//     asynchronous copies (cp.async.bulk, the 1-D TMA path) completing on mbarriers; the ring
//     runs continuously across tile boundaries;
//   * eight converter warps (thread = TMEM lane = sample) read their samples of the 16 channels
//     of a stage from shared memory, split them and write them to TMEM as the A operand
//     (tcgen05.st, K-major: one 32-bit column per input channel), double buffered;
//   * one thread issues tcgen05.mma kind::tf32 with A from TMEM and the gain matrix
//     (K-major, no swizzle, in shared memory) as the B descriptor, into double-buffered TMEM
//     accumulators, and commits to mbarriers that release the A buffer / publish the tile;
//   * four epilogue warps read both accumulators back (tcgen05.ld), apply the ramp, write the
//     M output rows with coalesced streaming stores, and build the gain matrix two tiles ahead.
#include <cuda.h>
#ifndef EARB_RENDER_EXPERIMENT
#define EARB_RENDER_EXPERIMENT 0
#endif
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>

#include "kernels.hpp"

namespace earb {
  namespace {
    constexpr int kTile = RENDER_TC_TILE;  // samples per CTA (2 UMMA M-tiles)
    constexpr int kMT = kTile / 128;       // M-tiles
    constexpr int kCh = 16;                // channels per stage = two K = 8 UMMA steps
    constexpr int kNP = 32;                // UMMA N: 16 columns for S, 16 for D

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
    // 2-D tiled TMA load: box {kTile samples, kCh channels} at (sample x, channel y)
    __device__ __forceinline__ void tma_load_2d(void *dst, const CUtensorMap *map, int x, int y, uint64_t *bar) {
      asm volatile(
          "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
              smem_u32(dst)),
          "l"(map), "r"(x), "r"(y), "r"(smem_u32(bar))
          : "memory");
    }
    // round-to-nearest (ties away) onto the 10-bit TF32 mantissa with two integer operations;
    // cvt.rna.tf32.f32 adds NaN/Inf handling the audio path does not need (Inf stays Inf here)
    __device__ __forceinline__ float round_tf32(float x) {
      return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u);
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
      return (k >> 2) * (2 * kNP * 4) + row * 4 + (k & 3);
    }
    __device__ __forceinline__ uint64_t b_desc(uint32_t smem_addr) {
      const uint64_t lbo = (2 * kNP * 16) >> 4;  // next 4-wide K chunk (64 rows x 16 bytes)
      const uint64_t sbo = 128 >> 4;         // next 8-row group
      return (uint64_t)((smem_addr >> 4) & 0x3FFF) | (lbo << 16) | (sbo << 32) | (1ull << 46);
    }

    __device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
      asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
    }

    // ------------------------------------------------------------------------------
    // Persistent, software-pipelined kernel: one CTA per SM loops over 256-sample tiles
    // (tile = blockIdx.x, + gridDim.x, ...); every role runs ahead of the others through
    // mbarriers only, so the input stream never drains between tiles.
    //
    //   warps 0..7   converters   warp w owns TMEM lanes 32*(w&3).. of M-tile (w>>2): shared-memory
    //                             stage -> hi/lo split -> tcgen05.st (A operand)
    //   warps 8..11  epilogue     accumulators of tile L -> ramp -> global stores, then build the
    //                             gain matrix B of tile L+2 into the buffer tile L just released
    //   warp  12     TMA producer cp.async.bulk of [16 channels x 256 samples] stages, ring of
    //                             n_stages, continuous across tiles
    //   warp  13     MMA issuer   owns the TMEM allocation, issues tcgen05.mma kind::tf32
    //
    //   full[s]     producer   -> converters  stage s has landed               (tx bytes)
    //   empty[s]    converters -> producer    stage s has been read           (8 warps)
    //   a_full[b]   converters -> MMA         A buffer b is complete in TMEM  (8 warps)
    //   a_free[b]   MMA commit -> converters  the UMMAs reading A buffer b are done
    //   b_full[t]   epilogue   -> MMA         gain matrix buffer t is built   (4 warps)
    //   acc_full[t] MMA commit -> epilogue    accumulators t are complete (B buffer t is free too)
    //   acc_free[t] epilogue   -> MMA         accumulators t have been read out (4 warps)
    constexpr int kConvWarps = 8, kEpiWarps = 4;
    // two MMA issuer warps, one per M-tile: a single thread cannot issue the UMMAs of both fast
    // enough (each tcgen05.mma costs the issuing thread ~100 cycles of scalar work)
    constexpr int kProducerWarp = kConvWarps + kEpiWarps, kMmaWarp = kProducerWarp + 1;
    constexpr int kThreadsP = (kMmaWarp + kMT) * 32;  // 480
    constexpr int kMaxStages = 8;
    constexpr int kTmemColsP = 512;
    // TMEM columns: accumulators [2 tiles][main: kMT x 32 | corr: kMT x 32], then
    // A [2 buffers][kMT][hi 16 | lo 16]
    constexpr int kAccStride = 2 * kMT * kNP;
    constexpr int kColAP = 2 * kAccStride;
    constexpr int kBBufs = 2;  // gain-matrix buffers in shared memory
    constexpr int kABufs = 4;  // A-operand ring in TMEM (power of two)
    static_assert(kColAP + kABufs * kMT * 2 * kCh <= kTmemColsP, "TMEM budget");

    __global__ void __launch_bounds__(kThreadsP, 1)
        render_tc_persistent_kernel(const __grid_constant__ CUtensorMap in_map, int N, int M,
                                    float *__restrict__ out, long out_stride,
                                    const RenderItem *__restrict__ items, int n_tiles,
                                    const RenderSegment *__restrict__ segs,
                                    const float *__restrict__ gains, int n_stages) {
      extern __shared__ __align__(128) unsigned char smem_raw[];
      const int Kp = (N + kCh - 1) / kCh * kCh;  // K padded to whole stages
      const int n_iters = Kp / kCh;
      float *sB = reinterpret_cast<float *>(smem_raw);       // [2 tiles][hi | lo][Kp/4][32][4]
      // raw gain matrices of the tile being built: [2 tiles][point a | point b][raw_floats]
      const int raw_floats = (N * M + 31) / 32 * 32;
      float *sraw = sB + (size_t)kBBufs * 2 * Kp * kNP;
      const bool raw_tma = ((N * M) & 3) == 0;  // bulk copies need 16-byte multiples
      float *sx = sraw + (size_t)4 * raw_floats;                 // [n_stages][kCh][kTile]
      __shared__ __align__(8) uint64_t full[kMaxStages], empty[kMaxStages];
      __shared__ __align__(8) uint64_t a_full[kABufs], a_free[kABufs], b_full[kBBufs], raw_full[2], raw_free[2], acc_full[2], acc_free[2];
      __shared__ uint32_t tmem_base_s;

      const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

      if (tid == 0) {
        for (int s = 0; s < kMaxStages; s++) {
          mbar_init(&full[s], 1);
          mbar_init(&empty[s], kConvWarps);
        }
        for (int b = 0; b < kABufs; b++) {
          mbar_init(&a_full[b], kConvWarps);
          mbar_init(&a_free[b], kMT);
        }
        for (int b = 0; b < kBBufs; b++) mbar_init(&b_full[b], kEpiWarps);
        for (int b = 0; b < 2; b++) {
          mbar_init(&raw_full[b], 1);
          mbar_init(&raw_free[b], kEpiWarps);
          mbar_init(&acc_full[b], kMT);
          mbar_init(&acc_free[b], kEpiWarps);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      }
      if (warp == kMmaWarp) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                         smem_u32(&tmem_base_s)),
                     "r"(kTmemColsP)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
      }
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncthreads();
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint32_t tmem_base = tmem_base_s;

      if (warp == kProducerWarp) {
        // ------------------------------------------------------------ TMA producer
        // one tiled TMA load per stage: box = [kCh channels][kTile samples]; channels past N are
        // zero-filled by the TMA unit (and count towards the transaction bytes)
        if (lane == 0) {
          asm volatile("prefetch.tensormap [%0];" ::"l"(&in_map) : "memory");
          int slot = 0, L = 0;
          uint32_t phase = 0;  // parity of the `empty` completion to wait for, once the ring wrapped
          bool wrapped = false;
          // raw gain matrices (points pa, pb of the tile's segment) for the epilogue warps' build
          auto load_raw = [&](int tile, int rb) {
            const RenderSegment seg = segs[items[tile].seg];
            const uint32_t bytes = (uint32_t)(N * M * sizeof(float));
            mbar_expect_tx(&raw_full[rb], 2 * bytes);
            bulk_g2s(sraw + (size_t)(rb * 2) * raw_floats, gains + (size_t)seg.pa * N * M, bytes, &raw_full[rb]);
            bulk_g2s(sraw + (size_t)(rb * 2 + 1) * raw_floats, gains + (size_t)seg.pb * N * M, bytes,
                     &raw_full[rb]);
          };
          if (raw_tma)
            for (int b = 0; b < 2; b++)
              if (blockIdx.x + b * gridDim.x < n_tiles) load_raw(blockIdx.x + b * gridDim.x, b);
          for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, L++) {
            const int t0 = (int)items[tile].t0;
            for (int it = 0; it < n_iters; it++) {
              if (wrapped) mbar_wait(&empty[slot], phase);
              mbar_expect_tx(&full[slot], (uint32_t)(kCh * kTile * sizeof(float)));
              tma_load_2d(sx + (size_t)slot * kCh * kTile, &in_map, t0, it * kCh, &full[slot]);
              if (++slot == n_stages) {
                slot = 0;
                if (wrapped) phase ^= 1;
                wrapped = true;
              }
            }
            // raw gains of tile L + 2 go into the buffer the build of tile L has released (that
            // build finished before the MMAs of tile L could start, so this does not stall)
            const int next = tile + 2 * gridDim.x;
            if (raw_tma && next < n_tiles) {
              mbar_wait(&raw_free[L & 1], (uint32_t)((L >> 1) & 1));
              load_raw(next, L & 1);
            }
          }
        }
      } else if (warp >= kMmaWarp) {
        // ------------------------------------------------------------ MMA issuers (one per M-tile)
        // The whole warp runs the loop so that every operand is warp-uniform (uniform registers);
        // only the tcgen05 instructions themselves are predicated on one elected lane.
        const int j = __shfl_sync(0xffffffffu, warp, 0) - kMmaWarp;
        const uint32_t tm = __shfl_sync(0xffffffffu, tmem_base, 0);
        uint32_t leader;
        asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(leader));
        // instruction descriptors: D f32, A/B tf32, both K-major, M = 128, N = 64 / 32
        const uint32_t idesc_base = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(128 >> 4) << 24);
        const uint32_t idesc64 = idesc_base | ((uint32_t)(2 * kNP >> 3) << 17);
        const uint32_t idesc32 = idesc_base | ((uint32_t)(kNP >> 3) << 17);
        const uint64_t bdesc_base = b_desc(smem_u32(sB));
        const uint32_t b_tile_step = (uint32_t)((2 * Kp * kNP * sizeof(float)) >> 4);  // next B buffer
        int g = 0, L = 0;
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, L++) {
          const int ab = L & 1;
          // B of this tile: [Kp/4][64 rows][4]; rows 0..31 = hi, 32..63 = lo
          const int bb = L & 1;
          uint64_t bd = bdesc_base + (uint64_t)(bb * b_tile_step);
          const uint32_t acc = tm + ab * kAccStride + j * 2 * kNP;  // [main 32 | corr 32]
          mbar_wait(&b_full[bb], (uint32_t)((L >> 1) & 1));
          if (L >= 2) mbar_wait(&acc_free[ab], (uint32_t)(((L >> 1) - 1) & 1));
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          for (int it = 0; it < n_iters; it++, g++) {
            const int abuf = g & (kABufs - 1);
            mbar_wait(&a_full[abuf], (uint32_t)((g / kABufs) & 1));
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t a_hi = tm + kColAP + (abuf * kMT + j) * 2 * kCh;
            if (leader) {
#pragma unroll
              for (int ks = 0; ks < (EARB_RENDER_EXPERIMENT >= 2 ? 0 : kCh / 8); ks++) {
                // one K = 8 step covers two 4-wide chunks of B (2 x 1024 bytes)
                // [main | corr] (+)= x_hi * [g_hi | g_lo];  corr += x_lo * g_hi
                umma_tf32_ts(acc, a_hi + ks * 8, bd, idesc64, (it | ks) != 0);
                umma_tf32_ts(acc + kNP, a_hi + kCh + ks * 8, bd, idesc32, 1);
                bd += (2 * 2 * kNP * 16) >> 4;
              }
              umma_commit(&a_free[abuf]);  // arrives when every UMMA issued so far has finished
              if (it == n_iters - 1) umma_commit(&acc_full[ab]);
            }
            __syncwarp();
          }
        }
      } else if (warp < kConvWarps) {
        // ------------------------------------------------------------ converters
        const int q = warp & 3, j = warp >> 2;
        const uint32_t lane_base = tmem_base + ((uint32_t)(q * 32) << 16);
        const int row = j * 128 + q * 32 + lane;
        int g = 0, slot = 0;
        uint32_t full_phase = 0;
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
          for (int it = 0; it < n_iters; it++, g++) {
            const int abuf = g & (kABufs - 1);
            mbar_wait(&full[slot], full_phase);
            const float *src = sx + (size_t)slot * kCh * kTile + row;
#if EARB_RENDER_EXPERIMENT >= 1  // stream-only ceiling: converters consume stages without work
            if (g >= kABufs) mbar_wait(&a_free[abuf], (uint32_t)(((g / kABufs) - 1) & 1));
            __syncwarp();
            if (lane == 0) {
              mbar_arrive(&a_full[abuf]);
              mbar_arrive(&empty[slot]);
            }
            if (++slot == n_stages) {
              slot = 0;
              full_phase ^= 1;
            }
            continue;
#endif
            float hi[kCh], lo[kCh];
#pragma unroll
            for (int c = 0; c < kCh; c++) {
              // hi = x truncated to TF32, lo = x - hi exactly (13 significant bits; the tensor core
              // reads the top 19 bits of each operand).  Split error <= 2^-21 |x|, see DESIGN.md
              const float x = src[c * kTile];
              hi[c] = __uint_as_float(__float_as_uint(x) & 0xffffe000u);
              lo[c] = x - hi[c];
            }
            // the UMMAs of iteration g - kABufs must have finished reading this A buffer
            if (g >= kABufs) {
              mbar_wait(&a_free[abuf], (uint32_t)(((g / kABufs) - 1) & 1));
              asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            }
            const uint32_t a_col = kColAP + (abuf * kMT + j) * 2 * kCh;
            tmem_st16(lane_base + a_col, hi);
            tmem_st16(lane_base + a_col + kCh, lo);
            asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) {
              mbar_arrive(&a_full[abuf]);
              mbar_arrive(&empty[slot]);
            }
            if (++slot == n_stages) {
              slot = 0;
              full_phase ^= 1;
            }
          }
        }
      } else {
        // ------------------------------------------------------------ epilogue + gain matrices
        const int q = warp & 3;
        const int et = (warp - kConvWarps) * 32 + lane;  // 0..127
        const uint32_t lane_base = tmem_base + ((uint32_t)(q * 32) << 16);
        // Gain matrix B of a tile: rows 0..15 = start gains s, rows 16..31 = e - s, split hi/lo.
        // Padding rows (m >= M) and channels (k >= N) are zeroed once and never written.
        {
          float4 *z = reinterpret_cast<float4 *>(sB);
          const float4 zero = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
          for (int i = et; i < kBBufs * 2 * Kp * kNP / 4; i += kEpiWarps * 32) z[i] = zero;
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");
        // Lb = local index of the tile (its B, raw and accumulator buffers are Lb & 1)
        auto build_b = [&](int tile, int Lb) {
          const int ab = Lb & 1;
          const RenderSegment seg = segs[items[tile].seg];
          float *bhi = sB + (size_t)ab * 2 * Kp * kNP, *blo = bhi + kNP * 4;  // lo = rows 32..63
          const float *ga, *gb;
          if (raw_tma) {
            ga = sraw + (size_t)(ab * 2) * raw_floats;
            gb = ga + raw_floats;
            mbar_wait(&raw_full[ab], (uint32_t)((Lb >> 1) & 1));
          } else {
            ga = gains + (size_t)seg.pa * N * M;
            gb = gains + (size_t)seg.pb * N * M;
          }
          const int total = N * M;
          const bool ramp = seg.ramp != 0;
          // element r = k * M + m of the [N][M] matrices; (k, m) advance incrementally by 128
          constexpr int kStep = kEpiWarps * 32;
          const int step_k = kStep / M, step_m = kStep % M;
          int k = et / M, m = et - k * M;
#pragma unroll 4
          for (int r = et; r < total; r += kStep) {
            const float sv = ga[r];
            const float dv = ramp ? gb[r] - sv : 0.0f;
            const float shi = round_tf32(sv), dhi = round_tf32(dv);
            const int o = b_offset_floats(m, k);
            bhi[o] = shi;
            blo[o] = round_tf32(sv - shi);
            bhi[o + 16 * 4] = dhi;  // row 16 + m
            blo[o + 16 * 4] = round_tf32(dv - dhi);
            k += step_k;
            m += step_m;
            if (m >= M) {
              m -= M;
              k++;
            }
          }
          // make the generic-proxy writes visible to the tensor core (async proxy)
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          __syncwarp();
          if (lane == 0) {
            mbar_arrive(&b_full[ab]);
            if (raw_tma) mbar_arrive(&raw_free[ab]);
          }
        };
        for (int b = 0; b < kBBufs; b++)
          if (blockIdx.x + b * gridDim.x < n_tiles) build_b((int)(blockIdx.x + b * gridDim.x), b);

        int L = 0;
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, L++) {
          const int ab = L & 1;
          const RenderItem item = items[tile];
          const RenderSegment seg = segs[item.seg];
          const bool ramp = seg.ramp != 0;
          mbar_wait(&acc_full[ab], (uint32_t)((L >> 1) & 1));
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          const uint32_t acc = lane_base + ab * kAccStride;
#pragma unroll
          for (int j = 0; j < kMT; j++) {
            float a[32], b[32];
            tmem_ld32(acc + j * 2 * kNP, a);
            tmem_ld32(acc + j * 2 * kNP + kNP, b);
            const long t = item.t0 + j * 128 + q * 32 + lane;
            const float p = ramp ? (float)(t - seg.ramp_start) * seg.scale : 0.0f;
#pragma unroll
            for (int m = 0; m < 16; m++) {
              if (m < M) {
                const float sv = a[m] + b[m];
                const float dv = a[16 + m] + b[16 + m];
                if (EARB_RENDER_EXPERIMENT < 3) __stcs(out + (long)m * out_stride + t, ramp ? fmaf(p, dv, sv) : sv);
              }
            }
          }
          asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
          __syncwarp();
          if (lane == 0) mbar_arrive(&acc_free[ab]);
          // every UMMA of tile L has finished (acc_full), so its B buffer can take tile L + 2
          const int next = tile + kBBufs * gridDim.x;
          if (next < n_tiles) build_b(next, L + kBBufs);
        }
      }
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncthreads();
      if (warp == kMmaWarp) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemColsP)
                     : "memory");
      }
    }
  }  // namespace

  namespace {
    constexpr size_t kSmemBudget = 224 * 1024;  // dynamic shared memory we ask for at most
    // gain matrices: kBBufs tiles x (hi | lo), plus the raw staging [2 tiles][2 points][N x M]
    size_t b_bytes(int N, int M) {
      const int Kp = (N + kCh - 1) / kCh * kCh;
      const int raw_floats = (N * M + 31) / 32 * 32;
      return sizeof(float) * ((size_t)kBBufs * 2 * Kp * kNP + (size_t)4 * raw_floats);
    }
    int stages_for(int N, int M) {
      const size_t stage = sizeof(float) * kCh * kTile;
      const size_t left = kSmemBudget - b_bytes(N, M);
      return (int)std::min<size_t>(kMaxStages, left / stage);
    }
  }  // namespace

  size_t render_tc_smem_bytes(int N, int M) {
    return b_bytes(N, M) + sizeof(float) * (size_t)stages_for(N, M) * kCh * kTile;
  }

  bool render_tc_supported(int N, int M) { return M <= 16 && N <= 256; }

  namespace {
    typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                      const cuuint64_t *, const cuuint32_t *, const cuuint32_t *,
                                      CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                      CUtensorMapFloatOOBfill);
    // the driver entry point, resolved through the runtime (no link-time dependency on libcuda)
    EncodeTiledFn encode_tiled() {
      static EncodeTiledFn fn = [] {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
          p = nullptr;
        return reinterpret_cast<EncodeTiledFn>(p);
      }();
      return fn;
    }
  }  // namespace

  // d_items: the ramp tiles followed by the constant tiles (one persistent launch covers both).
  // Returns false when the TMA descriptor cannot be built (the caller then uses the FFMA path).
  bool launch_render_tc(int N, int M, const float *in, long in_stride, long n_samples, float *out,
                        long out_stride, const RenderItem *d_items, long n_tiles,
                        const RenderSegment *d_segs, const float *d_gains, int n_sms,
                        cudaStream_t stream) {
    if (n_tiles <= 0) return true;
    EncodeTiledFn encode = encode_tiled();
    if (!encode) return false;
    // the input as a [N channels][n_samples] float tensor with a row pitch of in_stride floats
    CUtensorMap map;
    const cuuint64_t dims[2] = {(cuuint64_t)n_samples, (cuuint64_t)N};
    const cuuint64_t strides[1] = {(cuuint64_t)in_stride * sizeof(float)};
    const cuuint32_t box[2] = {(cuuint32_t)kTile, (cuuint32_t)kCh};
    const cuuint32_t estr[2] = {1, 1};
    if (encode(&map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float *>(in), dims, strides, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return false;
    const size_t smem = render_tc_smem_bytes(N, M);
    static bool configured = false;
    if (!configured) {
      cudaFuncSetAttribute(render_tc_persistent_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           (int)kSmemBudget);
      configured = true;
    }
    const unsigned grid = (unsigned)std::min<long>(n_tiles, n_sms > 0 ? n_sms : 148);
    render_tc_persistent_kernel<<<grid, kThreadsP, smem, stream>>>(
        map, N, M, out, out_stride, d_items, (int)n_tiles, d_segs, d_gains, stages_for(N, M));
    return true;
  }
}  // namespace earb
