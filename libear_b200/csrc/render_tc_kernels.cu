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
#include <cstdlib>

#include "kernels.hpp"

namespace earb {
  namespace {
    constexpr int kTile = RENDER_TC_TILE;  // samples per CTA (2 UMMA M-tiles)
    constexpr int kMT = kTile / 128;       // M-tiles
    constexpr int kCh = 16;                // channels per stage = two K = 8 UMMA steps
    // NP = UMMA N of one product: NP/2 columns for S, NP/2 for D.  NP = 32: up to 16 outputs per
    // work unit, separate main / correction accumulators; NP = 64 ("wide"): up to 32 outputs, the
    // three products into one accumulator

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
    // four 8-column loads in flight, one wait
    __device__ __forceinline__ void tmem_ld8x4(uint32_t ta, uint32_t tb, uint32_t tc, uint32_t td, float (&a)[8],
                                               float (&b)[8], float (&c)[8], float (&d)[8]) {
      uint32_t r[4][8];
      const uint32_t addr[4] = {ta, tb, tc, td};
#pragma unroll
      for (int k = 0; k < 4; k++)
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                     : "=r"(r[k][0]), "=r"(r[k][1]), "=r"(r[k][2]), "=r"(r[k][3]), "=r"(r[k][4]), "=r"(r[k][5]),
                       "=r"(r[k][6]), "=r"(r[k][7])
                     : "r"(addr[k])
                     : "memory");
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
      for (int i = 0; i < 8; i++) {
        a[i] = __uint_as_float(r[0][i]);
        b[i] = __uint_as_float(r[1][i]);
        c[i] = __uint_as_float(r[2][i]);
        d[i] = __uint_as_float(r[3][i]);
      }
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

    // Gain-matrix operand B of a tile, K-major, non-swizzled canonical layout (core matrices of
    // 8 rows x 16 bytes), stored group-major: [8 row groups][Kp/4 K-chunks][8 rows][4 floats].
    // Row (= UMMA N index) r: 0..15 point a hi, 16..31 point b hi, 32..47 point a lo, 48..63
    // point b lo -- so each quarter is one contiguous block that is copied verbatim from the
    // per-point images render_prepare_points_kernel writes.
    __device__ __forceinline__ uint64_t b_desc(uint32_t smem_addr, int Kp) {
      const uint64_t lbo = 128 >> 4;                      // next 4-wide K chunk
      const uint64_t sbo = (uint64_t)(Kp / 4 * 128) >> 4;  // next 8-row group
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
    // EPI epilogue warps: 4 (one per TMEM lane quarter, draining both M-tiles of a tile in turn) or 8
    // (one per lane quarter and M-tile).  The read-out is a fixed cost per tile, so with few input
    // channels it sets the pace and the second set of warps pays (N = 16: 0.148 -> 0.102 ms); at
    // N >= 128 the 96-register cap of 640 threads costs 2 %, and the wide units spill under it.
    constexpr int kConvWarps = 8;
    // two MMA issuer warps, one per M-tile: a single thread cannot issue the UMMAs of both fast
    // enough (each tcgen05.mma costs the issuing thread ~100 cycles of scalar work)
    constexpr int threads_for(int epi) { return (kConvWarps + epi + 1 + kMT + 1) * 32; }  // 512 or 640
    constexpr int kMaxStages = 10;
    constexpr int kTmemColsP = 512;
    // TMEM columns: accumulators [2 tiles][main: kMT x 32 | corr: kMT x 32], then
    // A [2 buffers][kMT][hi 16 | lo 16]
    // TMEM accumulator columns.  NP = 32: per M-tile [main 32 | corr 32], two tile buffers.
    // NP = 64: per M-tile [main 64 | corr 64], ONE tile buffer (the MMAs of a tile then wait for the
    // epilogue of the tile before; keeping the two small cross terms out of the main accumulator is
    // what holds the 1e-6 bar -- accumulating all three products into one accumulator measured
    // 1.7e-6 of peak at N = 128).  Either way 256 columns, and the A ring takes the other 256.
    // NP = 48 (17..24 outputs: 9+10+3's feeds): per M-tile [main 48 | corr 48], TWO tile buffers =
    // 384 columns -- the MMAs of a tile run under the read-out of the tile before, which is what
    // paces the NP = 64 form -- and an A ring of two buffers in the remaining 128.
    constexpr int kBBufs = 2;  // gain-matrix buffers in shared memory
    template <int NP>
    struct TmemPlan {
      static constexpr int kAccBufs = NP == 64 ? 1 : 2;           // accumulator tile buffers
      static constexpr int kAccTile = kMT * 2 * NP;               // columns of one tile's accumulators
      static constexpr int kColA = kAccBufs * kAccTile;           // first column of the A ring
      static constexpr int kABufs = (kTmemColsP - kColA) / (kMT * 2 * kCh);  // A-operand ring (power of two)
      static_assert(kABufs == 2 || kABufs == 4, "A ring");
      static_assert(kColA + kABufs * kMT * 2 * kCh <= kTmemColsP, "TMEM budget");
    };
    constexpr int kMaxABufs = 4;

    template <int NP, int EPI>
    __global__ void __launch_bounds__(threads_for(EPI), 1)
        render_tc_persistent_kernel(const __grid_constant__ CUtensorMap in_map, int N, int M,
                                    float *__restrict__ out, long out_stride,
                                    const RenderItem *__restrict__ items, int n_tiles,
                                    const RenderSegment *__restrict__ segs,
                                    const float *__restrict__ images, int n_stages, int accumulate,
                                    int n_groups, const RenderBus bus, const RenderPair pair) {
      extern __shared__ __align__(128) unsigned char smem_raw[];
      constexpr int kAccBufs = TmemPlan<NP>::kAccBufs, kAccTile = TmemPlan<NP>::kAccTile;
      constexpr int kColAP = TmemPlan<NP>::kColA, kABufs = TmemPlan<NP>::kABufs;
      constexpr int kEpiWarps = EPI;
      constexpr int kProducerWarp = kConvWarps + kEpiWarps, kMmaWarp = kProducerWarp + 1;
      constexpr int kBLoadWarp = kMmaWarp + kMT;  // streams the gain-matrix images
      const int Kp = (N + kCh - 1) / kCh * kCh;  // K padded to whole stages
      const int n_iters = Kp / kCh;
      // bus mode: this rank's walk over the tiles starts rank / n_ranks of the way in (RenderBus)
      const int tile_shift = bus.n_ranks > 0 ? (int)((long)bus.rank * n_tiles / bus.n_ranks) : 0;
      // a work unit = one 256-sample tile x one group of 16 output channels; the groups of a tile
      // are adjacent units, so they run on neighbouring CTAs at the same time and the second read
      // of the tile's input is served by L2
      const int n_units = n_tiles * n_groups;
      // output groups per interpolator: with a second one riding along (RenderPair) the groups
      // [gpi, 2 gpi) of a tile are its
      const int gpi = pair.out2 ? n_groups >> 1 : n_groups;
      float *sB = reinterpret_cast<float *>(smem_raw);       // [2 tiles][8 groups][Kp/4][8][4]
      float *sx = sB + (size_t)kBBufs * 2 * Kp * NP;         // [n_stages][kCh][kTile]
      __shared__ __align__(8) uint64_t full[kMaxStages], empty[kMaxStages];
      __shared__ __align__(8) uint64_t a_full[kMaxABufs], a_free[kMaxABufs], b_full[kBBufs], acc_full[2], acc_free[2];
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
        for (int b = 0; b < kBBufs; b++) mbar_init(&b_full[b], 1);
        for (int b = 0; b < 2; b++) {
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
          for (int unit = blockIdx.x; unit < n_units; unit += gridDim.x, L++) {
          const int tile0 = unit / n_groups, grp = unit - tile0 * n_groups;
          const int tile = tile0 + tile_shift < n_tiles ? tile0 + tile_shift : tile0 + tile_shift - n_tiles;
          (void)grp;
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
          }
        }
      } else if (warp == kBLoadWarp) {
        // ------------------------------------------------------------ gain-matrix loader
        // B of tile L = four contiguous quarters: point a hi, point b hi, point a lo, point b lo,
        // copied verbatim from the per-point images.  Buffer L & 1 is free once every UMMA of tile
        // L - 2 has finished (acc_full).
        if (lane == 0) {
          const int qf = Kp * (NP / 2);                                  // floats per quarter: NP/16 row groups
          const uint32_t quarter = (uint32_t)(qf * sizeof(float));
          const size_t img = (size_t)Kp * NP;                            // floats per point image
          int L = 0;
          for (int unit = blockIdx.x; unit < n_units; unit += gridDim.x, L++) {
          const int tile0 = unit / n_groups, grp = unit - tile0 * n_groups;
          const int tile = tile0 + tile_shift < n_tiles ? tile0 + tile_shift : tile0 + tile_shift - n_tiles;
          (void)grp;
            const int ab = L & 1;
            if (L >= 2) mbar_wait(&acc_full[ab], (uint32_t)(((L >> 1) - 1) & 1));
            const RenderSegment seg = segs[items[tile].seg];
            const bool second = grp >= gpi;
            const int gl = second ? grp - gpi : grp;
            const float *im = second ? pair.images2 : images;
            const float *pa = im + ((size_t)seg.pa * gpi + gl) * img, *pb = im + ((size_t)seg.pb * gpi + gl) * img;
            float *dst = sB + (size_t)ab * 2 * Kp * NP;
            mbar_expect_tx(&b_full[ab], 4 * quarter);
            bulk_g2s(dst, pa, quarter, &b_full[ab]);
            bulk_g2s(dst + qf, pb, quarter, &b_full[ab]);
            bulk_g2s(dst + 2 * qf, pa + qf, quarter, &b_full[ab]);
            bulk_g2s(dst + 3 * qf, pb + qf, quarter, &b_full[ab]);
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
        const uint32_t idesc64 = idesc_base | ((uint32_t)(64 >> 3) << 17);
        const uint32_t idescW = idesc_base | ((uint32_t)(NP >> 3) << 17);  // wide units: N = NP
        const uint32_t idescW2 = idesc_base | ((uint32_t)((2 * NP) >> 3) << 17);
        const uint32_t idesc32 = idesc_base | ((uint32_t)(32 >> 3) << 17);
        const uint64_t bdesc_base = b_desc(smem_u32(sB), Kp);
        const uint32_t b_tile_step = (uint32_t)((2 * Kp * NP * sizeof(float)) >> 4);  // next B buffer
        int g = 0, L = 0;
        for (int unit = blockIdx.x; unit < n_units; unit += gridDim.x, L++) {
          const int tile0 = unit / n_groups, grp = unit - tile0 * n_groups;
          (void)grp;
          const int ab = L & 1;
          // B of this tile: rows 0..31 = [a | b] hi, 32..63 = [a | b] lo
          const int bb = L & 1;
          uint64_t bd = bdesc_base + (uint64_t)(bb * b_tile_step);
          const uint32_t acc = tm + (kAccBufs == 2 ? ab * kAccTile : 0) + j * 2 * NP;
          mbar_wait(&b_full[bb], (uint32_t)((L >> 1) & 1));
          if (kAccBufs == 2) {
            if (L >= 2) mbar_wait(&acc_free[ab], (uint32_t)(((L >> 1) - 1) & 1));  // tile L - 2 read out
          } else {
            if (L >= 1) mbar_wait(&acc_free[(L - 1) & 1], (uint32_t)(((L - 1) >> 1) & 1));  // tile L - 1
          }
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          for (int it = 0; it < n_iters; it++, g++) {
            const int abuf = g & (kABufs - 1);
            mbar_wait(&a_full[abuf], (uint32_t)((g / kABufs) & 1));
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t a_hi = tm + kColAP + (abuf * kMT + j) * 2 * kCh;
            if (leader) {
#pragma unroll
              for (int ks = 0; ks < (EARB_RENDER_EXPERIMENT >= 2 ? 0 : kCh / 8); ks++) {
                // one K = 8 step covers two 4-wide chunks of B (2 x 128 bytes within a row group)
                // [main | corr] (+)= x_hi * [g_hi | g_lo];  corr += x_lo * g_hi
                if (NP == 32) {
                  umma_tf32_ts(acc, a_hi + ks * 8, bd, idesc64, (it | ks) != 0);
                  umma_tf32_ts(acc + 32, a_hi + kCh + ks * 8, bd, idesc32, 1);
                } else {
                  // wide: [main | corr] (+)= x_hi * [g_hi | g_lo] in ONE UMMA of N = 2 NP (the lo rows
                  // follow the hi rows in the B image, the correction columns the main ones in TMEM),
                  // then corr += x_lo * g_hi.  Two instructions per K step instead of three: the issuing
                  // thread, not the tensor pipe, paces this kernel.
                  umma_tf32_ts(acc, a_hi + ks * 8, bd, idescW2, (it | ks) != 0);
                  umma_tf32_ts(acc + NP, a_hi + kCh + ks * 8, bd, idescW, 1);
                }
                bd += (2 * 128) >> 4;
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
        for (int unit = blockIdx.x; unit < n_units; unit += gridDim.x) {
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
              // reads the top 19 bits of each operand).  Split error <= 2^-21 |x|, see DESIGN.md.
              // (Rounding hi instead was measured: -2.7 % throughput, same error against libear --
              // at N >= 128 the difference is dominated by the fp32 summation order on both sides.)
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
        // EPI == 8: this warp drains M-tile j0 only; EPI == 4: both, one after the other
        const int q = warp & 3, j0 = (warp - kConvWarps) >> 2;
        constexpr int kJ = kMT * 4 / EPI;  // M-tiles per epilogue warp
        const uint32_t lane_base = tmem_base + ((uint32_t)(q * 32) << 16);
        int L = 0;
        for (int unit = blockIdx.x; unit < n_units; unit += gridDim.x, L++) {
          const int tile0 = unit / n_groups, grp = unit - tile0 * n_groups;
          const int tile = tile0 + tile_shift < n_tiles ? tile0 + tile_shift : tile0 + tile_shift - n_tiles;
          (void)grp;
          const int ab = L & 1;
          const RenderItem item = items[tile];
          const RenderSegment seg = segs[item.seg];
          const bool ramp = seg.ramp != 0;
          constexpr int MO = NP / 2;               // output channels per work unit
          const bool second = grp >= gpi;
          const int gl = second ? grp - gpi : grp;
          const int Mg = min(MO, M - gl * MO);     // output channels of this group
          float *const og = (second ? pair.out2 : out) + (long)gl * MO * out_stride;
          const bool to_bus = bus.n_ranks > 0;
          // bus mode: owner and offset of the tile's first sample, once per tile in 32-bit arithmetic
          // (a 64-bit division per sample group doubled the epilogue's instruction count: the last
          // group of an accumulating mix ran 0.92 instead of 0.65 ms)
          uint32_t bus_owner0 = 0, bus_off0 = 0;
          if (to_bus) {
            const uint32_t bt0 = (uint32_t)((long long)item.t0 + bus.t_origin), sl = (uint32_t)bus.slice_len;
            bus_owner0 = bt0 / sl;
            bus_off0 = bt0 - bus_owner0 * sl;
          }
          // accumulate mode: the bus values this thread will add to are fetched before waiting for
          // the accumulators, so their latency hides behind the MMAs of this tile
          // (wide units read the bus inside the loop instead: 64 more registers would not fit)
          float prev[kJ][16];
          if (accumulate && NP == 32) {
#pragma unroll
            for (int jj = 0; jj < kJ; jj++) {
              const long tp = item.t0 + (j0 + jj) * 128 + q * 32 + lane;
#pragma unroll
              for (int m = 0; m < 16; m++) prev[jj][m] = (m < Mg) ? __ldcs(og + (long)m * out_stride + tp) : 0.0f;
            }
          }
          mbar_wait(&acc_full[ab], (uint32_t)((L >> 1) & 1));
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          const uint32_t acc = lane_base + (kAccBufs == 2 ? ab * kAccTile : 0);
#pragma unroll
          for (int jj = 0; jj < kJ; jj++) {
            const int j = j0 + jj;
            const long t = item.t0 + j * 128 + q * 32 + lane;
            const float p = ramp ? (float)(t - seg.ramp_start) * seg.scale : 0.0f;
            // where this sample's results go: the caller's buffer, or this rank's slot in the
            // staging memory of the GPU that owns the sample's time slice (a peer store)
            float *dst = og + t;
            long dst_stride = out_stride;
            if (to_bus) {
              // a tile is at most one slice long: it wraps into the next owner's slice at most once
              uint32_t off = bus_off0 + (uint32_t)(j * 128 + q * 32 + lane), owner = bus_owner0;
              if (off >= (uint32_t)bus.slice_len) off -= (uint32_t)bus.slice_len, owner++;
              dst = bus.slot[owner] + (long long)grp * MO * bus.slice_len + off;
              dst_stride = (long)bus.slice_len;
            }
            if (NP == 32) {
              // main and correction accumulators, each [16 x g_a (start point) | 16 x g_b (end point)]
              float a[32], b[32];
              tmem_ld32(acc + j * 64, a);
              tmem_ld32(acc + j * 64 + 32, b);
#pragma unroll
              for (int m = 0; m < 16; m++) {
                if (m < Mg) {
                  const float sa = a[m] + b[m];
                  const float sb = a[16 + m] + b[16 + m];
                  float r = ramp ? fmaf(p, sb - sa, sa) : sa;
                  if (accumulate) r += prev[jj][m];  // mix into the bus
                  if (EARB_RENDER_EXPERIMENT < 3) __stcs(dst + (long)m * dst_stride, r);
                }
              }
            } else {
              // wide: main and correction accumulators, each [32 x g_a | 32 x g_b], read out 8 output
              // channels at a time (32 live values: the 96-register form of the kernel has room for it)
#pragma unroll
              for (int h = 0; h < NP / 16; h++) {
                if (h * 8 < Mg) {
                  float mix[8];
                  if (accumulate) {
#pragma unroll
                    for (int m = 0; m < 8; m++)
                      mix[m] = (h * 8 + m < Mg) ? __ldcs(og + (long)(h * 8 + m) * out_stride + t) : 0.0f;
                  }
                  float sa[8], sb[8], ca[8], cb[8];
                  tmem_ld8x4(acc + j * 2 * NP + h * 8, acc + j * 2 * NP + NP / 2 + h * 8, acc + j * 2 * NP + NP + h * 8,
                             acc + j * 2 * NP + NP + NP / 2 + h * 8, sa, sb, ca, cb);
#pragma unroll
                  for (int m = 0; m < 8; m++) {
                    if (h * 8 + m < Mg) {
                      const float va = sa[m] + ca[m], vb = sb[m] + cb[m];
                      float r = ramp ? fmaf(p, vb - va, va) : va;
                      if (accumulate) r += mix[m];
                      if (EARB_RENDER_EXPERIMENT < 3) __stcs(dst + (long)(h * 8 + m) * dst_stride, r);
                    }
                  }
                }
              }
            }
          }
          asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
          __syncwarp();
          if (lane == 0) mbar_arrive(&acc_free[ab]);
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
    // Per-point image of a gain matrix g[N][M] in the layout the UMMA reads (see b_desc):
    // [hi | lo][2 row groups][Kp/4][8 rows][4], hi = g rounded to TF32, lo = TF32(g - hi); rows
    // m >= M and channels k >= N are zero.  Written once per set_points.
    __global__ void render_prepare_points_kernel(int N_total, int k0, int N, int M, int Kp, int NP, int n_groups,
                                                 long n_points,
                                                 const float *__restrict__ gains,
                                                 float *__restrict__ images) {
      const int half = Kp * (NP / 2);  // floats of the hi (or lo) part of one image
      const long total = n_points * n_groups * half;
      for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        // i enumerates (point, output group, row group, chunk, row-in-group, k-in-chunk) of the hi half
        const int e = (int)(i % half);
        const long pg = i / half;  // point * n_groups + output group
        const long p = pg / n_groups;
        const int og = (int)(pg - p * n_groups);
        const int grp = e / (Kp * 8), rem = e % (Kp * 8);
        const int chunk = rem / 32, r8 = (rem % 32) / 4, k4 = rem % 4;
        const int m = og * (NP / 2) + grp * 8 + r8, k = chunk * 4 + k4;
        float hi = 0.0f, lo = 0.0f;
        if (m < M && k < N) {
          const float g = gains[(p * N_total + k0 + k) * M + m];
          hi = round_tf32(g);
          lo = round_tf32(g - hi);
        }
        float *img = images + pg * ((long)Kp * NP);
        img[e] = hi;
        img[half + e] = lo;
      }
    }
  }  // namespace

  // How a render of N inputs into M outputs is cut up for the kernel:
  //   M <= 16: NP = 32 units (main + correction accumulators), input groups of up to 256 channels;
  //   M  > 16: NP = 64 "wide" units (up to 32 outputs, the input is converted once; the gain
  //            matrices of one tile are 2 * Kp * NP floats and two of them must leave room for the
  //            input ring, hence input groups of up to 128 channels).  With eight epilogue warps
  //            this beats NP = 32 with two output groups at every N (N = 256, M = 24: 0.80 vs
  //            1.07 ms; N = 512: 1.78 vs 2.15 ms).
  // Input groups run as one launch each, the later ones accumulating into the output.
  // EARB200_RENDER_WIDE=0 keeps NP = 32 for every shape (for comparison).
  namespace {
    struct TcShape {
      int NP, max_in, n_out_groups;
    };
    TcShape tc_shape(int N, int M) {
      static const int wide = [] {
        const char *e = std::getenv("EARB200_RENDER_WIDE");
        return e ? std::atoi(e) : 2;
      }();
      if (M > 16 && M <= 24 && wide > 1) return {48, 128, 1};
      if (M > 16 && wide > 0) return {64, 128, (M + 31) / 32};
      return {32, 256, (M + 15) / 16};
    }
    int padded_k(int n) { return (n + kCh - 1) / kCh * kCh; }
  }  // namespace

  // floats per interpolation point: one image per (input group, output group)
  size_t render_tc_image_floats(int N, int M) {
    const TcShape sh = tc_shape(N, M);
    size_t total = 0;
    for (int k0 = 0; k0 < N; k0 += sh.max_in)
      total += (size_t)padded_k(std::min(sh.max_in, N - k0)) * sh.NP * sh.n_out_groups;
    return total;
  }

  // image layout: [input group][point][output group][Kp_g * NP]
  void launch_render_prepare_points(int N, int M, long n_points, const float *d_gains, float *d_images,
                                    cudaStream_t stream) {
    if (n_points <= 0) return;
    const TcShape sh = tc_shape(N, M);
    for (int k0 = 0; k0 < N; k0 += sh.max_in) {
      const int Ng = std::min(sh.max_in, N - k0), Kp = padded_k(Ng);
      const long total = n_points * sh.n_out_groups * Kp * (sh.NP / 2);
      const unsigned blocks = (unsigned)std::min<long>((total + 255) / 256, 148 * 16);
      render_prepare_points_kernel<<<blocks, 256, 0, stream>>>(N, k0, Ng, M, Kp, sh.NP, sh.n_out_groups, n_points,
                                                                d_gains, d_images);
      d_images += (size_t)n_points * sh.n_out_groups * Kp * sh.NP;
    }
  }

  namespace {
    constexpr size_t kSmemBudget = 224 * 1024;  // dynamic shared memory we ask for at most
    // gain matrices of one input group: kBBufs tiles x 2 NP rows x Kp
    size_t b_bytes(int Ng, int NP) { return sizeof(float) * (size_t)kBBufs * 2 * padded_k(Ng) * NP; }
    int stages_for(int Ng, int NP) {
      const size_t stage = sizeof(float) * kCh * kTile;
      const size_t left = kSmemBudget - b_bytes(Ng, NP);
      return (int)std::min<size_t>(kMaxStages, left / stage);
    }
    size_t smem_for(int Ng, int NP) {
      return b_bytes(Ng, NP) + sizeof(float) * (size_t)stages_for(Ng, NP) * kCh * kTile;
    }
  }  // namespace

  size_t render_tc_smem_bytes(int N, int M) {
    const TcShape sh = tc_shape(N, M);
    return smem_for(std::min(N, sh.max_in), sh.NP);
  }

  bool render_tc_supported(int N, int M) { return M <= EARB_MAX_OUT && N >= 1; }

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
                        const RenderSegment *d_segs, const float *d_images, long n_points, int n_sms,
                        int accumulate, const RenderBus &bus, cudaStream_t stream, const RenderPair &pair) {
    if (n_tiles <= 0) return true;
    EncodeTiledFn encode = encode_tiled();
    if (!encode) return false;
    static std::atomic<unsigned long long> configured{0};
    if (first_use_on_current_device(configured)) {
      cudaFuncSetAttribute(render_tc_persistent_kernel<32, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           (int)kSmemBudget);
      cudaFuncSetAttribute(render_tc_persistent_kernel<32, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           (int)kSmemBudget);
      cudaFuncSetAttribute(render_tc_persistent_kernel<64, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           (int)kSmemBudget);
      cudaFuncSetAttribute(render_tc_persistent_kernel<48, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           (int)kSmemBudget);
    }
    const TcShape sh = tc_shape(N, M);
    const int n_groups = sh.n_out_groups * (pair.out2 ? 2 : 1);
    RenderPair gpair = pair;
    const unsigned grid = (unsigned)std::min<long>(n_tiles * n_groups, n_sms > 0 ? n_sms : 148);
    for (int k0 = 0; k0 < N; k0 += sh.max_in) {
      const int Ng = std::min(sh.max_in, N - k0);
      // the input group as a [Ng channels][n_samples] float tensor with a row pitch of in_stride floats
      CUtensorMap map;
      const cuuint64_t dims[2] = {(cuuint64_t)n_samples, (cuuint64_t)Ng};
      const cuuint64_t strides[1] = {(cuuint64_t)in_stride * sizeof(float)};
      const cuuint32_t box[2] = {(cuuint32_t)kTile, (cuuint32_t)kCh};
      const cuuint32_t estr[2] = {1, 1};
      if (encode(&map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float *>(in + (size_t)k0 * in_stride), dims,
                 strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                 CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
        return false;
      const int acc_mode = (accumulate || k0 > 0) ? 1 : 0;
      // input groups accumulate into `out`; only the last one stores to the bus slots
      RenderBus gbus = bus;
      if (k0 + sh.max_in < N) gbus.n_ranks = 0;
      if (sh.NP == 48)
        render_tc_persistent_kernel<48, 8><<<grid, threads_for(8), smem_for(Ng, 48), stream>>>(
            map, Ng, M, out, out_stride, d_items, (int)n_tiles, d_segs, d_images, stages_for(Ng, 48), acc_mode,
            n_groups, gbus, gpair);
      else if (sh.NP == 64)  // wide units: the read-out of 2 x 64 accumulator columns always wants eight warps
        render_tc_persistent_kernel<64, 8><<<grid, threads_for(8), smem_for(Ng, 64), stream>>>(
            map, Ng, M, out, out_stride, d_items, (int)n_tiles, d_segs, d_images, stages_for(Ng, 64), acc_mode,
            n_groups, gbus, gpair);
      else if (Ng <= 64)  // few input channels: the read-out sets the pace, eight epilogue warps
        render_tc_persistent_kernel<32, 8><<<grid, threads_for(8), smem_for(Ng, 32), stream>>>(
            map, Ng, M, out, out_stride, d_items, (int)n_tiles, d_segs, d_images, stages_for(Ng, 32), acc_mode,
            n_groups, gbus, gpair);
      else
        render_tc_persistent_kernel<32, 4><<<grid, threads_for(4), smem_for(Ng, 32), stream>>>(
            map, Ng, M, out, out_stride, d_items, (int)n_tiles, d_segs, d_images, stages_for(Ng, 32), acc_mode,
            n_groups, gbus, gpair);
      d_images += (size_t)n_points * sh.n_out_groups * padded_k(Ng) * sh.NP;
      if (gpair.images2) gpair.images2 += (size_t)n_points * sh.n_out_groups * padded_k(Ng) * sh.NP;
    }
    return true;
  }
}  // namespace earb
