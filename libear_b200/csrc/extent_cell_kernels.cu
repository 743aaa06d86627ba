// K2 core, second generation: cell-culled weight generation feeding tcgen05 (kind::f16) UMMAs.
//
// Replaces the hot loop of the extent panner, PolarExtentCoreSimd::run
// (ref src/object_based/polar_extent_simd.hpp:97-135): for every spreading pass
//     results[s] = sum_p w_p * G[s][p],   p over the virtual sources,
// i.e. R[passes x S] = W[passes x P] * G^T[P x S] with W generated on the fly.
//
// The weight of a virtual source is a function of its angular distance d to the "spine" of the
// stadium (an arc of the equator of the pass's rotated frame; a point for circular extents):
//     w = clamp(m * d + c, 0, 1)        (ref polar_extent_simd.hpp:39-95, polar_extent.cpp:176-245)
// and d is 1-Lipschitz on the sphere.  The virtual sources are therefore clustered (once, on the
// host) into compact CELLS of 16 with a bounding cap (centre, radius r): if the weight function
// evaluated at the centre is >= 1 + |m| r every weight of the cell is exactly 1, if it is
// <= -|m| r every weight is exactly 0 -- the reference's own all-ones / all-zeros block shortcut
// (polar_extent_simd.hpp:111-126), decided from one evaluation instead of sixteen.  On random
// metadata three quarters of all (pass, cell) pairs are uniform (tools/exp_cells.py).
//
// One persistent CTA per SM walks its passes in tiles of 512 (four M = 128 UMMA tiles):
//   warps 0..15  phase 1  thread = pass: pass constants -> shared memory, the weight function at
//                         the 104 cell centres -> per-cell bit masks (mixed / all-ones) by ballot.
//                phase 2  "evaluators": a batch = 32 (pass, cell) items of ONE cell whose class is
//                         mixed (lane = pass, taken from the cell's list); 16 weights per lane,
//                         split into fp16 hi + lo and stored into the A stage of the cell in
//                         shared memory (UMMA canonical K-major layout).
//                epilogue thread = pass: tcgen05.ld of the accumulator row, normalise, store.
//   warps 16,17  fillers  per cell: write the rows of the uniform passes (0 or 1) into the A
//                         stage and the list of mixed passes for the evaluators.
//   warp  18     MMA      per cell and M-tile: [main | corr] (+)= A_hi * [G_hi | G_lo]  (N = 48)
//                                              corr          +=  A_lo * G_hi           (N = 32)
//   warp  19     producer streams [gain image | point coordinates] per cell through a ring
//                         (cp.async.bulk); evaluators read the coordinates from the same stage.
// fp16 hi + lo carries 22 bits of the weight and of the gain (|error| <= 6e-8 absolute; weights
// and gains lie in [0, 1]); the three products accumulate in fp32 in TMEM.
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

#include "kernels.hpp"

namespace earb {
  namespace {
    constexpr int kCell = EXTENT_CELL_POINTS;  // 16 points per cell = K of one kind::f16 UMMA
    constexpr int kMT = 4;                     // M-tiles per tile
    constexpr int kTileRows = kMT * 128;
    constexpr int kEvalWarps = kTileRows / 32;  // 16
    constexpr int kFillWarps = 2;
    constexpr int kMmaWarp = kEvalWarps + kFillWarps, kProdWarp = kMmaWarp + 1;
    constexpr int kThreads = (kProdWarp + 1) * 32;  // 640
    constexpr int kD = 4;                            // A ring depth in cells
    constexpr int kBCells = 4;                       // cells per stream stage
    constexpr int kBStages = 4;
    constexpr int kRows = 48;                        // B rows: 24 G_hi + 24 G_lo
    constexpr int kCellBBytes = kRows * kCell * 2;   // 1536
    constexpr int kCellPtBytes = (kCell / 2) * 24;   // 8 x (float4 xy pair + float2 z pair) = 192
    constexpr int kCellBytes = kCellBBytes + kCellPtBytes;  // 1728
    constexpr int kBStageBytes = kBCells * kCellBytes;      // 6912
    constexpr int kAPlaneBytes = 128 * kCell * 2;           // one M-tile, one plane: 4096
    constexpr int kAStageBytes = kMT * 2 * kAPlaneBytes;    // 32768
    constexpr int kConstBytes = 64;                         // pass constants per row
    constexpr int kAccCols = 64;                            // per M-tile: [main 24 | corr 24 | spill 8 | -]
    constexpr int kTmemCols = kMT * kAccCols;               // 256
    static_assert(EARB_MAX_SPK <= 24, "B row packing assumes at most 24 loudspeakers");
    static_assert(kCellBytes % 16 == 0, "stage pieces must stay 16-byte aligned");
    static_assert(EXTENT_TC_STAGE_POINTS % (kBCells * kCell) == 0, "grid padding covers whole stages");
    // classification margin in weight units (|m| * 0.1 degree): covers the float rounding of the
    // centre evaluation, which is worst where acos is steep (angle ~ 0 or pi: 4.5e-4 rad)
    constexpr double kClassMargin = 0.01;

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
    __device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
      asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
    }
    __device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
      asm volatile(
          "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
              smem_u32(dst)),
          "l"(src), "r"(bytes), "r"(smem_u32(bar))
          : "memory");
    }
    // waits for the phase with the given parity; a protocol error traps after ~2 s instead of
    // hanging the GPU
    __device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
      uint32_t done;
      asm volatile(
          "{\n\t.reg .pred p;\n\t"
          "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
          "selp.u32 %0, 1, 0, p;\n\t}"
          : "=r"(done)
          : "r"(smem_u32(bar)), "r"(parity)
          : "memory");
      if (done) return;
      const long long t0 = clock64();
      do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, 0x989680;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
        if (!done && clock64() - t0 > 4000000000LL) __trap();
      } while (!done);
    }
    __device__ __forceinline__ void fence_proxy_async() {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
      uint32_t r[16];
      asm volatile(
          "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
          "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
          : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
            "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]),
            "=r"(r[15])
          : "r"(taddr)
          : "memory");
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
      for (int i = 0; i < 16; i++) v[i] = __uint_as_float(r[i]);
    }
    // D[tmem] (+)= A[smem desc] * B[smem desc], kind::f16 (fp16 operands, fp32 accumulate)
    __device__ __forceinline__ void umma_f16_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc,
                                                uint32_t idesc, uint32_t accumulate) {
      asm volatile(
          "{\n\t.reg .pred p;\n\t"
          "setp.ne.b32 p, %4, 0;\n\t"
          "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
          ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
          : "memory");
    }
    __device__ __forceinline__ void umma_commit(uint64_t *bar) {
      asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                       smem_u32(bar))
                   : "memory");
    }
    // K-major, no swizzle, 16-bit elements: a core matrix is 8 rows x 16 bytes (8 elements of K);
    // the two core matrices of one K = 16 step are 128 bytes apart (LBO), 8-row groups 256 bytes
    // (SBO).  Same layout for A (128 rows) and B (48 rows).
    __device__ __forceinline__ uint64_t kmajor_desc(uint32_t smem_addr) {
      const uint64_t lbo = 128 >> 4, sbo = 256 >> 4;
      return (uint64_t)((smem_addr >> 4) & 0x3FFF) | (lbo << 16) | (sbo << 32) | (1ull << 46);
    }

    // pass constants (PolarExtentCoreContext, ref polar_extent_core.hpp:12-42).  Circular extents
    // (ref weight_circle, polar_extent_simd.hpp:61-66) run through the stadium code with constants
    // that reduce it to the circle test exactly: sin_thr = +inf disables the straight-line part
    // and (cc0, cc1) = (0, 1) makes circle_dot = fma(|x'|, 0, y' * 1) = y'.
    struct PassRegs {
      float b[9];
      float ct0, ct1, cc0, cc1;
      float sin_thr;  // straight-line part when circle_test_dot >= sin_thr (0, or +inf if circular)
      float m, c;     // fade: weight = m * angle + c between the start and end angles
      float a_sin;    // m * pi/2 + c
    };

    typedef unsigned long long f32x2;
    __device__ __forceinline__ f32x2 pk(float a, float b) {
      f32x2 r;
      asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
      return r;
    }
    __device__ __forceinline__ void upk(f32x2 v, float &a, float &b) {
      asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v));
    }
    __device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
      f32x2 r;
      asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
      return r;
    }
    __device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) {
      f32x2 r;
      asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
      return r;
    }
    __device__ __forceinline__ f32x2 bc(float a) { return pk(a, a); }

    // The weight function m * angle + c of two virtual sources, BEFORE the clamp to [0, 1]
    // (ref polar_extent_simd.hpp:39-95).  The reference returns 1 for angle <= start, 0 for
    // angle >= end and the fade m * angle + c in between; the fade is 1 at `start`, 0 at `end`
    // and monotone, and the sentinels libear uses when an angle passes pi (pi/2 for the sine test)
    // only ever disable a comparison whose fade value is already inside [0, 1]
    // (polar_extent.cpp:176-199), so the whole function is clamp(m * angle + c, 0, 1).
    // angle = acos(v) on the circle part, asin(|z'|) on the straight part, both from
    // acos(|v|) = sqrt(u) * Q7(u), u = 1 - |v| (Abramowitz & Stegun 4.4.46 re-expanded in u,
    // |error| <= 2.2e-8); the straight part enters as v = -|z'| so that both parts are
    //     m * (-copysign(asin|v|, v)) + k,   k = c (straight) or m pi/2 + c (circle).
    // Packed fp32 (FFMA2 / FMUL2): two sources per issue slot for the rotation and the polynomial.
    __device__ __forceinline__ void weight_fn2(const PassRegs &c, f32x2 X, f32x2 Y, f32x2 Z, float &t0_out,
                                               float &t1_out) {
      const f32x2 xt = fma2(X, bc(c.b[0]), fma2(Y, bc(c.b[1]), mul2(Z, bc(c.b[2]))));
      const f32x2 yt = fma2(X, bc(c.b[3]), fma2(Y, bc(c.b[4]), mul2(Z, bc(c.b[5]))));
      const f32x2 zt = fma2(X, bc(c.b[6]), fma2(Y, bc(c.b[7]), mul2(Z, bc(c.b[8]))));
      float x0, x1, ty0, ty1, tc0, tc1, z0, z1;
      upk(xt, x0, x1);
      upk(mul2(yt, bc(c.ct1)), ty0, ty1);
      upk(mul2(yt, bc(c.cc1)), tc0, tc1);
      upk(zt, z0, z1);
      const float ctd0 = fmaf(fabsf(x0), c.ct0, ty0), ctd1 = fmaf(fabsf(x1), c.ct0, ty1);
      const float cd0 = fmaf(fabsf(x0), c.cc0, tc0), cd1 = fmaf(fabsf(x1), c.cc0, tc1);
      const bool s0 = ctd0 >= c.sin_thr, s1 = ctd1 >= c.sin_thr;
      const float v0 = s0 ? -fabsf(z0) : cd0, v1 = s1 ? -fabsf(z1) : cd1;
      // rounding can leave |v| a few ulp above 1: u saturates
      const float u0 = __saturatef(__fsub_rn(1.0f, fabsf(v0))), u1 = __saturatef(__fsub_rn(1.0f, fabsf(v1)));
      const f32x2 U = pk(u0, u1);
      f32x2 p = fma2(bc(0.0012628264f), U, bc(-0.0021685062f));
      p = fma2(p, U, bc(0.0035814755f));
      p = fma2(p, U, bc(0.0013143348f));
      p = fma2(p, U, bc(0.0082738195f));
      p = fma2(p, U, bc(0.0264650658f));
      p = fma2(p, U, bc(0.1178537458f));
      p = fma2(p, U, bc(1.4142135382f));
      float q0, q1;
      asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(q0) : "f"(u0));
      asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(q1) : "f"(u1));
      // nq = sqrt(u) * p - pi/2 = -asin(|v|)
      const f32x2 nq = fma2(pk(q0, q1), p, bc(-1.57079632679490f));
      float n0, n1;
      upk(nq, n0, n1);
      // asin(|v|) carrying the inverted sign of v: one XOR with the sign bit of v
      const float a0 = __uint_as_float(__float_as_uint(n0) ^ (__float_as_uint(v0) & 0x80000000u));
      const float a1 = __uint_as_float(__float_as_uint(n1) ^ (__float_as_uint(v1) & 0x80000000u));
      t0_out = fmaf(c.m, a0, s0 ? c.c : c.a_sin);
      t1_out = fmaf(c.m, a1, s1 ? c.c : c.a_sin);
    }

    // shared-memory record of the pass constants: 16 floats; the 16-byte pieces are XOR-swizzled
    // with the row so that lanes reading scattered rows spread over the banks
    __device__ __forceinline__ uint32_t const_piece(uint32_t row, uint32_t i) {
      return row * kConstBytes + ((i ^ ((row >> 1) & 3u)) << 4);
    }

    // Barriers (parities follow counters that run on across tiles)
    //   filled[s]    filler -> evaluators        uniform rows written, list of mixed passes ready
    //   a_full[s]    16 evaluator warps -> MMA   every mixed row of the cell is in the A stage
    //   a_free[s]    MMA commit -> filler        the UMMAs reading the A stage have finished
    //   b_full[t]    producer -> MMA, evaluators stream stage t has landed (tx bytes)
    //   b_empty[t]   MMA commit -> producer      the UMMAs reading stream stage t have finished
    //   acc_full     MMA commit -> epilogue      every UMMA of the tile has finished
    //   masks_ready  16 evaluator warps -> fillers  phase 1 of the tile is complete
    __global__ void __launch_bounds__(kThreads, 1)
        extent_cell_kernel(int n_cells, int n_spk, const unsigned char *__restrict__ g_stream,
                           const float *__restrict__ g_centres, const unsigned int *__restrict__ n_active_ptr,
                           const ExtentPass *__restrict__ passes, float *__restrict__ results) {
      extern __shared__ __align__(128) unsigned char smem_raw[];
      unsigned char *sA = smem_raw;                                       // [kD][kAStageBytes]
      unsigned char *sB = sA + kD * kAStageBytes;                         // [kBStages][kBStageBytes]
      unsigned char *sConst = sB + kBStages * kBStageBytes;               // [kTileRows][64]
      uint32_t *sMixed = reinterpret_cast<uint32_t *>(sConst + kTileRows * kConstBytes);  // [n_cells][16]
      uint32_t *sOnes = sMixed + n_cells * kEvalWarps;                    // [n_cells][16]
      float *sCtr = reinterpret_cast<float *>(sOnes + n_cells * kEvalWarps);  // [n_cells / 2][12]
      unsigned short *sList = reinterpret_cast<unsigned short *>(sCtr + (n_cells / 2) * 12);  // [kD][kTileRows]
      __shared__ __align__(8) uint64_t filled[kD], a_full[kD], a_free[kD];
      __shared__ __align__(8) uint64_t b_full[kBStages], b_empty[kBStages], acc_full, masks_ready;
      __shared__ uint32_t s_nmixed[kD];
      __shared__ uint32_t tmem_base_s;

      const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
      const int n_active = (int)*n_active_ptr;
      // every CTA owns one contiguous range of passes, equal to within one warp
      const int per_cta = (((n_active + (int)gridDim.x - 1) / (int)gridDim.x) + 31) & ~31;
      const int p_begin = min(n_active, (int)blockIdx.x * per_cta), p_end = min(n_active, p_begin + per_cta);
      const int n_tiles = (p_end - p_begin + kTileRows - 1) / kTileRows;
      const int n_bstages = n_cells / kBCells;
      if (n_tiles == 0) return;  // no passes for this CTA (small batches): nothing to set up

      for (int j = tid; j < (n_cells / 2) * 12; j += kThreads) sCtr[j] = g_centres[j];
      if (tid == 0) {
        for (int s = 0; s < kD; s++) {
          mbar_init(&filled[s], 1);
          mbar_init(&a_full[s], kEvalWarps);
          mbar_init(&a_free[s], 1);
        }
        for (int s = 0; s < kBStages; s++) {
          mbar_init(&b_full[s], 1);
          mbar_init(&b_empty[s], 1);
        }
        mbar_init(&acc_full, 1);
        mbar_init(&masks_ready, kEvalWarps);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      }
      if (warp == kMmaWarp) {
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

      if (warp == kProdWarp) {
        // ------------------------------------------------------------ stream producer
        if (lane == 0) {
          int g = 0;
          for (int tile = 0; tile < n_tiles; tile++)
            for (int st = 0; st < n_bstages; st++, g++) {
              const int bs = g % kBStages;
              if (g >= kBStages) mbar_wait(&b_empty[bs], (uint32_t)(((g / kBStages) - 1) & 1));
              mbar_expect_tx(&b_full[bs], (uint32_t)kBStageBytes);
              bulk_g2s(sB + bs * kBStageBytes, g_stream + (size_t)st * kBStageBytes, (uint32_t)kBStageBytes,
                       &b_full[bs]);
            }
        }
      } else if (warp == kMmaWarp) {
        // ------------------------------------------------------------ MMA issuer
        const uint32_t tm = __shfl_sync(0xffffffffu, tmem_base, 0);
        uint32_t leader;
        asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(leader));
        // instruction descriptors: D f32, A/B f16, both K-major, M = 128, N = 48 / 32
        const uint32_t idesc_base = (1u << 4) | ((uint32_t)(128 >> 4) << 24);
        const uint32_t idesc48 = idesc_base | ((uint32_t)(48 >> 3) << 17);
        const uint32_t idesc32 = idesc_base | ((uint32_t)(32 >> 3) << 17);
        const uint32_t sA_addr = smem_u32(sA), sB_addr = smem_u32(sB);
        int g = 0;  // cell counter, continuous across tiles
        for (int tile = 0; tile < n_tiles; tile++) {
          const int n_mt = min(kMT, (p_end - (p_begin + tile * kTileRows) + 127) / 128);
          for (int k = 0; k < n_cells; k++, g++) {
            const int s = g % kD, bsg = g / kBCells, bs = bsg % kBStages, sub = g % kBCells;
            if (sub == 0) mbar_wait(&b_full[bs], (uint32_t)((bsg / kBStages) & 1));
            mbar_wait(&a_full[s], (uint32_t)((g / kD) & 1));
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            if (leader) {
              const uint64_t bd = kmajor_desc(sB_addr + bs * kBStageBytes + sub * kCellBytes);
#pragma unroll
              for (int mt = 0; mt < kMT; mt++) {
                if (mt >= n_mt) break;
                const uint32_t acc = tm + mt * kAccCols;
                const uint32_t a_hi = sA_addr + s * kAStageBytes + mt * 2 * kAPlaneBytes;
                umma_f16_ss(acc, kmajor_desc(a_hi), bd, idesc48, k != 0);
                umma_f16_ss(acc + 24, kmajor_desc(a_hi + kAPlaneBytes), bd, idesc32, 1);
              }
              umma_commit(&a_free[s]);
              if (sub == kBCells - 1) umma_commit(&b_empty[bs]);
              if (k == n_cells - 1) umma_commit(&acc_full);
            }
            __syncwarp();
          }
        }
      } else if (warp >= kEvalWarps) {
        // ------------------------------------------------------------ fillers
        const int f = warp - kEvalWarps;
        const uint32_t lt_mask = (1u << lane) - 1u;
        for (int tile = 0; tile < n_tiles; tile++) {
          mbar_wait(&masks_ready, (uint32_t)(tile & 1));
          for (int k = f; k < n_cells; k += kFillWarps) {
            const int g = tile * n_cells + k, s = g % kD, u = g / kD;
            if (u >= 1) mbar_wait(&a_free[s], (uint32_t)((u - 1) & 1));
            unsigned char *stage = sA + s * kAStageBytes;
            unsigned short *list = sList + s * kTileRows;
            uint32_t base = 0;
#pragma unroll 4
            for (int w = 0; w < kEvalWarps; w++) {
              const uint32_t mm = sMixed[k * kEvalWarps + w], mo = sOnes[k * kEvalWarps + w];
              const uint32_t row = (uint32_t)(w * 32 + lane);
              if ((mm >> lane) & 1u) {
                list[base + __popc(mm & lt_mask)] = (unsigned short)row;
              } else {
                const uint32_t v = ((mo >> lane) & 1u) ? 0x3C003C00u : 0u;  // fp16 1.0 twice, or zeros
                unsigned char *hi = stage + (row >> 7) * (2 * kAPlaneBytes) + ((row & 127u) >> 3) * 256 +
                                    (row & 7u) * 16;
                *reinterpret_cast<uint4 *>(hi) = make_uint4(v, v, v, v);
                *reinterpret_cast<uint4 *>(hi + 128) = make_uint4(v, v, v, v);
                *reinterpret_cast<uint4 *>(hi + kAPlaneBytes) = make_uint4(0u, 0u, 0u, 0u);
                *reinterpret_cast<uint4 *>(hi + kAPlaneBytes + 128) = make_uint4(0u, 0u, 0u, 0u);
              }
              base += (uint32_t)__popc(mm);
            }
            if (lane == 0) s_nmixed[s] = base;
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) mbar_arrive(&filled[s]);
          }
        }
      } else {
        // ------------------------------------------------------------ evaluators (phase 1, 2, epilogue)
        const int mt = warp >> 2, q = warp & 3;
        const uint32_t lane_base = tmem_base + ((uint32_t)(q * 32) << 16);
        const uint32_t row_own = (uint32_t)(warp * 32 + lane);
        const uint32_t sConst_addr = smem_u32(sConst);
        for (int tile = 0; tile < n_tiles; tile++) {
          const int idx = p_begin + tile * kTileRows + (int)row_own;
          const bool valid = idx < p_end;
          int slot;
          // ---- phase 1: pass constants to shared memory, cell classes from the centres
          {
            PassRegs c;
            {
              ExtentPass e;
              const float4 *src = reinterpret_cast<const float4 *>(passes + (valid ? idx : p_end - 1));
              float4 *dst = reinterpret_cast<float4 *>(&e);
#pragma unroll
              for (int k = 0; k < (int)(sizeof(ExtentPass) / 16); k++) dst[k] = __ldg(src + k);
#pragma unroll
              for (int k = 0; k < 9; k++) c.b[k] = e.basis[k];
              c.ct0 = e.circle_test[0];
              c.ct1 = e.circle_test[1];
              c.cc0 = e.circle_centre[0];
              c.cc1 = e.circle_centre[1];
              c.m = e.m;
              c.c = e.c;
              c.sin_thr = 0.0f;
              if (e.is_circular) {
                c.sin_thr = __int_as_float(0x7f800000);
                c.cc0 = 0.0f;
                c.cc1 = 1.0f;
              }
              c.a_sin = fmaf(c.m, 1.57079632679490f, c.c);
              slot = e.slot;
            }
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(sConst_addr + const_piece(row_own, 0)),
                         "f"(c.b[0]), "f"(c.b[1]), "f"(c.b[2]), "f"(c.b[3])
                         : "memory");
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(sConst_addr + const_piece(row_own, 1)),
                         "f"(c.b[4]), "f"(c.b[5]), "f"(c.b[6]), "f"(c.b[7])
                         : "memory");
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(sConst_addr + const_piece(row_own, 2)),
                         "f"(c.b[8]), "f"(c.ct0), "f"(c.ct1), "f"(c.cc0)
                         : "memory");
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(sConst_addr + const_piece(row_own, 3)),
                         "f"(c.cc1), "f"(c.sin_thr), "f"(c.m), "f"(c.c)
                         : "memory");
            for (int cp = 0; cp < n_cells / 2; cp++) {
              // centres of cells 2cp, 2cp + 1: (x0, x1, y0, y1), (z0, z1, -, -), (one0, zero0, one1, zero1)
              const float4 xy = *reinterpret_cast<const float4 *>(sCtr + cp * 12);
              const float4 zz = *reinterpret_cast<const float4 *>(sCtr + cp * 12 + 4);
              const float4 th = *reinterpret_cast<const float4 *>(sCtr + cp * 12 + 8);
              float t0, t1;
              weight_fn2(c, pk(xy.x, xy.y), pk(xy.z, xy.w), pk(zz.x, zz.y), t0, t1);
              const unsigned one0 = __ballot_sync(0xffffffffu, valid && t0 >= th.x);
              const unsigned zero0 = __ballot_sync(0xffffffffu, !valid || t0 <= th.y);
              const unsigned one1 = __ballot_sync(0xffffffffu, valid && t1 >= th.z);
              const unsigned zero1 = __ballot_sync(0xffffffffu, !valid || t1 <= th.w);
              if (lane == 0) {
                sMixed[(2 * cp) * kEvalWarps + warp] = ~(one0 | zero0);
                sOnes[(2 * cp) * kEvalWarps + warp] = one0;
                sMixed[(2 * cp + 1) * kEvalWarps + warp] = ~(one1 | zero1);
                sOnes[(2 * cp + 1) * kEvalWarps + warp] = one1;
              }
            }
          }
          __syncwarp();
          if (lane == 0) mbar_arrive(&masks_ready);
          // every evaluator's constants are in shared memory before any batch reads them
          asm volatile("bar.sync 1, %0;" ::"n"(kEvalWarps * 32) : "memory");

          // ---- phase 2: batches of mixed (pass, cell) items
          for (int k = 0; k < n_cells; k++) {
            const int g = tile * n_cells + k, s = g % kD, bsg = g / kBCells, bs = bsg % kBStages;
            if ((g % kBCells) == 0) mbar_wait(&b_full[bs], (uint32_t)((bsg / kBStages) & 1));
            mbar_wait(&filled[s], (uint32_t)((g / kD) & 1));
            const uint32_t n_mixed = s_nmixed[s];
            const uint32_t first = (uint32_t)((warp - 5 * k) & (kEvalWarps - 1)) * 32u;
            if (first < n_mixed) {
              const uint32_t item = first + (uint32_t)lane;
              const bool active = item < n_mixed;
              const uint32_t row = sList[s * kTileRows + (active ? item : first)];
              PassRegs c;
              {
                float4 p0, p1, p2, p3;
                asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                             : "=f"(p0.x), "=f"(p0.y), "=f"(p0.z), "=f"(p0.w)
                             : "r"(sConst_addr + const_piece(row, 0)));
                asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                             : "=f"(p1.x), "=f"(p1.y), "=f"(p1.z), "=f"(p1.w)
                             : "r"(sConst_addr + const_piece(row, 1)));
                asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                             : "=f"(p2.x), "=f"(p2.y), "=f"(p2.z), "=f"(p2.w)
                             : "r"(sConst_addr + const_piece(row, 2)));
                asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                             : "=f"(p3.x), "=f"(p3.y), "=f"(p3.z), "=f"(p3.w)
                             : "r"(sConst_addr + const_piece(row, 3)));
                c.b[0] = p0.x, c.b[1] = p0.y, c.b[2] = p0.z, c.b[3] = p0.w;
                c.b[4] = p1.x, c.b[5] = p1.y, c.b[6] = p1.z, c.b[7] = p1.w;
                c.b[8] = p2.x, c.ct0 = p2.y, c.ct1 = p2.z, c.cc0 = p2.w;
                c.cc1 = p3.x, c.sin_thr = p3.y, c.m = p3.z, c.c = p3.w;
                c.a_sin = fmaf(c.m, 1.57079632679490f, c.c);
              }
              const unsigned char *pts = sB + bs * kBStageBytes + (g % kBCells) * kCellBytes + kCellBBytes;
              uint32_t hi[kCell / 2], lo[kCell / 2];
#pragma unroll
              for (int i = 0; i < kCell / 2; i++) {
                const float4 xy = reinterpret_cast<const float4 *>(pts)[i];  // warp-uniform broadcasts
                const float2 zz = reinterpret_cast<const float2 *>(pts + 128)[i];
                float t0, t1;
                weight_fn2(c, pk(xy.x, xy.y), pk(xy.z, xy.w), pk(zz.x, zz.y), t0, t1);
                const float w0 = __saturatef(t0), w1 = __saturatef(t1);
                const __half2 h = __floats2half2_rn(w0, w1);
                const float2 hf = __half22float2(h);
                const __half2 l = __floats2half2_rn(__fsub_rn(w0, hf.x), __fsub_rn(w1, hf.y));
                hi[i] = *reinterpret_cast<const uint32_t *>(&h);
                lo[i] = *reinterpret_cast<const uint32_t *>(&l);
              }
              if (active) {
                unsigned char *dst = sA + s * kAStageBytes + (row >> 7) * (2 * kAPlaneBytes) +
                                     ((row & 127u) >> 3) * 256 + (row & 7u) * 16;
                *reinterpret_cast<uint4 *>(dst) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                *reinterpret_cast<uint4 *>(dst + 128) = make_uint4(hi[4], hi[5], hi[6], hi[7]);
                *reinterpret_cast<uint4 *>(dst + kAPlaneBytes) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
                *reinterpret_cast<uint4 *>(dst + kAPlaneBytes + 128) = make_uint4(lo[4], lo[5], lo[6], lo[7]);
              }
              fence_proxy_async();
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&a_full[s]);
          }

          // ---- epilogue: this thread's row of [main | corr]; results /= ||results|| in float
          // (ref polar_extent.cpp:271-273)
          mbar_wait(&acc_full, (uint32_t)(tile & 1));
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          if (p_begin + tile * kTileRows + mt * 128 < p_end) {  // M-tiles without passes were not computed
            float r[48];
            {
              float t[16];
#pragma unroll
              for (int k = 0; k < 3; k++) {
                tmem_ld16(lane_base + mt * kAccCols + k * 16, t);
#pragma unroll
                for (int i = 0; i < 16; i++) r[k * 16 + i] = t[i];
              }
            }
            float v[EARB_MAX_SPK];
            float sq = 0.0f;
#pragma unroll
            for (int sp = 0; sp < EARB_MAX_SPK; sp++) {
              v[sp] = (sp < n_spk) ? __fadd_rn(r[sp], r[24 + sp]) : 0.0f;
              sq = __fadd_rn(sq, __fmul_rn(v[sp], v[sp]));
            }
            const float scale = (float)(1.0 / (double)sqrtf(sq));
            if (valid) {
              float4 *dst = reinterpret_cast<float4 *>(results + (size_t)slot * EARB_MAX_SPK);
#pragma unroll
              for (int k = 0; k < EARB_MAX_SPK / 4; k++)
                dst[k] = make_float4(__fmul_rn(v[4 * k], scale), __fmul_rn(v[4 * k + 1], scale),
                                     __fmul_rn(v[4 * k + 2], scale), __fmul_rn(v[4 * k + 3], scale));
            }
          }
          asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        }
      }
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncthreads();
      if (warp == kMmaWarp) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols)
                     : "memory");
      }
    }

    // ---------------------------------------------------------------- host: cell clustering
    struct V3 {
      double x, y, z;
    };
    // recursive bisection along the principal axis of the subset; left halves are whole cells
    void bisect(const std::vector<V3> &pts, std::vector<int> idx, std::vector<std::vector<int>> &cells) {
      const int n = (int)idx.size();
      if (n <= kCell) {
        cells.push_back(idx);
        return;
      }
      double mean[3] = {0, 0, 0};
      for (int i : idx) mean[0] += pts[i].x, mean[1] += pts[i].y, mean[2] += pts[i].z;
      for (double &m : mean) m /= n;
      double cov[3][3] = {};
      for (int i : idx) {
        const double d[3] = {pts[i].x - mean[0], pts[i].y - mean[1], pts[i].z - mean[2]};
        for (int a = 0; a < 3; a++)
          for (int b = 0; b < 3; b++) cov[a][b] += d[a] * d[b];
      }
      // principal axis by power iteration (deterministic start)
      double ax[3] = {0.5773502691896258, 0.5773502691896258 * 1.1, 0.5773502691896258 * 0.9};
      for (int it = 0; it < 200; it++) {
        double t[3];
        for (int a = 0; a < 3; a++) t[a] = cov[a][0] * ax[0] + cov[a][1] * ax[1] + cov[a][2] * ax[2];
        const double nrm = std::sqrt(t[0] * t[0] + t[1] * t[1] + t[2] * t[2]);
        if (nrm < 1e-300) break;
        for (int a = 0; a < 3; a++) ax[a] = t[a] / nrm;
      }
      std::vector<std::pair<double, int>> proj;
      proj.reserve(n);
      for (int i : idx)
        proj.push_back({(pts[i].x - mean[0]) * ax[0] + (pts[i].y - mean[1]) * ax[1] + (pts[i].z - mean[2]) * ax[2], i});
      std::stable_sort(proj.begin(), proj.end());
      const int n_cells = (n + kCell - 1) / kCell, left = (n_cells / 2) * kCell;
      std::vector<int> a, b;
      for (int i = 0; i < n; i++) (i < left ? a : b).push_back(proj[i].second);
      bisect(pts, a, cells);
      bisect(pts, b, cells);
    }
  }  // namespace

  void extent_cells_build(int n_points, const float *xyz, int n_points_pad, ExtentCells &out) {
    std::vector<V3> pts(n_points);
    for (int p = 0; p < n_points; p++) pts[p] = {xyz[p], xyz[n_points_pad + p], xyz[2 * n_points_pad + p]};
    std::vector<int> all(n_points);
    for (int p = 0; p < n_points; p++) all[p] = p;
    std::vector<std::vector<int>> cells;
    bisect(pts, all, cells);
    const int n_cells = n_points_pad / kCell;
    while ((int)cells.size() < n_cells) cells.push_back({});
    out.n_cells = n_cells;
    out.order.assign((size_t)n_cells * kCell, -1);
    out.centre.assign((size_t)n_cells * 3, 0.0f);
    out.radius.assign(n_cells, 0.0);
    for (int c = 0; c < n_cells; c++) {
      const std::vector<int> &cell = cells[c];
      for (size_t i = 0; i < cell.size(); i++) out.order[(size_t)c * kCell + i] = cell[i];
      if (cell.empty()) {
        out.centre[3 * c + 1] = 1.0f;  // an empty padding cell: any direction, zero gains
        continue;
      }
      // bounding cap: start at the normalised mean, walk towards the farthest point
      double ctr[3] = {0, 0, 0};
      for (int i : cell) ctr[0] += pts[i].x, ctr[1] += pts[i].y, ctr[2] += pts[i].z;
      auto normalise = [](double *v) {
        const double n = std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
        if (n > 0) v[0] /= n, v[1] /= n, v[2] /= n;
      };
      normalise(ctr);
      for (int it = 0; it < 400; it++) {
        int far = cell[0];
        double dmin = 2.0;
        for (int i : cell) {
          const double d = pts[i].x * ctr[0] + pts[i].y * ctr[1] + pts[i].z * ctr[2];
          if (d < dmin) dmin = d, far = i;
        }
        const double step = 0.02;
        ctr[0] += step * (pts[far].x - ctr[0]);
        ctr[1] += step * (pts[far].y - ctr[1]);
        ctr[2] += step * (pts[far].z - ctr[2]);
        normalise(ctr);
      }
      // the radius is taken from the centre as the kernel sees it (float)
      const float cf[3] = {(float)ctr[0], (float)ctr[1], (float)ctr[2]};
      double c64[3] = {cf[0], cf[1], cf[2]};
      normalise(c64);
      double r = 0.0;
      for (int i : cell) {
        double p[3] = {pts[i].x, pts[i].y, pts[i].z};
        normalise(p);
        const double cr[3] = {p[1] * c64[2] - p[2] * c64[1], p[2] * c64[0] - p[0] * c64[2], p[0] * c64[1] - p[1] * c64[0]};
        const double s = std::sqrt(cr[0] * cr[0] + cr[1] * cr[1] + cr[2] * cr[2]);
        const double d = p[0] * c64[0] + p[1] * c64[1] + p[2] * c64[2];
        r = std::max(r, std::atan2(s, d));
      }
      out.centre[3 * c] = cf[0];
      out.centre[3 * c + 1] = cf[1];
      out.centre[3 * c + 2] = cf[2];
      out.radius[c] = r;
    }
  }

  size_t extent_cell_stream_bytes(const ExtentCells &cells) { return (size_t)cells.n_cells * kCellBytes; }
  size_t extent_cell_centre_floats(const ExtentCells &cells) { return (size_t)(cells.n_cells / 2) * 12; }

  // per cell: the UMMA B image of [G_hi (24 rows) | G_lo (24 rows)] x 16 points in fp16 --
  // [6 row groups][2 K-chunks][8 rows][8 halves] -- followed by the point coordinates as eight
  // (x, x', y, y') and eight (z, z') pairs.  gains is speaker-major [S][Pp] float in grid order.
  void extent_cell_build_stream(const ExtentCells &cells, int n_spk, int n_points_pad, const float *xyz,
                                const float *gains, unsigned char *stream) {
    memset(stream, 0, extent_cell_stream_bytes(cells));
    for (int c = 0; c < cells.n_cells; c++) {
      unsigned char *cell = stream + (size_t)c * kCellBytes;
      __half *b = reinterpret_cast<__half *>(cell);
      float *pxy = reinterpret_cast<float *>(cell + kCellBBytes);
      float *pz = reinterpret_cast<float *>(cell + kCellBBytes + 128);
      // padding slots repeat a real point of the cell (or the cell's centre), with zero gains
      int fallback = -1;
      for (int i = 0; i < kCell; i++)
        if (cells.order[(size_t)c * kCell + i] >= 0) fallback = cells.order[(size_t)c * kCell + i];
      for (int i = 0; i < kCell; i++) {
        const int p = cells.order[(size_t)c * kCell + i];
        const int q = p >= 0 ? p : fallback;
        const float x = q >= 0 ? xyz[q] : cells.centre[3 * c];
        const float y = q >= 0 ? xyz[n_points_pad + q] : cells.centre[3 * c + 1];
        const float z = q >= 0 ? xyz[2 * n_points_pad + q] : cells.centre[3 * c + 2];
        pxy[(i / 2) * 4 + (i & 1)] = x;
        pxy[(i / 2) * 4 + 2 + (i & 1)] = y;
        pz[(i / 2) * 2 + (i & 1)] = z;
        if (p < 0) continue;
        for (int s = 0; s < n_spk; s++) {
          const float g = gains[(size_t)s * n_points_pad + p];
          const __half hi = __float2half_rn(g);
          const __half lo = __float2half_rn(g - __half2float(hi));
          auto at = [&](int row) { return (row / 8) * 128 + (i / 8) * 64 + (row % 8) * 8 + (i % 8); };
          b[at(s)] = hi;
          b[at(24 + s)] = lo;
        }
      }
    }
  }

  // centres of cells 2j, 2j + 1 as (x0, x1, y0, y1), (z0, z1, 0, 0) and the class thresholds
  // (one0, zero0, one1, zero1): all weights are 1 when the weight function at the centre is
  // >= one, all 0 when it is <= zero (|m| = 1 / radians(10), ref polar_extent.cpp:194)
  void extent_cell_build_centres(const ExtentCells &cells, float *out) {
    const double M = 1.0 / (10.0 * 3.141592653589793238462643383279502884 / 180.0);
    for (int j = 0; j < cells.n_cells / 2; j++) {
      float *o = out + (size_t)j * 12;
      for (int h = 0; h < 2; h++) {
        const int c = 2 * j + h;
        o[h] = cells.centre[3 * c];
        o[2 + h] = cells.centre[3 * c + 1];
        o[4 + h] = cells.centre[3 * c + 2];
        o[8 + 2 * h] = (float)(1.0 + M * cells.radius[c] + kClassMargin);
        o[9 + 2 * h] = (float)(-(M * cells.radius[c] + kClassMargin));
      }
      o[6] = o[7] = 0.0f;
    }
  }

  size_t extent_cell_smem_bytes(int n_cells) {
    return (size_t)kD * kAStageBytes + (size_t)kBStages * kBStageBytes + (size_t)kTileRows * kConstBytes +
           (size_t)n_cells * kEvalWarps * 2 * sizeof(uint32_t) + (size_t)(n_cells / 2) * 12 * sizeof(float) +
           (size_t)kD * kTileRows * sizeof(unsigned short);
  }

  void launch_extent_cell(int n_cells, int n_spk, const unsigned char *d_stream, const float *d_centres,
                          const unsigned int *d_n_active, const ExtentPass *d_passes, float *d_results,
                          int n_sms, cudaStream_t stream) {
    const size_t smem = extent_cell_smem_bytes(n_cells);
    // the attribute is per device: set it on every launch (a few hundred nanoseconds)
    cudaFuncSetAttribute(extent_cell_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    extent_cell_kernel<<<(unsigned)n_sms, kThreads, smem, stream>>>(n_cells, n_spk, d_stream, d_centres,
                                                                      d_n_active, d_passes, d_results);
  }
}  // namespace earb
