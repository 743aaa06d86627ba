// K2 core on the 5th-generation tensor cores (tcgen05 + TMEM), sm_100a only.
//
// Replaces the hot loop of the extent panner, PolarExtentCoreSimd::run
// (ref src/object_based/polar_extent_simd.hpp:97-135): for every spreading pass
//     results[s] = sum_p w_p * G[s][p],   p over the 1,652 virtual sources,
// i.e. R[passes x S] = W[passes x P] * G^T[P x S] with W generated on the fly.
//
// One persistent CTA per SM works on tiles of 512 passes (four M = 128 UMMA tiles):
//   warps 0..15  generators   thread = pass = TMEM lane.  Each thread keeps its pass context
//                             (the 21 floats of PolarExtentCoreContext) in registers and
//                             evaluates the weight of 16 virtual sources at a time -- the point
//                             coordinates are warp-uniform shared-memory broadcasts -- splits
//                             each weight into TF32 hi + lo and writes them to TMEM as the A
//                             operand (tcgen05.st).  After the last chunk the same threads read
//                             their row of the accumulator back, normalise it and store it.
//   warp  16     MMA issuer   tcgen05.mma kind::tf32, M = 128:
//                               [main | corr] (+)= w_hi * [G_hi | G_lo]   (N = 48)
//                               corr          +=  w_lo * G_hi             (N = 32)
//   warp  17     B producer   streams the gain table (G_hi rows 0..23, G_lo rows 24..47, K-major
//                             canonical layout, prepared once per layout) through a
//                             shared-memory ring with cp.async.bulk, 64 points per stage.
//
// The weight is NOT computed with the reference's threshold comparisons
// (polar_extent_simd.hpp:39-95: 1 for angle <= start, 0 for angle >= end, the fade m * angle + c in
// between) but as clamp(m * angle + c, 0, 1) with one branch-free polynomial asin -- the fade is 1
// at `start`, 0 at `end` and monotone, so the two agree everywhere except for rounding at the two
// ends of the fade (measured 1.5e-7 on the gains, bar 1e-5).  3xTF32 keeps the contraction at fp32
// accuracy (weights off the fade are exactly 0 or 1, exact in TF32).
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <cstring>

#include "kernels.hpp"

namespace earb {
  namespace {
    constexpr int kKC = EXTENT_TC_STAGE_POINTS;  // points per B stage
    constexpr int kChunk = 16;                   // points per A hand-off
    constexpr int kMTiles = 4;
    constexpr int kGenWarps = 4 * kMTiles;
    constexpr int kMmaWarp = kGenWarps, kProdWarp = kGenWarps + 1;
    constexpr int kThreads = (kProdWarp + 1) * 32;  // 576
    constexpr int kBStages = 4;
    constexpr int kRows = 48;  // B rows: 24 hi + 24 lo
    constexpr int kStageFloats = kRows * kKC;
    constexpr int kTmemCols = 512;
    constexpr int kAccCols = 64;            // per M-tile: [main 24 | corr 24 | spill-over 8 | -]
    constexpr int kColA = kMTiles * kAccCols;  // A: [M-tile][2 buffers][hi 16 | lo 16]
    static_assert(kColA + kMTiles * 2 * 2 * kChunk <= kTmemCols, "TMEM budget");
    static_assert(EARB_MAX_SPK <= 24, "B row packing assumes at most 24 loudspeakers");

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
    // spins on the barrier phase; a protocol error traps after ~2 s instead of hanging the GPU
    __device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
      uint32_t done;
      const long long t0 = clock64();
      do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
        if (!done && clock64() - t0 > 4000000000LL) __trap();
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
    // B stage in shared memory: [6 row groups][kKC/4 K-chunks][8 rows][4 floats] -- K-major,
    // no swizzle; core matrices adjacent in K are 128 bytes apart, row groups kKC/4*128 bytes
    __device__ __forceinline__ uint64_t b_desc(uint32_t smem_addr) {
      const uint64_t lbo = 128 >> 4;
      const uint64_t sbo = (uint64_t)(kKC / 4 * 128) >> 4;
      return (uint64_t)((smem_addr >> 4) & 0x3FFF) | (lbo << 16) | (sbo << 32) | (1ull << 46);
    }

    // pass context in registers (PolarExtentCoreContext, ref polar_extent_core.hpp:12-42).
    // Circular extents (ref weight_circle, polar_extent_simd.hpp:61-66) run through the stadium
    // code with constants that reduce it to the circle test exactly: sin_thr = +inf disables the
    // straight-line part and (cc0, cc1) = (0, 1) makes circle_dot = fma(|x'|, 0, y' * 1) = y'.
    struct PassRegs {
      float b[9];
      float ct0, ct1, cc0, cc1;
      float sin_thr;  // straight-line part when circle_test_dot >= sin_thr (0, or +inf if circular)
      float m, c;     // fade: weight = m * angle + c between the start and end angles
      float a_sin;    // m * pi/2 + c
    };

    // Weight of one virtual source (ref polar_extent_simd.hpp:39-95), branch-free.
    // The reference returns 1 for angle <= start, 0 for angle >= end and the fade
    // m * angle + c in between; the fade is 1 at `start`, 0 at `end` and monotone, and the
    // sentinels libear uses when an angle passes pi (pi/2 for the sine test) only ever disable a
    // comparison whose fade value is already inside [0, 1] (polar_extent.cpp:176-199), so the whole
    // function is clamp(m * angle + c, 0, 1) -- one saturating FMA instead of two comparisons and
    // two selects (weights differ from the reference's by rounding at the fade ends only).
    // angle = acos(v) on the circle part, asin(|z'|) on the straight part, both from
    // r = acos(|v|) = sqrt(u) * Q7(u), u = 1 - |v|: Abramowitz & Stegun 4.4.46 (|error| <= 2.2e-8)
    // re-expanded in u, so that |v| is only ever an operand modifier of the subtraction -- a
    // separate |.| is an ALU instruction, and those hold the dispatch port two cycles
    // (tools/ubench_fp32x2.cu):
    // With q = pi/2 - r = asin(|v|) (folded into the last FMA) the three cases
    //   straight:        angle = q                         weight = sat(m q + c)
    //   circle:          angle = pi/2 - copysign(q, v)     weight = sat(m (-copysign(q, v)) + (m pi/2 + c))
    // share one saturating FMA whose operand is q with the inverted sign of v (one LOP3).

    // ---- packed fp32 (sm_100 FFMA2 / FMUL2 / FADD2: two FMAs per lane per issue slot).  The kernel
    // is issue-bound, so the weights of two virtual sources are evaluated side by side: the FMA-pipe
    // work keeps its cost, the issue slots it needs halve.  Scalar operands (the pass constants)
    // are given as broadcast pairs, which ptxas folds into the instruction's scalar-operand form.
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
    __device__ __forceinline__ f32x2 sub2(f32x2 a, f32x2 b) {
      f32x2 r;
      asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
      return r;
    }
    __device__ __forceinline__ f32x2 bc(float a) { return pk(a, a); }

    // the weight of two virtual sources at once: X, Y, Z hold the coordinates of both (the same
    // operations in the same order per half)
    __device__ __forceinline__ void point_weight2(const PassRegs &c, f32x2 X, f32x2 Y, f32x2 Z, float &w0,
                                                  float &w1) {
      const f32x2 xt = fma2(X, bc(c.b[0]), fma2(Y, bc(c.b[1]), mul2(Z, bc(c.b[2]))));
      const f32x2 yt = fma2(X, bc(c.b[3]), fma2(Y, bc(c.b[4]), mul2(Z, bc(c.b[5]))));
      const f32x2 zt = fma2(X, bc(c.b[6]), fma2(Y, bc(c.b[7]), mul2(Z, bc(c.b[8]))));
      // |x'| rides as an operand modifier of scalar FMAs (same operations as point_weight)
      float x0, x1, ty0, ty1, tc0, tc1, z0, z1;
      upk(xt, x0, x1);
      upk(mul2(yt, bc(c.ct1)), ty0, ty1);
      upk(mul2(yt, bc(c.cc1)), tc0, tc1);
      upk(zt, z0, z1);
      const float ctd0 = fmaf(fabsf(x0), c.ct0, ty0), ctd1 = fmaf(fabsf(x1), c.ct0, ty1);
      const float cd0 = fmaf(fabsf(x0), c.cc0, tc0), cd1 = fmaf(fabsf(x1), c.cc0, tc1);
      const bool s0 = ctd0 >= c.sin_thr, s1 = ctd1 >= c.sin_thr;
      // the straight part enters as -|z'|: with the inverted sign of v on q below, both parts are
      // weight = sat(m * (-copysign(q, v)) + k)
      const float v0 = s0 ? -fabsf(z0) : cd0, v1 = s1 ? -fabsf(z1) : cd1;
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
      // nq = sq * p - pi/2 = -asin(|v|)
      const f32x2 nq = fma2(pk(q0, q1), p, bc(-1.57079632679490f));
      float n0, n1;
      upk(nq, n0, n1);
      // asin(|v|) carrying the inverted sign of v: n = -asin(|v|) has its sign bit set, so one
      // XOR with the sign of v does it (n can round to +1e-7 at |v| < 2^-25, where the sign of an
      // angle of 1e-7 does not matter)
      const float t0 = __uint_as_float(__float_as_uint(n0) ^ (__float_as_uint(v0) & 0x80000000u));
      const float t1 = __uint_as_float(__float_as_uint(n1) ^ (__float_as_uint(v1) & 0x80000000u));
      w0 = __saturatef(fmaf(c.m, t0, s0 ? c.c : c.a_sin));
      w1 = __saturatef(fmaf(c.m, t1, s1 ? c.c : c.a_sin));
    }

    // Barriers
    //   a_full[mt][b]  generators -> MMA         A buffer b of M-tile mt is complete (4 warps)
    //   a_free[mt][b]  MMA commit -> generators  the UMMAs reading it have finished
    //   b_full[s]      producer   -> MMA         B stage s has landed (tx bytes)
    //   b_empty[s]     MMA commit -> producer    the UMMAs reading B stage s have finished
    //   acc_full       MMA commit -> generators  every UMMA of the tile has finished
    __global__ void __launch_bounds__(kThreads, 1)
        extent_core_tc_kernel(ExtentGrid grid, const float4 *__restrict__ g_points,
                              const float *__restrict__ g_bimage,
                              const unsigned int *__restrict__ n_active_ptr,
                              const ExtentPass *__restrict__ passes, float *__restrict__ results) {
      extern __shared__ __align__(128) unsigned char smem_raw[];
      const int Pp = grid.n_points_pad;  // multiple of kKC
      const int S = grid.n_spk;
      float *sB = reinterpret_cast<float *>(smem_raw);                 // [kBStages][kStageFloats]
      // virtual sources in pairs: sxy[j] = (x, x', y, y'), sz[j] = (z, z') of points 2j and 2j + 1
      float4 *sxy = reinterpret_cast<float4 *>(sB + kBStages * kStageFloats);  // [Pp / 2]
      float2 *sz = reinterpret_cast<float2 *>(sxy + Pp / 2);                     // [Pp / 2]
      __shared__ __align__(8) uint64_t a_full[kMTiles][2], a_free[kMTiles][2];
      __shared__ __align__(8) uint64_t b_full[kBStages], b_empty[kBStages], acc_full;
      __shared__ uint32_t tmem_base_s;

      const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
      const int n_active = (int)*n_active_ptr;
      // every CTA owns one contiguous range of passes, equal to within one warp (32 passes): the
      // CTAs finish together whatever n_active is (whole 512-pass tiles dealt out round-robin left
      // 2.6 tiles per CTA at the 131,072-object chunk of the batched path, i.e. 3 for most).  Only
      // the last tile of a CTA is partial: its empty M-tiles are skipped, its empty warps only keep
      // the barriers moving.
      const int per_cta = (((n_active + (int)gridDim.x - 1) / (int)gridDim.x) + 31) & ~31;
      const int p_begin = min(n_active, (int)blockIdx.x * per_cta), p_end = min(n_active, p_begin + per_cta);
      const int n_tiles = (p_end - p_begin + kMTiles * 128 - 1) / (kMTiles * 128);
      const int n_chunks = Pp / kChunk, n_stages = Pp / kKC;
      if (n_tiles == 0) return;  // no passes for this CTA (small batches): nothing to set up

      for (int j = tid; j < Pp / 2; j += kThreads) {
        const float4 a = g_points[2 * j], b = g_points[2 * j + 1];
        sxy[j] = make_float4(a.x, b.x, a.y, b.y);
        sz[j] = make_float2(a.z, b.z);
      }
      if (tid == 0) {
        for (int m = 0; m < kMTiles; m++)
          for (int b = 0; b < 2; b++) {
            mbar_init(&a_full[m][b], 4);
            mbar_init(&a_free[m][b], 1);
          }
        for (int s = 0; s < kBStages; s++) {
          mbar_init(&b_full[s], 1);
          mbar_init(&b_empty[s], 1);
        }
        mbar_init(&acc_full, 1);
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
        // ------------------------------------------------------------ B producer
        if (lane == 0) {
          int g = 0;
          for (int tile = 0; tile < n_tiles; tile++)
            for (int s = 0; s < n_stages; s++, g++) {
              const int bs = g % kBStages;
              if (g >= kBStages) mbar_wait(&b_empty[bs], (uint32_t)(((g / kBStages) - 1) & 1));
              mbar_expect_tx(&b_full[bs], (uint32_t)(kStageFloats * sizeof(float)));
              bulk_g2s(sB + bs * kStageFloats, g_bimage + (size_t)s * kStageFloats,
                       (uint32_t)(kStageFloats * sizeof(float)), &b_full[bs]);
            }
        }
      } else if (warp == kMmaWarp) {
        // ------------------------------------------------------------ MMA issuer
        // the whole warp runs the loop (warp-uniform operands); one elected lane issues
        const uint32_t tm = __shfl_sync(0xffffffffu, tmem_base, 0);
        uint32_t leader;
        asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(leader));
        // instruction descriptors: D f32, A/B tf32, both K-major, M = 128, N = 48 / 32
        const uint32_t idesc_base = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(128 >> 4) << 24);
        const uint32_t idesc48 = idesc_base | ((uint32_t)(48 >> 3) << 17);
        const uint32_t idesc32 = idesc_base | ((uint32_t)(32 >> 3) << 17);
        const uint64_t bdesc0 = b_desc(smem_u32(sB));
        int gs = 0, gc = 0;  // B stage / chunk counters, continuous across tiles
        for (int tile = 0; tile < n_tiles; tile++) {
          // M-tiles of this tile that hold passes
          const int n_mt = min(kMTiles, (p_end - (p_begin + tile * (kMTiles * 128)) + 127) / 128);
          for (int c = 0; c < n_chunks; c++, gc++) {
            const int sub = c % (kKC / kChunk);
            const int bs = gs % kBStages;
            if (sub == 0) mbar_wait(&b_full[bs], (uint32_t)((gs / kBStages) & 1));
            const int buf = gc & 1;
            // chunk `sub` of the stage: kChunk points = 4 K-chunks of 128 bytes
            const uint64_t bd = bdesc0 + (uint64_t)((bs * kStageFloats * 4 + sub * (kChunk / 4) * 128) >> 4);
#pragma unroll
            for (int mt = 0; mt < kMTiles; mt++) {
              if (mt >= n_mt) break;
              mbar_wait(&a_full[mt][buf], (uint32_t)((gc >> 1) & 1));
              asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
              if (leader) {
                const uint32_t acc = tm + mt * kAccCols;
                const uint32_t a_hi = tm + kColA + (mt * 2 + buf) * 2 * kChunk;
#pragma unroll
                for (int ks = 0; ks < kChunk / 8; ks++) {
                  umma_tf32_ts(acc, a_hi + ks * 8, bd + (uint64_t)(ks * ((2 * 128) >> 4)), idesc48,
                               (c | ks) != 0);
                  umma_tf32_ts(acc + 24, a_hi + kChunk + ks * 8, bd + (uint64_t)(ks * ((2 * 128) >> 4)),
                               idesc32, 1);
                }
                umma_commit(&a_free[mt][buf]);
              }
              __syncwarp();
            }
            if (sub == kKC / kChunk - 1) {
              if (leader) umma_commit(&b_empty[bs]);
              __syncwarp();
              gs++;
            }
          }
          if (leader) umma_commit(&acc_full);
          __syncwarp();
        }
      } else {
        // ------------------------------------------------------------ generators + epilogue
        const int mt = warp >> 2, q = warp & 3;
        const uint32_t lane_base = tmem_base + ((uint32_t)(q * 32) << 16);
        int gc = 0, L = 0;
        for (int tile = 0; tile < n_tiles; tile++, L++) {
          const int base = p_begin + tile * (kMTiles * 128) + mt * 128;
          if (base >= p_end) break;  // an empty M-tile (last tile of the CTA only)
          const int idx = base + q * 32 + lane;
          const bool valid = idx < p_end;
          if (base + q * 32 >= p_end) {
            // no passes in this warp (last tile of the CTA only): keep a_full / a_free in step
            for (int ch = 0; ch < n_chunks; ch++, gc++) {
              const int buf = gc & 1;
              if (gc >= 2) mbar_wait(&a_free[mt][buf], (uint32_t)(((gc >> 1) - 1) & 1));
              __syncwarp();
              if (lane == 0) mbar_arrive(&a_full[mt][buf]);
            }
            break;
          }
          const ExtentPass *pp = passes + (valid ? idx : n_active - 1);
          PassRegs c;
          int slot;
          {
            ExtentPass e;
            const float4 *src = reinterpret_cast<const float4 *>(pp);
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
            c.a_sin = fmaf(e.m, 1.57079632679490f, e.c);
            c.sin_thr = 0.0f;
            if (e.is_circular) {
              c.sin_thr = __int_as_float(0x7f800000);
              c.cc0 = 0.0f;
              c.cc1 = 1.0f;
            }
            slot = e.slot;
          }
          for (int ch = 0; ch < n_chunks; ch++, gc++) {
            const int buf = gc & 1;
            float hi[kChunk], lo[kChunk];
#pragma unroll
            for (int i = 0; i < kChunk; i += 2) {
              const float4 xy = sxy[(ch * kChunk + i) >> 1];  // warp-uniform broadcasts
              const float2 zz = sz[(ch * kChunk + i) >> 1];
              float w0, w1;
              point_weight2(c, pk(xy.x, xy.y), pk(xy.z, xy.w), pk(zz.x, zz.y), w0, w1);
              hi[i] = __uint_as_float(__float_as_uint(w0) & 0xffffe000u);
              hi[i + 1] = __uint_as_float(__float_as_uint(w1) & 0xffffe000u);
              upk(sub2(pk(w0, w1), pk(hi[i], hi[i + 1])), lo[i], lo[i + 1]);
            }
            if (gc >= 2) {
              mbar_wait(&a_free[mt][buf], (uint32_t)(((gc >> 1) - 1) & 1));
              asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            }
            const uint32_t a_col = kColA + (mt * 2 + buf) * 2 * kChunk;
            tmem_st16(lane_base + a_col, hi);
            tmem_st16(lane_base + a_col + kChunk, lo);
            asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(&a_full[mt][buf]);
          }
          // epilogue: this thread's row of [main | corr]; results /= ||results|| in float
          // (ref polar_extent.cpp:271-273)
          mbar_wait(&acc_full, (uint32_t)(L & 1));
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
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
          asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
          float v[EARB_MAX_SPK];
          float sq = 0.0f;
#pragma unroll
          for (int s = 0; s < EARB_MAX_SPK; s++) {
            v[s] = (s < S) ? __fadd_rn(r[s], r[24 + s]) : 0.0f;
            sq = __fadd_rn(sq, __fmul_rn(v[s], v[s]));
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
      }
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncthreads();
      if (warp == kMmaWarp) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols)
                     : "memory");
      }
    }
  }  // namespace

  size_t extent_tc_smem_bytes(const ExtentGrid &grid) {
    return sizeof(float) * (size_t)kBStages * kStageFloats + sizeof(float4) * (size_t)grid.n_points_pad;
  }

  size_t extent_tc_bimage_floats(const ExtentGrid &grid) { return (size_t)kRows * grid.n_points_pad; }

  // host: K-major canonical image of [G_hi (24 rows) | G_lo (24 rows)] per stage of kKC points:
  // image[stage][row group 0..5][K-chunk][8 rows][4]; gains is speaker-major [S][Pp] float
  void extent_tc_build_bimage(const ExtentGrid &grid, const float *gains, float *image) {
    const int Pp = grid.n_points_pad, S = grid.n_spk;
    auto tf32 = [](float x) {
      uint32_t u;
      memcpy(&u, &x, 4);
      u = (u + 0x1000u) & 0xffffe000u;
      float r;
      memcpy(&r, &u, 4);
      return r;
    };
    std::fill(image, image + (size_t)kRows * Pp, 0.0f);
    for (int s = 0; s < S; s++)
      for (int p = 0; p < Pp; p++) {
        const float g = gains[(size_t)s * Pp + p];
        const float hi = tf32(g), lo = tf32(g - hi);
        const int stage = p / kKC, k = p % kKC;
        auto at = [&](int row) {
          return (size_t)stage * kStageFloats + (size_t)(row / 8) * (kKC / 4 * 32) + (size_t)(k / 4) * 32 +
                 (row % 8) * 4 + (k % 4);
        };
        image[at(s)] = hi;
        image[at(24 + s)] = lo;
      }
  }

  void launch_extent_core_tc(const ExtentGrid &grid, const float4 *d_points, const float *d_bimage,
                             const unsigned int *d_n_active, const ExtentPass *d_passes, float *d_results,
                             int n_sms, cudaStream_t stream) {
    const size_t smem = extent_tc_smem_bytes(grid);
    // the attribute is per device: set it on every launch (a few hundred nanoseconds)
    cudaFuncSetAttribute(extent_core_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    extent_core_tc_kernel<<<(unsigned)n_sms, kThreads, smem, stream>>>(grid, d_points, d_bimage, d_n_active,
                                                                         d_passes, d_results);
  }
}  // namespace earb
