// K1 device code: point-source panning of one position over the flat region tables.
//
// FP64 throughout, compiled with --fmad=false so that every multiply/add rounds
// exactly as libear's (-ffp-contract=off) CPU build does; the region acceptance tests
// (>= -1e-11 for triplets, (-1e-10, 1+1e-10) for quad roots) therefore make the same
// decisions, which is what "bit-exact region index" rests on.
//
// Follows: PolarPointSourcePanner::handle (ref src/common/point_source_panner.cpp:201-211),
// Triplet::handle (:41-50), VirtualNgon::handle (:84-99), QuadRegion::handle (:113-130),
// QuadRegion::_pan (:174-186), real_quadratic_roots (:159-172),
// PointSourcePannerDownmix::handle (:234-244), StereoPannerDownmix::handle (:369-392).
#pragma once
#include "tables.h"

namespace earb {

  struct PspHit {
    int region;  // index into PspTables::regions, -1 if no region accepted the position
    int inner;   // inner triplet index for virtual n-gons, else -1
  };

  __device__ __forceinline__ bool psp_triplet(const PspTriplet &t, double x, double y,
                                              double z, double pv[3]) {
#pragma unroll
    for (int j = 0; j < 3; j++)
      pv[j] = (x * t.basis[j] + y * t.basis[3 + j]) + z * t.basis[6 + j];
    const double epsilon = -1e-11;
    if (pv[0] >= epsilon && pv[1] >= epsilon && pv[2] >= epsilon) {
      double norm = sqrt((pv[0] * pv[0] + pv[1] * pv[1]) + pv[2] * pv[2]);
#pragma unroll
      for (int j = 0; j < 3; j++) pv[j] = fmin(fmax(pv[j] / norm, 0.0), 1.0);
      return true;
    }
    return false;
  }

  // one bilinear coordinate of a position inside a quad: the root in [0, 1] of
  // (polyBasis * pos) . (r^2, r, 1)
  __device__ __forceinline__ bool psp_quad_pan(const double *pb, double x, double y,
                                               double z, double &out) {
    const double eps = 1e-10;
    double a = (pb[0] * x + pb[1] * y) + pb[2] * z;
    double b = (pb[3] * x + pb[4] * y) + pb[5] * z;
    double c = (pb[6] * x + pb[7] * y) + pb[8] * z;
    double roots[2];
    int n_roots = 0;
    if (fabs(c) < eps) {
      roots[n_roots++] = 0.0;
    } else if (fabs(a) < eps) {
      roots[n_roots++] = -c / b;
    } else {
      double det = b * b - 4.0 * a * c;
      if (det > eps) {
        double sq = sqrt(det);
        roots[n_roots++] = (-b + sq) / (2.0 * a);
        roots[n_roots++] = (-b - sq) / (2.0 * a);
      } else if (det > -eps) {
        roots[n_roots++] = -b / (2.0 * a);
      }
    }
    for (int i = 0; i < n_roots; i++) {
      if (-eps < roots[i] && roots[i] < 1.0 + eps) {
        out = fmin(fmax(roots[i], 0.0), 1.0);
        return true;
      }
    }
    return false;
  }

  __device__ __forceinline__ bool psp_quad(const PspQuad &q, double x, double y, double z,
                                           double pvs[4]) {
    double u, v;
    if (!psp_quad_pan(q.polyX, x, y, z, u)) return false;
    if (!psp_quad_pan(q.polyY, x, y, z, v)) return false;
    double g[4] = {(1 - u) * (1 - v), u * (1 - v), u * v, (1 - u) * v};
#pragma unroll
    for (int i = 0; i < 4; i++) pvs[i] = 0.0;
#pragma unroll
    for (int i = 0; i < 4; i++) {
      // pvs[order[i]] = g[i] without dynamic register indexing
      int o = q.order[i];
#pragma unroll
      for (int k = 0; k < 4; k++)
        if (o == k) pvs[k] = g[i];
    }
    // the panned direction must point towards the source
    double vx = 0.0, vy = 0.0, vz = 0.0;
#pragma unroll
    for (int i = 0; i < 4; i++) {
      vx += pvs[i] * q.pos[3 * i];
      vy += pvs[i] * q.pos[3 * i + 1];
      vz += pvs[i] * q.pos[3 * i + 2];
    }
    if ((vx * x + vy * y) + vz * z <= 0) return false;
    double norm = sqrt(((pvs[0] * pvs[0] + pvs[1] * pvs[1]) + pvs[2] * pvs[2]) + pvs[3] * pvs[3]);
#pragma unroll
    for (int i = 0; i < 4; i++) pvs[i] /= norm;
    return true;
  }

  // One candidate of the first-match scan: region r (for a virtual n-gon: its inner triplet k).
  // On acceptance the gains of the region's vertices are written to aug[] (which the caller has
  // zeroed) and true is returned.
  __device__ inline bool psp_try_candidate(const PspTables &T, int r, int k, double x, double y, double z,
                                           double *aug) {
    const int kind = T.regions[r].kind;
    const int idx = T.regions[r].index;
    if (kind == EARB_REGION_TRIPLET) {
      double pv[3];
      const PspTriplet &t = T.tri[idx];
      if (!psp_triplet(t, x, y, z, pv)) return false;
      aug[t.ch[0]] = pv[0];
      aug[t.ch[1]] = pv[1];
      aug[t.ch[2]] = pv[2];
      return true;
    }
    if (kind == EARB_REGION_QUAD) {
      double pvs[4];
      const PspQuad &q = T.quad[idx];
      if (!psp_quad(q, x, y, z, pvs)) return false;
#pragma unroll
      for (int i = 0; i < 4; i++) aug[q.ch[i]] = pvs[i];
      return true;
    }
    const PspNgon &g = T.ngon[idx];
    double pv[3];
    const PspTriplet &t = T.tri[g.first_tri + k];
    if (!psp_triplet(t, x, y, z, pv)) return false;
    // spread the virtual centre speaker's gain over all n real vertices and
    // renormalise; the two edge vertices keep their own gain as well
    const double centre = g.centre_downmix * pv[2];
    double sumsq = 0.0;
    for (int i = 0; i < g.n; i++) {
      double v = 0.0;
      if (i == t.ch[0]) v = pv[0];
      if (i == t.ch[1]) v = pv[1];
      v = v + centre;
      aug[g.ch[i]] = v;
      sumsq += v * v;
    }
    const double norm = sqrt(sumsq);
    for (int i = 0; i < g.n; i++) aug[g.ch[i]] /= norm;
    return true;
  }

  // 0/1 downmix of the extra loudspeakers into their mid-layer parents, power normalisation, and
  // the stereo downmix of 0+2+0: aug[] -> out[T.n_real]
  __device__ inline void psp_finish(const PspTables &T, const double *aug, double *out) {
    const int A = T.n_aug;
    const int S5 = T.stereo ? 5 : T.n_real;
    double full[EARB_MAX_SPK];
    for (int s = 0; s < S5; s++) full[s] = 0.0;
    for (int a = 0; a < A; a++) full[T.parent[a]] += aug[a];
    double sumsq = 0.0;
    for (int s = 0; s < S5; s++) sumsq += full[s] * full[s];
    const double norm = sqrt(sumsq);
    for (int s = 0; s < S5; s++) full[s] /= norm;

    if (!T.stereo) {
      for (int s = 0; s < S5; s++) out[s] = full[s];
    } else {
      // 0+5+0 order: M+030, M-030, M+000, M+110, M-110
      const double k3 = sqrt(3.0) / 3.0, k5 = sqrt(0.5);
      double l = (((1.0 * full[0] + 0.0 * full[1]) + k3 * full[2]) + k5 * full[3]) + 0.0 * full[4];
      double r = (((0.0 * full[0] + 1.0 * full[1]) + k3 * full[2]) + 0.0 * full[3]) + k5 * full[4];
      double n2 = sqrt(l * l + r * r);
      l /= n2;
      r /= n2;
      // 0 dB at the front to -3 dB at the back
      double front = fmax(fmax(full[0], full[1]), full[2]);
      double back = fmax(full[3], full[4]);
      double scale = pow(0.5, 0.5 * back / (front + back));
      out[0] = l * scale;
      out[1] = r * scale;
    }
  }

  // Pan one position. `out` receives T.n_real gains (S for a full layout, 2 for 0+2+0).
  // `aug` is caller-provided scratch of EARB_MAX_AUG doubles.
  __device__ inline PspHit psp_handle(const PspTables &T, double x, double y, double z,
                                      double *out, double *aug) {
    PspHit hit;
    hit.region = -1;
    hit.inner = -1;
    const int A = T.n_aug;
    for (int a = 0; a < A; a++) aug[a] = 0.0;
    for (int r = 0; r < T.n_regions && hit.region < 0; r++) {
      if (T.regions[r].kind == EARB_REGION_NGON) {
        const int n = T.ngon[T.regions[r].index].n;
        for (int k = 0; k < n; k++)
          if (psp_try_candidate(T, r, k, x, y, z, aug)) {
            hit.region = r;
            hit.inner = k;
            break;
          }
      } else if (psp_try_candidate(T, r, -1, x, y, z, aug)) {
        hit.region = r;
      }
    }
    psp_finish(T, aug, out);
    return hit;
  }

  // The same scan by a whole warp for ONE position: lane c tests candidate c of the scan order (the
  // inner triplets of a virtual n-gon count one candidate each), the first accepting candidate --
  // lowest index, as in the serial scan -- wins and its lane finishes.  Every candidate test is the
  // serial code on the same inputs, so region, inner index and gains are bit-identical to
  // psp_handle; what changes is the latency: two rounds of tests instead of up to 44 in a row.
  // All 32 lanes must call it; `out` is written by one lane only.
  __device__ inline PspHit psp_handle_warp(const PspTables &T, double x, double y, double z, double *out,
                                           double *aug) {
    const int lane = (int)(threadIdx.x & 31u);
    PspHit hit;
    hit.region = -1;
    hit.inner = -1;
    const int A = T.n_aug;
    int n_cand = 0;
    for (int r = 0; r < T.n_regions; r++)
      n_cand += T.regions[r].kind == EARB_REGION_NGON ? T.ngon[T.regions[r].index].n : 1;
    for (int c0 = 0; c0 < n_cand; c0 += 32) {
      const int c = c0 + lane;
      // candidate c -> (region, inner)
      int r = 0, k = -1, acc = 0;
      bool valid = false;
      for (int rr = 0; rr < T.n_regions; rr++) {
        const int cnt = T.regions[rr].kind == EARB_REGION_NGON ? T.ngon[T.regions[rr].index].n : 1;
        if (c < acc + cnt) {
          r = rr;
          k = T.regions[rr].kind == EARB_REGION_NGON ? c - acc : -1;
          valid = true;
          break;
        }
        acc += cnt;
      }
      bool ok = false;
      if (valid) {
        for (int a = 0; a < A; a++) aug[a] = 0.0;
        ok = psp_try_candidate(T, r, k, x, y, z, aug);
      }
      const unsigned won = __ballot_sync(0xffffffffu, ok);
      if (won) {
        const int first = __ffs((int)won) - 1;
        if (lane == first) psp_finish(T, aug, out);
        hit.region = __shfl_sync(0xffffffffu, r, first);
        hit.inner = __shfl_sync(0xffffffffu, k, first);
        return hit;
      }
    }
    // no region accepted the position (never happens on the built-in layouts): the serial scan's
    // result for that case is the normalisation of an all-zero vector
    if (lane == 0) {
      for (int a = 0; a < A; a++) aug[a] = 0.0;
      psp_finish(T, aug, out);
    }
    return hit;
  }

}  // namespace earb
