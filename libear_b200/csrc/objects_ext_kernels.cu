// OPT-IN extensions of the batched Objects calculator: objectDivergence and channelLock for polar
// metadata.  libear 0.9 has NO arithmetic for them -- GainCalculatorObjectsImpl::calculate throws
// ear::not_implemented (ref src/object_based/gain_calculator_objects.cpp:37-43) and its tests
// require the throws -- so nothing here can be compared with the reference: PARITY UNPINNED.
// The default behaviour of the library stays the reference's (the throws); these run only after
// earb200_objects_enable_extensions.  The arithmetic restates Rec. ITU-R BS.2127-0 as the EBU ADM
// Renderer implements it (the renderer libear's docs point to, docs/index.rst); the ITU text is not
// in the reference tree, so the restatement is from the published algorithm, by hand:
//   channelLock   the object snaps to the nearest loudspeaker (Euclidean distance between the
//                 object's Cartesian position and the loudspeakers' unit-distance positions) if one
//                 lies within maxDistance; ties go to the smaller |elevation|, then the smaller
//                 |azimuth|, then the lower channel index.
//   divergence    value x in [0, 1], azimuthRange r: the object becomes three sources -- itself
//                 with power weight (1 - x) / (1 + x) and two at +-x r degrees of azimuth in the
//                 object's local frame (rows right / front / up at its position) with weight
//                 x / (1 + x) each; every source goes through the extent panner and the gains are
//                 combined in the power domain, g = sqrt(sum_k w_k g_k^2).
// FP64, compiled without FMA contraction like the rest of the decision code.
#include <cuda_runtime.h>

#include "kernels.hpp"

namespace earb {
  namespace {
    __device__ __forceinline__ double rad(double d) { return d * (3.141592653589793238462643383279502884 / 180.0); }
    __device__ __forceinline__ void cart(double az, double el, double dist, double &x, double &y, double &z) {
      // ref src/common/geom.cpp:82-87
      x = sin(rad(-az)) * cos(rad(el)) * dist;
      y = cos(rad(-az)) * cos(rad(el)) * dist;
      z = sin(rad(el)) * dist;
    }

    // one thread per object: 3 sources [3 i + k]: xyz, power weight; extents copied, gain 1, diffuse 0
    __global__ void objects_ext_expand_kernel(long n, ObjectsSoA in, const double *__restrict__ divergence,
                                              const double *__restrict__ az_range,
                                              const double *__restrict__ lock_max_distance, int n_spk,
                                              const double *__restrict__ spk /* [S][5]: x y z az el */,
                                              double *__restrict__ xyz3, double *__restrict__ w3,
                                              double *__restrict__ h3, double *__restrict__ d3,
                                              double *__restrict__ ones, double *__restrict__ zeros,
                                              double *__restrict__ weights) {
      const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
      if (i >= n) return;
      double az = in.azimuth[i], el = in.elevation[i], dist = in.distance[i];
      double px, py, pz;
      cart(az, el, dist, px, py, pz);
      // ---- channelLock
      if (lock_max_distance && !(lock_max_distance[i] < 0.0)) {  // negative: no lock; +inf: no limit
        const double tol = 1e-10, max_d = lock_max_distance[i];
        int best = -1;
        double best_d = 0.0;
        for (int s = 0; s < n_spk; s++) {
          const double dx = px - spk[5 * s], dy = py - spk[5 * s + 1], dz = pz - spk[5 * s + 2];
          const double d = sqrt((dx * dx + dy * dy) + dz * dz);
          if (!(d < max_d + tol)) continue;
          bool take = best < 0 || d < best_d - tol;
          if (!take && d < best_d + tol) {  // a tie: |elevation|, then |azimuth|, then the index
            const double ea = fabs(spk[5 * s + 4]), eb = fabs(spk[5 * best + 4]);
            const double aa = fabs(spk[5 * s + 3]), ab = fabs(spk[5 * best + 3]);
            take = ea < eb - tol || (ea < eb + tol && aa < ab - tol);
          }
          if (take) {
            best = s;
            best_d = d;
          }
        }
        if (best >= 0) {
          px = spk[5 * best], py = spk[5 * best + 1], pz = spk[5 * best + 2];
          az = spk[5 * best + 3], el = spk[5 * best + 4], dist = 1.0;
        }
      }
      // ---- divergence
      const double x = divergence ? fmin(fmax(divergence[i], 0.0), 1.0) : 0.0;
      const double r = az_range ? az_range[i] : 45.0;
      double lx, ly, lz, rx, ry, rz;
      cart(x * r, 0.0, dist, lx, ly, lz);
      cart(-x * r, 0.0, dist, rx, ry, rz);
      // rows of the local frame at (az, el): right, front, up (ref include/ear/common/geom.hpp:91-98)
      double m[9];
      cart(az - 90.0, 0.0, 1.0, m[0], m[1], m[2]);
      cart(az, el, 1.0, m[3], m[4], m[5]);
      cart(az, el + 90.0, 1.0, m[6], m[7], m[8]);
      double *o = xyz3 + 9 * i;
      o[0] = px, o[1] = py, o[2] = pz;
      for (int c = 0; c < 3; c++) {
        o[3 + c] = (lx * m[c] + ly * m[3 + c]) + lz * m[6 + c];
        o[6 + c] = (rx * m[c] + ry * m[3 + c]) + rz * m[6 + c];
      }
      weights[3 * i] = (1.0 - x) / (x + 1.0);
      weights[3 * i + 1] = weights[3 * i + 2] = x / (x + 1.0);
      for (int k = 0; k < 3; k++) {
        w3[3 * i + k] = in.width[i];
        h3[3 * i + k] = in.height[i];
        d3[3 * i + k] = in.depth[i];
        ones[3 * i + k] = 1.0;
        zeros[3 * i + k] = 0.0;
      }
    }

    // thread = (object, output channel): g = gain sqrt(sum_k w_k g_k^2), then the direct / diffuse
    // split (ref gain_calculator_objects.cpp:49-56)
    template <typename OutT>
    __global__ void objects_ext_combine_kernel(long n, int n_out, const double *__restrict__ g3 /* [3n][n_out] */,
                                               const double *__restrict__ weights, const double *__restrict__ gain,
                                               const double *__restrict__ diffuse, OutT *__restrict__ out_direct,
                                               OutT *__restrict__ out_diffuse) {
      const long e = (long)blockIdx.x * blockDim.x + threadIdx.x;
      if (e >= n * n_out) return;
      const long i = e / n_out;
      const int c = (int)(e - i * n_out);
      double acc = 0.0;
      for (int k = 0; k < 3; k++) {
        const double g = g3[(3 * i + k) * n_out + c];
        acc += weights[3 * i + k] * (g * g);
      }
      const double v = sqrt(acc) * gain[i];
      out_direct[e] = (OutT)(v * sqrt(1.0 - diffuse[i]));
      out_diffuse[e] = (OutT)(v * sqrt(diffuse[i]));
    }
  }  // namespace

  void launch_objects_ext_expand(long n, const ObjectsSoA &in, const double *divergence, const double *az_range,
                                 const double *lock_max_distance, int n_spk, const double *d_spk, double *xyz3,
                                 double *w3, double *h3, double *d3, double *ones, double *zeros, double *weights,
                                 cudaStream_t stream) {
    if (n <= 0) return;
    objects_ext_expand_kernel<<<(unsigned)((n + 127) / 128), 128, 0, stream>>>(
        n, in, divergence, az_range, lock_max_distance, n_spk, d_spk, xyz3, w3, h3, d3, ones, zeros, weights);
  }

  template <typename OutT>
  void launch_objects_ext_combine(long n, int n_out, const double *g3, const double *weights, const double *gain,
                                  const double *diffuse, OutT *out_direct, OutT *out_diffuse, cudaStream_t stream) {
    if (n <= 0) return;
    const long total = n * n_out;
    objects_ext_combine_kernel<OutT><<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(
        n, n_out, g3, weights, gain, diffuse, out_direct, out_diffuse);
  }
  template void launch_objects_ext_combine<float>(long, int, const double *, const double *, const double *,
                                                  const double *, float *, float *, cudaStream_t);
  template void launch_objects_ext_combine<double>(long, int, const double *, const double *, const double *,
                                                   const double *, double *, double *, cudaStream_t);
}  // namespace earb
