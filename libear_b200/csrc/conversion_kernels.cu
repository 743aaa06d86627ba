// Batched Cartesian <-> polar conversion of Objects metadata (position + extent), one
// thread per object, FP64.  Restates ear::conversion (ref src/conversion.cpp:24-263: the
// five azimuth sectors of the "cube to sphere" warp, the elevation warp at 30 / 45 degrees,
// whd2xyz / xyz2whd).  Compiled with --fmad=false so that products and sums round like the
// CPU build; the transcendental functions are CUDA's (<= 2 ulp), so results agree with
// libear to ~1e-12, not bit for bit.
//
// Layout: positions [n][3] and extents [n][3] (width, height, depth) as the ADM block
// formats carry them; 48 B in + 48 B out per object -- a pre-pass in front of K1/K2.
#include <cuda_runtime.h>

#include "kernels.hpp"

namespace earb {

  namespace {
    constexpr double kPi = 3.141592653589793238462643383279502884;
    constexpr double kElTop = 30.0, kElTopTilde = 45.0;  // ref conversion.cpp:75-76

    __device__ __forceinline__ double c_radians(double d) { return d * (kPi / 180.0); }
    __device__ __forceinline__ double c_degrees(double r) { return r * (180.0 / kPi); }
    __device__ __forceinline__ double c_sign(double x) { return x < 0.0 ? -1.0 : (x > 0.0 ? 1.0 : 0.0); }

    // angles this far out would spin the reference's wrap loops for minutes; refuse them
    __device__ __forceinline__ bool angle_ok(double a) { return fabs(a) <= 1.0e7; }

    // y moved by multiples of 360 into [x, x + 360)  (ref src/common/geom.hpp:31-39)
    __device__ __forceinline__ double relative_angle(double x, double y) {
      while (y - 360.0 >= x) y -= 360.0;
      while (y < x) y += 360.0;
      return y;
    }

    // is x inside [start, end] going clockwise?  (ref src/common/geom.cpp:7-28, tol = 0)
    __device__ __forceinline__ bool inside_angle_range(double x, double start, double end) {
      while (end - 360.0 > start) end -= 360.0;
      while (end < start) end += 360.0;
      while (x - 360.0 >= start) x -= 360.0;
      while (x < start) x += 360.0;
      return x <= end;
    }

    // ref conversion.cpp:96-123
    __device__ __forceinline__ double map_az_to_linear(double left_az, double right_az, double az) {
      const double mid_az = (left_az + right_az) / 2.0;
      const double az_range = right_az - mid_az;
      const double rel_az = az - mid_az;
      const double gain_r = 0.5 + 0.5 * tan(c_radians(rel_az)) / tan(c_radians(az_range));
      return atan2(gain_r, 1.0 - gain_r) * (2.0 / kPi);
    }
    __device__ __forceinline__ double map_linear_to_az(double left_az, double right_az, double x) {
      const double mid_az = (left_az + right_az) / 2.0;
      const double az_range = right_az - mid_az;
      const double gain_l_ = cos(x * (kPi / 2.0));
      const double gain_r_ = sin(x * (kPi / 2.0));
      const double gain_r = gain_r_ / (gain_l_ + gain_r_);
      const double rel_az = c_degrees(atan(2.0 * (gain_r - 0.5) * tan(c_radians(az_range))));
      return mid_az + rel_az;
    }

    // rows of localCoordinateSystem(az, el) (ref src/common/geom.hpp:91-98, geom.cpp:82-87)
    __device__ __forceinline__ void cart_row(double az, double el, double (&r)[3]) {
      const double saz = sin(c_radians(-az)), caz = cos(c_radians(-az));
      const double sel = sin(c_radians(el)), cel = cos(c_radians(el));
      r[0] = saz * cel * 1.0;
      r[1] = caz * cel * 1.0;
      r[2] = sel * 1.0;
    }
    __device__ __forceinline__ void local_coordinate_system(double az, double el, double (&m)[3][3]) {
      cart_row(az - 90.0, 0.0, m[0]);
      cart_row(az, el, m[1]);
      cart_row(az, el + 90.0, m[2]);
    }
    // Eigen's stableNorm-free .norm() of a 3-vector: sqrt of the sum in index order
    __device__ __forceinline__ double norm3(double a, double b, double c) { return sqrt(a * a + b * b + c * c); }

    // ref conversion.cpp:127-157; false = the reference's assertion / sector search fails
    __device__ bool point_polar_to_cart(const ConversionSectors &S, double az, double el, double dist,
                                        double (&out)[3]) {
      double r_xy, z;
      if (fabs(el) > kElTop) {
        const double el_tilde = kElTopTilde + (90.0 - kElTopTilde) * (fabs(el) - kElTop) / (90.0 - kElTop);
        z = dist * c_sign(el);
        r_xy = dist * tan(c_radians(90.0 - el_tilde));
      } else {
        const double el_tilde = kElTopTilde * el / kElTop;
        z = tan(c_radians(el_tilde)) * dist;
        r_xy = dist;
      }
      if (!angle_ok(az)) return false;
      int s = -1;
#pragma unroll
      for (int k = 4; k >= 0; k--)
        if (inside_angle_range(az, S.polar_az[(k + 1) % 5], S.polar_az[k])) s = k;
      if (s < 0) return false;
      const double p_start = S.polar_az[s], p_end = S.polar_az[(s + 1) % 5];
      const double rel_az = relative_angle(p_end, az);
      const double rel_left_az = relative_angle(p_end, p_start);
      const double p = map_az_to_linear(rel_left_az, p_end, rel_az);
      if (!(-1e-6 <= p && p <= 1.0 + 1e-6)) return false;
      const double sx = S.cart_pos[s][0], sy = S.cart_pos[s][1];
      const double ex = S.cart_pos[(s + 1) % 5][0], ey = S.cart_pos[(s + 1) % 5][1];
      out[0] = r_xy * (sx + (ex - sx) * p);
      out[1] = r_xy * (sy + (ey - sy) * p);
      out[2] = z;
      return true;
    }

    // ref conversion.cpp:159-194
    __device__ bool point_cart_to_polar(const ConversionSectors &S, double X, double Y, double Z,
                                        double (&out)[3]) {
      const double eps = 1e-10;
      if (fabs(X) < eps && fabs(Y) < eps) {
        if (fabs(Z) < eps) {
          out[0] = 0.0; out[1] = 0.0; out[2] = 0.0;
        } else {
          out[0] = 0.0; out[1] = c_sign(Z) * 90.0; out[2] = fabs(Z);
        }
        return true;
      }
      const double az_c = -c_degrees(atan2(X, Y));  // ref geom.cpp:71-73
      if (!(az_c == az_c)) return false;
      int s = -1;
#pragma unroll
      for (int k = 4; k >= 0; k--)
        if (inside_angle_range(az_c, S.cart_az[(k + 1) % 5], S.cart_az[k])) s = k;
      if (s < 0) return false;
      // g_lr = [X Y] * m
      const double g_l = X * S.m[s][0] + Y * S.m[s][2];
      const double g_r = X * S.m[s][1] + Y * S.m[s][3];
      const double r_xy = g_l + g_r;
      const double p_start = S.polar_az[s], p_end = S.polar_az[(s + 1) % 5];
      const double rel_left_az = relative_angle(p_end, p_start);
      double az = map_linear_to_az(rel_left_az, p_end, g_r / r_xy);
      if (!angle_ok(az)) return false;
      az = relative_angle(-180.0, az);
      const double el_tilde = c_degrees(atan(Z / r_xy));
      double d, el;
      if (fabs(el_tilde) > kElTopTilde) {
        const double abs_el = kElTop + (90.0 - kElTop) * (fabs(el_tilde) - kElTopTilde) / (90.0 - kElTopTilde);
        el = c_sign(el_tilde) * abs_el;
        d = fabs(Z);
      } else {
        el = kElTop * el_tilde / kElTopTilde;
        d = r_xy;
      }
      out[0] = az; out[1] = el; out[2] = d;
      return true;
    }

    // std::max(a, b) keeps a when the comparison is false (NaN b), unlike fmax
    __device__ __forceinline__ double std_max(double a, double b) { return a < b ? b : a; }

    // ref conversion.cpp:198-213; returns x, y, z sizes
    __device__ __forceinline__ void whd2xyz(double w, double h, double d, double (&o)[3]) {
      const double x_size_width = w < 180.0 ? sin(c_radians(w / 2.0)) : 1.0;
      const double y_size_width = (1.0 - cos(c_radians(w / 2.0))) / 2.0;
      const double z_size_height = h < 180.0 ? sin(c_radians(h / 2.0)) : 1.0;
      const double y_size_height = (1.0 - cos(c_radians(h / 2.0))) / 2.0;
      o[0] = x_size_width;
      o[1] = std_max(std_max(y_size_width, y_size_height), d);
      o[2] = z_size_height;
    }
    // ref conversion.cpp:215-234
    __device__ __forceinline__ void xyz2whd(double s_x, double s_y, double s_z, double (&o)[3]) {
      const double width_from_sx = 2.0 * c_degrees(asin(s_x));
      const double width_from_sy = 2.0 * c_degrees(acos(1.0 - 2.0 * s_y));
      const double width = width_from_sx + s_x * std_max(width_from_sy - width_from_sx, 0.0);
      const double height_from_sz = 2.0 * c_degrees(asin(s_z));
      const double height_from_sy = 2.0 * c_degrees(acos(1.0 - 2.0 * s_y));
      const double height = height_from_sz + s_z * std_max(height_from_sy - height_from_sz, 0.0);
      double eq[3];
      whd2xyz(width, height, 0.0, eq);
      o[0] = width;
      o[1] = height;
      o[2] = std_max(0.0, s_y - eq[1]);
    }

    __global__ void __launch_bounds__(128)
    conversion_kernel(ConversionSectors S, long n, int to_polar, const double *__restrict__ pos,
                      const double *__restrict__ ext, double *__restrict__ pos_out,
                      double *__restrict__ ext_out, unsigned int *__restrict__ n_failed) {
      for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
        const double a = pos[3 * i], b = pos[3 * i + 1], c = pos[3 * i + 2];
        double o[3];
        bool ok;
        if (to_polar) {
          ok = point_cart_to_polar(S, a, b, c, o);
        } else {
          ok = point_polar_to_cart(S, a, b, c, o);
        }
        if (!ok) {
          atomicAdd(n_failed, 1u);
          const double nan = __longlong_as_double(0x7ff8000000000000LL);
          o[0] = o[1] = o[2] = nan;
        }
        pos_out[3 * i] = o[0];
        pos_out[3 * i + 1] = o[1];
        pos_out[3 * i + 2] = o[2];
        if (!ext) continue;
        const double w = ext[3 * i], h = ext[3 * i + 1], d = ext[3 * i + 2];
        double e[3], m[3][3];
        if (to_polar) {
          // extentCartToPolar (ref conversion.cpp:236-250): columns of diag(w, d, h) * LCS^T
          local_coordinate_system(o[0], o[1], m);
          const double sx = norm3(m[0][0] * w, m[0][1] * d, m[0][2] * h);
          const double sy = norm3(m[1][0] * w, m[1][1] * d, m[1][2] * h);
          const double sz = norm3(m[2][0] * w, m[2][1] * d, m[2][2] * h);
          xyz2whd(sx, sy, sz, e);
        } else {
          // extentPolarToCart (ref conversion.cpp:252-263): columns of diag(front_size) * LCS
          double f[3];
          whd2xyz(w, h, d, f);
          local_coordinate_system(a, b, m);
          const double sx = norm3(m[0][0] * f[0], m[1][0] * f[1], m[2][0] * f[2]);
          const double sy = norm3(m[0][1] * f[0], m[1][1] * f[1], m[2][1] * f[2]);
          const double sz = norm3(m[0][2] * f[0], m[1][2] * f[1], m[2][2] * f[2]);
          e[0] = sx;  // width  = x size
          e[1] = sz;  // height = z size
          e[2] = sy;  // depth  = y size
        }
        ext_out[3 * i] = e[0];
        ext_out[3 * i + 1] = e[1];
        ext_out[3 * i + 2] = e[2];
      }
    }
  }  // namespace

  void launch_conversion(const ConversionSectors &sectors, long n, bool to_polar, const double *d_pos,
                         const double *d_ext, double *d_pos_out, double *d_ext_out,
                         unsigned int *d_n_failed, int n_sms, cudaStream_t stream) {
    if (n <= 0) return;
    const int threads = 128;
    long blocks = (n + threads - 1) / threads;
    const long cap = (long)n_sms * 16;  // grid-stride beyond 16 resident CTAs per SM
    if (blocks > cap) blocks = cap;
    conversion_kernel<<<(unsigned)blocks, threads, 0, stream>>>(sectors, n, to_polar ? 1 : 0, d_pos, d_ext,
                                                                d_pos_out, d_ext_out, d_n_failed);
  }

}  // namespace earb
