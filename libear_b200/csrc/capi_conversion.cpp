// C ABI: batched Cartesian <-> polar conversion of Objects metadata (ear::conversion,
// ref include/ear/conversion.hpp:14-71, src/conversion.cpp).  A pre-pass in front of the
// Objects gain calculator: Cartesian-authored blocks become the polar blocks it accepts.
#include <cmath>
#include <memory>

#include "capi_common.hpp"
#include "kernels.hpp"

using namespace earb;

struct earb200_conversion {
  int device = 0, n_sms = 0;
  ConversionSectors sectors;
  DeviceBuffer<double> d_pos, d_ext, d_pos_out, d_ext_out;
  DeviceBuffer<unsigned int> d_failed;
  int64_t launches = 0;
};

namespace {
  // the corner table of ref src/conversion.cpp:39-73: polar azimuth -> position on the square
  ConversionSectors make_sectors() {
    const double pi = 3.141592653589793238462643383279502884;
    const double corner_az[5] = {0.0, -30.0, -110.0, 110.0, 30.0};
    const double corner_xy[5][2] = {{0.0, 1.0}, {1.0, 1.0}, {1.0, -1.0}, {-1.0, -1.0}, {-1.0, 1.0}};
    ConversionSectors S;
    for (int i = 0; i < 5; i++) {
      S.polar_az[i] = corner_az[i];
      S.cart_pos[i][0] = corner_xy[i][0];
      S.cart_pos[i][1] = corner_xy[i][1];
      S.cart_az[i] = -(std::atan2(corner_xy[i][0], corner_xy[i][1]) * (180.0 / pi));  // ref geom.cpp:71-73
    }
    for (int i = 0; i < 5; i++) {
      const int j = (i + 1) % 5;
      // rows of the matrix to invert: start corner, end corner
      const double a = corner_xy[i][0], b = corner_xy[i][1], c = corner_xy[j][0], d = corner_xy[j][1];
      const double inv_det = 1.0 / (a * d - b * c);
      S.m[i][0] = d * inv_det;
      S.m[i][1] = -b * inv_det;
      S.m[i][2] = -c * inv_det;
      S.m[i][3] = a * inv_det;
    }
    return S;
  }

  void convert_device(earb200_conversion &h, int64_t n, bool to_polar, const double *d_pos, const double *d_ext,
                      double *d_pos_out, double *d_ext_out, cudaStream_t stream) {
    if (n < 0) throw Error(INVALID_ARGUMENT, "negative object count");
    if (n == 0) return;
    if (!d_pos || !d_pos_out) throw Error(INVALID_ARGUMENT, "null position buffer");
    if ((d_ext == nullptr) != (d_ext_out == nullptr))
      throw Error(INVALID_ARGUMENT, "extent input and output must both be given or both be null");
    use_device(h.device);
    launch_conversion(h.sectors, n, to_polar, d_pos, d_ext, d_pos_out, d_ext_out, h.d_failed.ptr, h.n_sms, stream);
    EARB_CUDA(cudaGetLastError());
    h.launches++;
  }

  // synchronises; throws what libear throws from find_*_sector / the p-range assertion
  void check_failed(earb200_conversion &h, cudaStream_t stream) {
    unsigned int failed = 0;
    EARB_CUDA(cudaMemcpyAsync(&failed, h.d_failed.ptr, sizeof(failed), cudaMemcpyDeviceToHost, stream));
    EARB_CUDA(cudaStreamSynchronize(stream));
    if (failed) {
      EARB_CUDA(cudaMemsetAsync(h.d_failed.ptr, 0, sizeof(unsigned int), stream));
      EARB_CUDA(cudaStreamSynchronize(stream));
      throw Error(INTERNAL_ERROR, "could not find sector / p should be 0 to 1 (" + std::to_string(failed) +
                                      " position(s) could not be converted; their outputs are NaN)");
    }
  }

  void convert_host(earb200_conversion &h, int64_t n, bool to_polar, const double *pos, const double *ext,
                    double *pos_out, double *ext_out) {
    if (n < 0) throw Error(INVALID_ARGUMENT, "negative object count");
    if (n == 0) return;
    if (!pos || !pos_out) throw Error(INVALID_ARGUMENT, "null position buffer");
    if ((ext == nullptr) != (ext_out == nullptr))
      throw Error(INVALID_ARGUMENT, "extent input and output must both be given or both be null");
    use_device(h.device);
    const size_t bytes = sizeof(double) * 3 * (size_t)n;
    double *dp = h.d_pos.reserve(3 * (size_t)n), *dpo = h.d_pos_out.reserve(3 * (size_t)n);
    double *de = nullptr, *deo = nullptr;
    EARB_CUDA(cudaMemcpyAsync(dp, pos, bytes, cudaMemcpyHostToDevice, 0));
    if (ext) {
      de = h.d_ext.reserve(3 * (size_t)n);
      deo = h.d_ext_out.reserve(3 * (size_t)n);
      EARB_CUDA(cudaMemcpyAsync(de, ext, bytes, cudaMemcpyHostToDevice, 0));
    }
    convert_device(h, n, to_polar, dp, de, dpo, deo, 0);
    EARB_CUDA(cudaMemcpyAsync(pos_out, dpo, bytes, cudaMemcpyDeviceToHost, 0));
    if (ext) EARB_CUDA(cudaMemcpyAsync(ext_out, deo, bytes, cudaMemcpyDeviceToHost, 0));
    check_failed(h, 0);
  }
}  // namespace

extern "C" {

int earb200_conversion_create(int device, earb200_conversion **out) {
  return guarded([&] {
    if (!out) throw Error(INVALID_ARGUMENT, "null argument");
    use_device(device);
    auto h = std::make_unique<earb200_conversion>();
    h->device = device;
    cudaDeviceProp prop;
    EARB_CUDA(cudaGetDeviceProperties(&prop, device));
    h->n_sms = prop.multiProcessorCount;
    h->sectors = make_sectors();
    EARB_CUDA(cudaMemset(h->d_failed.reserve(1), 0, sizeof(unsigned int)));
    *out = h.release();
  });
}

void earb200_conversion_destroy(earb200_conversion *conv) { delete conv; }

int64_t earb200_conversion_launch_count(const earb200_conversion *conv) { return conv ? conv->launches : 0; }

int earb200_conversion_cart_to_polar_host(earb200_conversion *conv, int64_t n, const double *pos,
                                          const double *extent, double *pos_out, double *extent_out) {
  return guarded([&] {
    if (!conv) throw Error(INVALID_ARGUMENT, "null handle");
    convert_host(*conv, n, true, pos, extent, pos_out, extent_out);
  });
}

int earb200_conversion_polar_to_cart_host(earb200_conversion *conv, int64_t n, const double *pos,
                                          const double *extent, double *pos_out, double *extent_out) {
  return guarded([&] {
    if (!conv) throw Error(INVALID_ARGUMENT, "null handle");
    convert_host(*conv, n, false, pos, extent, pos_out, extent_out);
  });
}

int earb200_conversion_cart_to_polar_device(earb200_conversion *conv, int64_t n, const double *d_pos,
                                            const double *d_extent, double *d_pos_out, double *d_extent_out,
                                            void *stream) {
  return guarded([&] {
    if (!conv) throw Error(INVALID_ARGUMENT, "null handle");
    convert_device(*conv, n, true, d_pos, d_extent, d_pos_out, d_extent_out, (cudaStream_t)stream);
  });
}

int earb200_conversion_polar_to_cart_device(earb200_conversion *conv, int64_t n, const double *d_pos,
                                            const double *d_extent, double *d_pos_out, double *d_extent_out,
                                            void *stream) {
  return guarded([&] {
    if (!conv) throw Error(INVALID_ARGUMENT, "null handle");
    convert_device(*conv, n, false, d_pos, d_extent, d_pos_out, d_extent_out, (cudaStream_t)stream);
  });
}

int earb200_conversion_check(earb200_conversion *conv, void *stream) {
  return guarded([&] {
    if (!conv) throw Error(INVALID_ARGUMENT, "null handle");
    use_device(conv->device);
    check_failed(*conv, (cudaStream_t)stream);
  });
}

}  // extern "C"
