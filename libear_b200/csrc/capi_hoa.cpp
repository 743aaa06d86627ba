// C ABI: HOA gain calculator (K5) and decorrelator design (setup-time, host).
#include <cmath>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "capi_common.hpp"
#include "kernels.hpp"

using namespace earb;

// implemented in capi_objects.cpp
int earb200_panner_handle_device(earb200_panner *, int64_t, const double *, double *, int32_t *,
                                 int32_t *, void *);

struct earb200_hoa {
  int device = 0;
  HostLayout layout;
  earb200_panner *panner = nullptr;
  int S = 0, n_out = 0, P = 0;
  std::vector<int> out_index;
  DeviceBuffer<double> d_xyz, d_G, d_Y, d_D, d_partial, d_norms;
  DeviceBuffer<int> d_orders, d_degrees;
  ~earb200_hoa() {
    if (panner) earb200_panner_destroy(panner);
  }
};

namespace {
  double factorial(int n) {
    double r = 1.0;
    for (int i = 2; i <= n; i++) r *= (double)i;
    return r;
  }
  // HOA normalisations, BS.2076-1 section 10.2 (ref src/hoa/hoa.hpp:38-65)
  double norm_N3D(int n, int abs_m) {
    return std::sqrt((2.0 * n + 1.0) * factorial(n - abs_m) / factorial(n + abs_m));
  }
  double norm_SN3D(int n, int abs_m) { return std::sqrt(factorial(n - abs_m) / factorial(n + abs_m)); }
  double norm_FuMa(int n, int abs_m) {
    static const double f[4][4] = {{1.0 / std::sqrt(2.0), 0, 0, 0},
                                   {1.0, 1.0, 0, 0},
                                   {1.0, 2.0 / std::sqrt(3.0), 2.0 / std::sqrt(3.0), 0},
                                   {1.0, std::sqrt(45.0 / 32.0), 3.0 / std::sqrt(5.0), std::sqrt(8.0 / 5.0)}};
    if (n > 3) throw Error(INVALID_ARGUMENT, "FuMa normalization is only defined up to order 3");
    return f[n][abs_m] * norm_SN3D(n, abs_m);
  }
}  // namespace

extern "C" {

int earb200_hoa_create(const earb200_layout *layout, int device, earb200_hoa **out) {
  return guarded([&] {
    if (!layout || !out) throw Error(INVALID_ARGUMENT, "null argument");
    use_device(device);
    auto h = std::make_unique<earb200_hoa>();
    h->device = device;
    h->layout = layout->layout;
    int rc = earb200_panner_create(layout, device, &h->panner);
    if (rc != EARB200_OK) throw Error((Status)rc, last_error());
    h->S = earb200_panner_num_channels(h->panner);
    h->n_out = (int)layout->layout.channels.size();
    for (int c = 0; c < h->n_out; c++)
      if (!layout->layout.channels[c].lfe) h->out_index.push_back(c);
    // G_virt: point-source gains at the t-design points (ref gain_calculator_hoa.cpp:12-14)
    const std::vector<Vec3> &pts = hoaDesignPoints();
    h->P = (int)pts.size();
    std::vector<double> xyz(3 * pts.size());
    for (size_t p = 0; p < pts.size(); p++) {
      xyz[3 * p] = pts[p].x;
      xyz[3 * p + 1] = pts[p].y;
      xyz[3 * p + 2] = pts[p].z;
    }
    EARB_CUDA(cudaMemcpy(h->d_xyz.reserve(xyz.size()), xyz.data(), sizeof(double) * xyz.size(),
                         cudaMemcpyHostToDevice));
    h->d_G.reserve((size_t)h->P * h->S);
    rc = earb200_panner_handle_device(h->panner, h->P, h->d_xyz.ptr, h->d_G.ptr, nullptr, nullptr, nullptr);
    if (rc != EARB200_OK) throw Error((Status)rc, last_error());
    EARB_CUDA(cudaDeviceSynchronize());
    *out = h.release();
  });
}
void earb200_hoa_destroy(earb200_hoa *hoa) { delete hoa; }

int earb200_hoa_calculate(earb200_hoa *hoa, int n_channels, const int *orders, const int *degrees,
                          const char *normalization, double nfc_ref_dist, int screen_ref,
                          double *decode, uint32_t *warnings) {
  return guarded([&] {
    if (!hoa || !normalization || n_channels < 0 || (n_channels > 0 && (!orders || !degrees || !decode)))
      throw Error(INVALID_ARGUMENT, "null argument");
    earb200_hoa &h = *hoa;
    const int C = n_channels;
    for (int i = 0; i < C; i++) {
      if (orders[i] < 0) throw Error(INVALID_ARGUMENT, "orders must not be negative");
      if (std::abs(degrees[i]) > orders[i])
        throw Error(INVALID_ARGUMENT, "magnitude of degree must not be greater than order");
    }
    const std::string norm_name(normalization);
    double (*norm)(int, int) = nullptr;
    if (norm_name == "N3D")
      norm = norm_N3D;
    else if (norm_name == "SN3D")
      norm = norm_SN3D;
    else if (norm_name == "FuMa")
      norm = norm_FuMa;
    else
      throw Error(ADM_ERROR, "ADM error: unknown normalization type: '" + norm_name + "'");
    uint32_t w = 0;
    if (screen_ref) w |= 1u << EARB200_WARN_HOA_SCREENREF;
    if (nfc_ref_dist != 0.0) w |= 1u << EARB200_WARN_HOA_NFCREFDIST;
    if (warnings) *warnings = w;
    if (C == 0) return;

    use_device(h.device);
    std::vector<double> y_norm(C), conv(C);
    for (int i = 0; i < C; i++) {
      const int am = std::abs(degrees[i]);
      y_norm[i] = norm_N3D(orders[i], am);
      conv[i] = norm_N3D(orders[i], am) / norm(orders[i], am);  // ref hoa.hpp:147-160
    }
    const int P = h.P, S = h.S, NB = hoa_power_blocks(P);
    EARB_CUDA(cudaMemcpy(h.d_orders.reserve(C), orders, sizeof(int) * C, cudaMemcpyHostToDevice));
    EARB_CUDA(cudaMemcpy(h.d_degrees.reserve(C), degrees, sizeof(int) * C, cudaMemcpyHostToDevice));
    EARB_CUDA(cudaMemcpy(h.d_norms.reserve(C), y_norm.data(), sizeof(double) * C, cudaMemcpyHostToDevice));
    h.d_Y.reserve((size_t)C * P);
    h.d_D.reserve((size_t)S * C);
    h.d_partial.reserve(NB);
    launch_hoa_design(P, S, C, h.d_xyz.ptr, h.d_G.ptr, h.d_orders.ptr, h.d_degrees.ptr, h.d_norms.ptr,
                      h.d_Y.ptr, h.d_D.ptr, h.d_partial.ptr, 0);
    EARB_CUDA(cudaGetLastError());
    std::vector<double> D((size_t)S * C), partial(NB);
    EARB_CUDA(cudaMemcpy(D.data(), h.d_D.ptr, sizeof(double) * D.size(), cudaMemcpyDeviceToHost));
    EARB_CUDA(cudaMemcpy(partial.data(), h.d_partial.ptr, sizeof(double) * NB, cudaMemcpyDeviceToHost));
    double power = 0.0;
    for (double v : partial) power += v;
    // D *= sqrt(P) / ||D Y||_F, then per-channel normalisation conversion, then LFE rows
    const double scale = std::sqrt((double)P) / std::sqrt(power);
    for (int c = 0; c < C; c++) {
      for (int o = 0; o < h.n_out; o++) decode[(size_t)c * h.n_out + o] = 0.0;
      for (int s = 0; s < S; s++)
        decode[(size_t)c * h.n_out + h.out_index[s]] = D[(size_t)s * C + c] * scale * conv[c];
    }
  });
}

}  // extern "C"
