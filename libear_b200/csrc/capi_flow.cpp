// C ABI: the Objects signal flow of libear's docs (ref docs/dsp.rst:40-71,
// include/ear/gain_calculators.hpp:43-56) as one device pipeline:
//
//   direct bus  = GainInterpolator<LinearInterpMatrix>(direct gains)   over the N object channels
//   diffuse bus = GainInterpolator<LinearInterpMatrix>(diffuse gains)
//   out = DelayBuffer(direct bus, decorrelatorCompensationDelay())
//       + BlockConvolver(diffuse bus, designDecorrelators(layout))      per loudspeaker channel
//
// Everything between the object audio and the loudspeaker feeds stays in HBM: two render
// launches and one convolver launch whose store adds the delayed direct bus.
#include <cstdlib>
#include <memory>
#include <vector>

#include "capi_common.hpp"

using namespace earb;

struct earb200_objects_renderer {
  int device = 0;
  int N = 0, C = 0;
  int64_t B = 512;
  earb200_interp *direct = nullptr, *diffuse = nullptr;
  earb200_convolver *conv = nullptr;
  earb200_delay *delay = nullptr;
  DeviceBuffer<float> d_direct, d_diffuse, d_in, d_out;
  ~earb200_objects_renderer() {
    if (direct) earb200_interp_destroy(direct);
    if (diffuse) earb200_interp_destroy(diffuse);
    if (conv) earb200_convolver_destroy(conv);
    if (delay) earb200_delay_destroy(delay);
  }
};

// capi_render.cpp
struct earb200_interp;
void earb_interp_share_segmentation(earb200_interp &a, earb200_interp &b);
void earb_interp_process_pair_device(earb200_interp &a, earb200_interp &b, int64_t block_start, int64_t nsamples,
                                     const float *d_in, int64_t in_stride, float *d_out_a, float *d_out_b,
                                     int64_t out_stride, cudaStream_t stream);

namespace {
  void chk(int rc) {
    if (rc != EARB200_OK) throw Error((Status)rc, last_error());
  }

  void flow_process_device(earb200_objects_renderer &h, int64_t block_start, int64_t n_blocks, const float *d_in,
                           int64_t in_stride, float *d_out, int64_t out_stride, cudaStream_t stream) {
    if (n_blocks < 0) throw Error(INVALID_ARGUMENT, "negative block count");
    if (n_blocks == 0) return;
    if (!d_in || !d_out) throw Error(INVALID_ARGUMENT, "null buffer");
    use_device(h.device);
    const int64_t T = n_blocks * h.B;
    const size_t bus = (size_t)h.C * (size_t)T;
    float *dir = h.d_direct.reserve(bus), *dif = h.d_diffuse.reserve(bus);
    // both renders on one launch: every input tile is read from HBM once (the second unit of a tile
    // runs on the neighbouring SM at the same time and finds it in L2)
    static const bool paired = [] {
      const char *e = std::getenv("EARB200_FLOW_PAIRED");
      return !e || std::atoi(e) != 0;
    }();
    if (paired) {
      earb_interp_process_pair_device(*h.direct, *h.diffuse, block_start, T, d_in, in_stride, dir, dif, T, stream);
    } else {
      chk(earb200_interp_process_device(h.direct, block_start, T, d_in, in_stride, dir, T, stream));
      chk(earb200_interp_process_device(h.diffuse, block_start, T, d_in, in_stride, dif, T, stream));
    }
    // decorrelators + compensation delay + sum in one pass: the delayed direct bus joins in the
    // convolver's store
    chk(earb200_convolver_process_add_delayed_device(h.conv, h.delay, n_blocks, dif, T, dir, T, d_out, out_stride,
                                                     stream));
  }
}  // namespace

extern "C" {

int earb200_objects_renderer_create(const earb200_layout *layout, int n_objects, int device,
                                    earb200_objects_renderer **out) {
  return guarded([&] {
    if (!layout || !out) throw Error(INVALID_ARGUMENT, "null argument");
    if (n_objects < 1) throw Error(INVALID_ARGUMENT, "invalid object count");
    use_device(device);
    auto h = std::make_unique<earb200_objects_renderer>();
    h->device = device;
    h->N = n_objects;
    h->C = (int)layout->layout.channels.size();
    chk(earb200_interp_create(h->N, h->C, device, &h->direct));
    chk(earb200_interp_create(h->N, h->C, device, &h->diffuse));
    // one decorrelation filter per loudspeaker channel (512 taps = one partition of B = 512)
    const int taps = earb200_decorrelator_size();
    std::vector<float> filt((size_t)h->C * taps);
    chk(earb200_design_decorrelators(layout, filt.data(), nullptr));
    chk(earb200_convolver_create(h->B, h->C, (taps + h->B - 1) / h->B, device, &h->conv));
    for (int c = 0; c < h->C; c++) {
      int id = -1, blocks = 0;
      chk(earb200_convolver_add_filter(h->conv, taps, filt.data() + (size_t)c * taps, &id, &blocks));
      chk(earb200_convolver_set_filter(h->conv, c, id));
    }
    chk(earb200_delay_create(h->C, earb200_decorrelator_compensation_delay(), device, &h->delay));
    *out = h.release();
  });
}

void earb200_objects_renderer_destroy(earb200_objects_renderer *r) { delete r; }

int earb200_objects_renderer_num_outputs(const earb200_objects_renderer *r) { return r ? r->C : 0; }
int64_t earb200_objects_renderer_launch_count(const earb200_objects_renderer *r) {
  if (!r) return 0;
  return earb200_interp_launch_count(r->direct) + earb200_interp_launch_count(r->diffuse) +
         earb200_convolver_launch_count(r->conv) + earb200_delay_launch_count(r->delay);
}
int64_t earb200_objects_renderer_block_size(const earb200_objects_renderer *r) { return r ? r->B : 0; }
int earb200_objects_renderer_delay(const earb200_objects_renderer *r) {
  return r ? earb200_decorrelator_compensation_delay() : 0;
}

int earb200_objects_renderer_set_gains(earb200_objects_renderer *r, int64_t n_points, const int64_t *sample_idx,
                                       const float *direct, const float *diffuse) {
  return guarded([&] {
    if (!r || !sample_idx || !direct || !diffuse) throw Error(INVALID_ARGUMENT, "null argument");
    chk(earb200_interp_set_points(r->direct, n_points, sample_idx, direct));
    chk(earb200_interp_set_points(r->diffuse, n_points, sample_idx, diffuse));
    earb_interp_share_segmentation(*r->direct, *r->diffuse);
  });
}

int earb200_objects_renderer_set_gains_device(earb200_objects_renderer *r, int64_t n_points,
                                              const int64_t *sample_idx, const float *d_direct,
                                              const float *d_diffuse, void *stream) {
  return guarded([&] {
    if (!r || !sample_idx || !d_direct || !d_diffuse) throw Error(INVALID_ARGUMENT, "null argument");
    chk(earb200_interp_set_points_device(r->direct, n_points, sample_idx, d_direct, stream));
    chk(earb200_interp_set_points_device(r->diffuse, n_points, sample_idx, d_diffuse, stream));
    earb_interp_share_segmentation(*r->direct, *r->diffuse);
  });
}

int earb200_objects_renderer_process_device(earb200_objects_renderer *r, int64_t block_start, int64_t n_blocks,
                                            const float *d_in, int64_t in_stride, float *d_out,
                                            int64_t out_stride, void *stream) {
  return guarded([&] {
    if (!r) throw Error(INVALID_ARGUMENT, "null handle");
    flow_process_device(*r, block_start, n_blocks, d_in, in_stride, d_out, out_stride, (cudaStream_t)stream);
  });
}

int earb200_objects_renderer_process_host(earb200_objects_renderer *r, int64_t block_start, int64_t n_blocks,
                                          const float *in, int64_t in_stride, float *out, int64_t out_stride) {
  return guarded([&] {
    if (!r) throw Error(INVALID_ARGUMENT, "null handle");
    if (n_blocks < 0) throw Error(INVALID_ARGUMENT, "negative block count");
    if (n_blocks == 0) return;
    if (!in || !out) throw Error(INVALID_ARGUMENT, "null buffer");
    use_device(r->device);
    const int64_t T = n_blocks * r->B;
    float *di = r->d_in.reserve((size_t)r->N * T), *dout = r->d_out.reserve((size_t)r->C * T);
    EARB_CUDA(cudaMemcpy2DAsync(di, sizeof(float) * T, in, sizeof(float) * in_stride, sizeof(float) * T, r->N,
                                cudaMemcpyHostToDevice, 0));
    flow_process_device(*r, block_start, n_blocks, di, T, dout, T, 0);
    EARB_CUDA(cudaMemcpy2DAsync(out, sizeof(float) * out_stride, dout, sizeof(float) * T, sizeof(float) * T, r->C,
                                cudaMemcpyDeviceToHost, 0));
    EARB_CUDA(cudaStreamSynchronize(0));
  });
}

}  // extern "C"
