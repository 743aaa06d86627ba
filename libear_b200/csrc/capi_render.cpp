// C ABI: gain interpolator (K3).  Host side = the segmentation logic of
// GainInterpolator::process / find_block (ref include/ear/dsp/gain_interpolator.hpp:53-129),
// turned into a table of segments and work items that one kernel launch consumes.
#include <algorithm>
#include <cstdint>
#include <cstdlib>
#include <memory>
#include <string>
#include <vector>

#include "capi_common.hpp"
#include "kernels.hpp"

using namespace earb;

struct earb200_interp {
  int device = 0;
  int N = 0, M = 0;
  // interp_points (host copy) + equality of neighbouring points (constant_interp, :139-141)
  std::vector<int64_t> sample_idx;
  std::vector<char> same_as_prev;
  bool sorted = true;
  DeviceBuffer<float> d_gains;
  DeviceBuffer<float> d_images;  // tcgen05 path: per-point TF32 hi/lo images in UMMA layout
  DeviceBuffer<char> d_same;     // set_points_device: equality flags of neighbouring points
  int64_t n_points = 0;
  // plan cache keyed by (block_start, nsamples, buffer alignment): a caller that walks the same
  // chunks every pass (chunked mixing with the bus reduce of chunk k behind the render of chunk
  // k + 1) pays the host-side segmentation and its upload once per chunk, not once per call
  struct Plan {
    int64_t start = 0, n = -1;
    int align = -1;
    long n_items[4] = {0, 0, 0, 0};  // fast ramp, fast const, slow ramp, slow const
    DeviceBuffer<RenderItem> d_items;
    DeviceBuffer<RenderSegment> d_segs;
    uint64_t last_use = 0;
    cudaEvent_t used = nullptr;  // recorded after the last launch that read the buffers
    ~Plan() {
      if (used) cudaEventDestroy(used);
    }
  };
  std::vector<std::unique_ptr<Plan>> plans;
  uint64_t plan_clock = 0;
  // staging for process_host
  DeviceBuffer<float> d_in, d_out;
  int64_t launches = 0;
  KernelTimer timer;
  int n_sms = 0;
  int accumulate = 0;  // process() adds into `out` instead of overwriting it
  int use_tc = 0;  // tensor-core (tcgen05) fast path; EARB200_RENDER_PATH=cuda selects the FFMA kernel
};

// capi_bus.cpp: the slots this rank stores to in the current epoch
struct earb200_bus;
earb::RenderBus earb200_bus_map(const earb200_bus &bus, int64_t t_origin);
int64_t earb200_bus_capacity(const earb200_bus *bus);
int earb200_bus_num_channels(const earb200_bus *bus);
void earb200_bus_count_launch(earb200_bus *bus);

namespace {
  // align: sample phase p in [0, 4) such that call-relative samples t with (t + p) % 4 == 0
  // are 16-byte aligned in both buffers, or -1 when no such phase exists
  constexpr size_t kMaxPlans = 64;

  earb200_interp::Plan &build_plan(earb200_interp &h, int64_t block_start, int64_t nsamples, int align,
                                   cudaStream_t stream) {
    for (auto &p : h.plans)
      if (p->n == nsamples && p->start == block_start && p->align == align) {
        p->last_use = ++h.plan_clock;
        // the tables were uploaded on the stream of an earlier call: order this one behind it
        if (p->used) EARB_CUDA(cudaStreamWaitEvent(stream, p->used, 0));
        return *p;
      }
    const int fast_tile = h.use_tc ? RENDER_TC_TILE : RENDER_PIPE_TILE;
    const int64_t P = h.n_points;
    if (P == 0) throw Error(INVALID_ARGUMENT, "no interpolation points");
    if (!h.sorted) throw Error(INVALID_ARGUMENT, "interpolation points are not sorted");
    std::vector<RenderSegment> segs;
    std::vector<RenderItem> items[4];
    const int tile = render_tile_samples(h.M);
    const int64_t block_end = block_start + nsamples;
    int64_t cur = block_start;
    while (cur < block_end) {
      // block index = number of points at or before `cur`
      const int64_t idx = std::upper_bound(h.sample_idx.begin(), h.sample_idx.end(), cur) -
                          h.sample_idx.begin();
      const int64_t end = (idx == P) ? block_end : std::min(h.sample_idx[idx], block_end);
      RenderSegment s{};
      const bool constant = (idx == 0 || idx == P || h.same_as_prev[idx]);
      if (constant) {
        s.ramp = 0;
        s.pa = s.pb = (int)((idx == P) ? idx - 1 : idx);
        s.ramp_start = 0;
        s.scale = 0.0f;
      } else {
        s.ramp = 1;
        s.pa = (int)(idx - 1);
        s.pb = (int)idx;
        s.ramp_start = h.sample_idx[idx - 1] - block_start;
        s.scale = 1.0f / (float)(h.sample_idx[idx] - h.sample_idx[idx - 1]);
      }
      const int seg_id = (int)segs.size();
      segs.push_back(s);
      auto emit = [&](int64_t t, int64_t len, bool fast) {
        RenderItem it;
        it.t0 = t - block_start;
        it.len = (int)len;
        it.seg = seg_id;
        items[(fast ? 0 : 2) + (constant ? 1 : 0)].push_back(it);
      };
      int64_t t = cur;
      const int64_t head = (4 - (((t - block_start) + align) & 3)) & 3;
      if (align >= 0 && end - (t + head) >= fast_tile) {
        // ragged head up to the first 16-byte aligned sample, then full TMA tiles
        if (head) emit(t, head, false);
        t += head;
        for (; end - t >= fast_tile; t += fast_tile) emit(t, fast_tile, true);
      }
      for (; t < end; t += tile) emit(t, std::min<int64_t>(tile, end - t), false);
      cur = end;
    }
    // a streaming caller never repeats a key (block_start advances): the least recently used plan is
    // recycled together with its device buffers (grow-only), so steady state allocates nothing, and
    // the tables go up with asynchronous copies on the caller's stream -- ordered before the launch
    // that reads them, and (through the plan's event) after the last launch that read the buffers'
    // previous contents, whatever stream that was on
    std::unique_ptr<earb200_interp::Plan> plan;
    if (h.plans.size() >= kMaxPlans) {
      size_t lru = 0;
      for (size_t i = 1; i < h.plans.size(); i++)
        if (h.plans[i]->last_use < h.plans[lru]->last_use) lru = i;
      plan = std::move(h.plans[lru]);
      h.plans.erase(h.plans.begin() + lru);
      if (plan->used) EARB_CUDA(cudaStreamWaitEvent(stream, plan->used, 0));
    } else {
      plan = std::make_unique<earb200_interp::Plan>();
    }
    std::vector<RenderItem> all;
    for (int k = 0; k < 4; k++) {
      plan->n_items[k] = (long)items[k].size();
      all.insert(all.end(), items[k].begin(), items[k].end());
    }
    plan->d_items.reserve(std::max<size_t>(all.size(), 1));
    plan->d_segs.reserve(std::max<size_t>(segs.size(), 1));
    // (pageable source: the runtime has staged the bytes when the call returns, so the vectors may
    // go out of scope)
    EARB_CUDA(cudaMemcpyAsync(plan->d_items.ptr, all.data(), sizeof(RenderItem) * all.size(),
                              cudaMemcpyHostToDevice, stream));
    EARB_CUDA(cudaMemcpyAsync(plan->d_segs.ptr, segs.data(), sizeof(RenderSegment) * segs.size(),
                              cudaMemcpyHostToDevice, stream));
    plan->start = block_start;
    plan->n = nsamples;
    plan->align = align;
    plan->last_use = ++h.plan_clock;
    h.plans.push_back(std::move(plan));
    return *h.plans.back();
  }

  void process_device(earb200_interp &h, int64_t block_start, int64_t nsamples, const float *d_in,
                      int64_t in_stride, float *d_out, int64_t out_stride, cudaStream_t stream,
                      earb200_bus *bus = nullptr, int64_t bus_t0 = 0) {
    if (nsamples < 0) throw Error(INVALID_ARGUMENT, "negative sample count");
    if (nsamples == 0) return;
    if (!d_in || (!d_out && !bus)) throw Error(INVALID_ARGUMENT, "null buffer");
    use_device(h.device);
    if (bus) {
      if (earb200_bus_num_channels(bus) != h.M) throw Error(INVALID_ARGUMENT, "the bus has another channel count");
      if (bus_t0 < 0 || bus_t0 + nsamples > earb200_bus_capacity(bus))
        throw Error(INVALID_ARGUMENT, "the call reaches outside the bus");
      if (!d_out) {  // no partial bus: the output imposes no alignment, take the input's
        out_stride = 4;
        d_out = reinterpret_cast<float *>(reinterpret_cast<uintptr_t>(d_in) & 15);
      }
    }
    // the TMA path needs 16-byte aligned rows in both buffers at a common sample phase
    int align = -1;
    const uintptr_t ia = reinterpret_cast<uintptr_t>(d_in), oa = reinterpret_cast<uintptr_t>(d_out);
    if (ia % 4 == 0 && oa % 4 == 0 && in_stride % 4 == 0 && out_stride % 4 == 0 &&
        ((ia / 4) & 3) == ((oa / 4) & 3))
      align = (int)((ia / 4) & 3);
    earb200_interp::Plan &pl = build_plan(h, block_start, nsamples, align, stream);
    RenderPlan plan{pl.d_items.ptr, pl.n_items[0], pl.n_items[1], pl.n_items[2], pl.n_items[3], h.use_tc,
                    h.n_sms, (long)nsamples, h.d_images.ptr, h.accumulate, (long)h.n_points, RenderBus{}};
    // Bus mode.  Fused: the epilogue of the tcgen05 kernel stores its tiles to the bus slots itself
    // (accumulate mode and the earlier input groups of N > 128 read / write the caller's partial
    // bus); the ragged runs between them (segment boundaries that are not tile-aligned: the FFMA
    // kernels) are rendered into the partial bus and only THEY are pushed.  Without a tcgen05 path or
    // without any tile: render into the partial bus, then push all of it.
    const bool have_partial = reinterpret_cast<uintptr_t>(d_out) > 15;
    const long n_slow = pl.n_items[2] + pl.n_items[3], n_fast = pl.n_items[0] + pl.n_items[1];
    const bool fused = bus && h.use_tc && n_fast > 0 &&
                       (have_partial || (n_slow == 0 && !h.accumulate && h.N <= 128));
    if (bus && !fused && !have_partial)
      throw Error(INVALID_ARGUMENT, "this call needs a partial-bus buffer (accumulate mode, more than 128 "
                                    "input channels, or a call that is not made of whole aligned tiles)");
    if (fused) plan.bus = earb200_bus_map(*bus, bus_t0);
    h.timer.start(stream);
    launch_render(h.N, h.M, d_in, (long)in_stride, d_out, (long)out_stride, plan, pl.d_segs.ptr,
                  h.d_gains.ptr, stream);
    if (bus && !fused) {
      launch_bus_push(d_out, out_stride, h.M, bus_t0, nsamples, earb200_bus_map(*bus, bus_t0), stream);
      earb200_bus_count_launch(bus);
    } else if (bus && n_slow > 0) {
      launch_bus_push_items(pl.d_items.ptr + n_fast, n_slow, d_out, out_stride, h.M, earb200_bus_map(*bus, bus_t0),
                            stream);
      earb200_bus_count_launch(bus);
    }
    h.timer.stop(stream);
    EARB_CUDA(cudaGetLastError());
    if (!pl.used) EARB_CUDA(cudaEventCreateWithFlags(&pl.used, cudaEventDisableTiming));
    EARB_CUDA(cudaEventRecord(pl.used, stream));
    if (h.use_tc) {
      if (pl.n_items[0] + pl.n_items[1] > 0) h.launches += (h.N + 255) / 256;  // one per group of 256 inputs
      h.launches += (pl.n_items[2] > 0) + (pl.n_items[3] > 0);
    } else {
      for (int k = 0; k < 4; k++) h.launches += pl.n_items[k] > 0;
    }
  }
}  // namespace

// ---- two interpolators over the same input in one launch (capi_flow.cpp: the direct and the
// diffuse render of the Objects signal flow, ref docs/dsp.rst:40-71)
//
// A segment between two points with equal gain matrices is "constant" (constant_interp, ref
// gain_interpolator.hpp:139-141); rendered as a ramp it gives the same samples, because the ramp's
// slope term is then exactly zero.  Two interpolators with the same sample indices therefore share
// one segmentation once a segment counts as constant only where it is for both.
void earb_interp_share_segmentation(earb200_interp &a, earb200_interp &b) {
  if (a.sample_idx != b.sample_idx) return;
  for (size_t i = 0; i < a.same_as_prev.size(); i++)
    a.same_as_prev[i] = b.same_as_prev[i] = (char)(a.same_as_prev[i] && b.same_as_prev[i]);
  a.plans.clear();
  b.plans.clear();
}

// process() of both: the aligned tiles of both on ONE launch of the tcgen05 kernel (RenderPair: the
// second read of every input tile is served by L2), the ragged runs per interpolator.  Falls back
// to two calls when the two do not share shape, segmentation or the tcgen05 path.
void earb_interp_process_pair_device(earb200_interp &a, earb200_interp &b, int64_t block_start, int64_t nsamples,
                                     const float *d_in, int64_t in_stride, float *d_out_a, float *d_out_b,
                                     int64_t out_stride, cudaStream_t stream) {
  const uintptr_t ia = reinterpret_cast<uintptr_t>(d_in), oa = reinterpret_cast<uintptr_t>(d_out_a),
                  ob = reinterpret_cast<uintptr_t>(d_out_b);
  const bool aligned = ia % 4 == 0 && oa % 4 == 0 && ob % 4 == 0 && in_stride % 4 == 0 && out_stride % 4 == 0 &&
                       ((ia / 4) & 3) == ((oa / 4) & 3) && ((ia / 4) & 3) == ((ob / 4) & 3);
  const bool pairable = a.use_tc && b.use_tc && a.N == b.N && a.M == b.M && a.device == b.device && !a.accumulate &&
                        !b.accumulate && a.n_points == b.n_points && a.n_points > 0 && a.sorted && aligned &&
                        nsamples > 0 && d_in && d_out_a && d_out_b && a.sample_idx == b.sample_idx &&
                        a.same_as_prev == b.same_as_prev;
  if (!pairable) {
    process_device(a, block_start, nsamples, d_in, in_stride, d_out_a, out_stride, stream);
    process_device(b, block_start, nsamples, d_in, in_stride, d_out_b, out_stride, stream);
    return;
  }
  use_device(a.device);
  const int align = (int)((ia / 4) & 3);
  earb200_interp::Plan &pl = build_plan(a, block_start, nsamples, align, stream);
  const long n_fast = pl.n_items[0] + pl.n_items[1];
  a.timer.start(stream);
  if (n_fast > 0) {
    RenderPair pair;
    pair.images2 = b.d_images.ptr;
    pair.out2 = d_out_b;
    if (!launch_render_tc(a.N, a.M, d_in, (long)in_stride, (long)nsamples, d_out_a, (long)out_stride, pl.d_items.ptr,
                          n_fast, pl.d_segs.ptr, a.d_images.ptr, (long)a.n_points, a.n_sms, 0, RenderBus{}, stream,
                          pair))
      throw Error(INTERNAL_ERROR, "cuTensorMapEncodeTiled failed for the render input");
    a.launches += (a.N + 255) / 256;
  }
  if (pl.n_items[2] + pl.n_items[3] > 0) {
    // the ragged runs: the FFMA kernels, per interpolator, on the same item list
    const RenderPlan slow{pl.d_items.ptr + n_fast, 0, 0, pl.n_items[2], pl.n_items[3], 0, a.n_sms, (long)nsamples,
                          nullptr, 0, (long)a.n_points, RenderBus{}};
    launch_render(a.N, a.M, d_in, (long)in_stride, d_out_a, (long)out_stride, slow, pl.d_segs.ptr, a.d_gains.ptr,
                  stream);
    launch_render(b.N, b.M, d_in, (long)in_stride, d_out_b, (long)out_stride, slow, pl.d_segs.ptr, b.d_gains.ptr,
                  stream);
    a.launches += (pl.n_items[2] > 0) + (pl.n_items[3] > 0);
    b.launches += (pl.n_items[2] > 0) + (pl.n_items[3] > 0);
  }
  a.timer.stop(stream);
  EARB_CUDA(cudaGetLastError());
  if (!pl.used) EARB_CUDA(cudaEventCreateWithFlags(&pl.used, cudaEventDisableTiming));
  EARB_CUDA(cudaEventRecord(pl.used, stream));
}

extern "C" {

int earb200_interp_set_accumulate(earb200_interp *interp, int enable) {
  return guarded([&] {
    if (!interp) throw Error(INVALID_ARGUMENT, "null argument");
    interp->accumulate = enable ? 1 : 0;
  });
}

int earb200_interp_create(int n_in, int n_out, int device, earb200_interp **out) {
  return guarded([&] {
    if (!out) throw Error(INVALID_ARGUMENT, "null argument");
    if (n_in < 1 || n_out < 1 || n_out > EARB_MAX_OUT)
      throw Error(INVALID_ARGUMENT, "unsupported channel counts");
    if (render_smem_bytes(n_in, n_out) > 200 * 1024)
      throw Error(INVALID_ARGUMENT, "too many input channels for one gain interpolator");
    use_device(device);
    auto h = std::make_unique<earb200_interp>();
    h->device = device;
    cudaDeviceProp prop;
    EARB_CUDA(cudaGetDeviceProperties(&prop, device));
    h->n_sms = prop.multiProcessorCount;
    // EARB200_RENDER_SM_RESERVE=k leaves k SMs free of the persistent render CTAs (each takes a
    // whole SM), e.g. for a collective running next to the render on another stream
    if (const char *r = std::getenv("EARB200_RENDER_SM_RESERVE")) {
      const int k = std::atoi(r);
      if (k > 0 && k < h->n_sms) h->n_sms -= k;
    }
    h->N = n_in;
    h->M = n_out;
    const char *path = std::getenv("EARB200_RENDER_PATH");
    h->use_tc = render_tc_supported(n_in, n_out) && !(path && std::string(path) == "cuda");
    *out = h.release();
  });
}
void earb200_interp_destroy(earb200_interp *interp) { delete interp; }
int64_t earb200_interp_launch_count(const earb200_interp *interp) {
  return interp ? interp->launches : 0;
}

int earb200_interp_set_profiling(earb200_interp *interp, int enable) {
  return guarded([&] {
    if (!interp) throw Error(INVALID_ARGUMENT, "null argument");
    interp->timer.enabled = enable != 0;
  });
}
int earb200_interp_kernel_time(earb200_interp *interp, double *ms, int64_t *launches) {
  return guarded([&] {
    if (!interp || !ms) throw Error(INVALID_ARGUMENT, "null argument");
    use_device(interp->device);
    *ms = interp->timer.collect(launches);
  });
}

int earb200_interp_set_points(earb200_interp *interp, int64_t n_points,
                              const int64_t *sample_idx, const float *gains) {
  return guarded([&] {
    if (!interp || n_points < 0 || (n_points > 0 && (!sample_idx || !gains)))
      throw Error(INVALID_ARGUMENT, "null argument");
    earb200_interp &h = *interp;
    use_device(h.device);
    const size_t mat = (size_t)h.N * h.M;
    h.sample_idx.assign(sample_idx, sample_idx + n_points);
    h.same_as_prev.assign(n_points, 0);
    h.sorted = std::is_sorted(h.sample_idx.begin(), h.sample_idx.end());
    for (int64_t i = 1; i < n_points; i++) {
      const float *a = gains + (size_t)(i - 1) * mat, *b = gains + (size_t)i * mat;
      bool same = true;
      for (size_t k = 0; k < mat && same; k++) same = (a[k] == b[k]);
      h.same_as_prev[i] = same;
    }
    h.n_points = n_points;
    h.plans.clear();
    if (n_points > 0) {
      h.d_gains.reserve(mat * n_points);
      EARB_CUDA(cudaMemcpy(h.d_gains.ptr, gains, sizeof(float) * mat * n_points,
                           cudaMemcpyHostToDevice));
      if (h.use_tc) {
        h.d_images.reserve(render_tc_image_floats(h.N, h.M) * n_points);
        launch_render_prepare_points(h.N, h.M, (long)n_points, h.d_gains.ptr, h.d_images.ptr, 0);
        EARB_CUDA(cudaGetLastError());
        EARB_CUDA(cudaStreamSynchronize(0));
        h.launches++;
      }
    }
  });
}

// interp_points whose gain matrices are already on the device (the output of the gain
// calculators' *_device entry points): no host round trip.  Synchronises `stream`.
int earb200_interp_set_points_device(earb200_interp *interp, int64_t n_points, const int64_t *sample_idx,
                                     const float *d_gains, void *stream) {
  return guarded([&] {
    if (!interp || n_points < 0 || (n_points > 0 && (!sample_idx || !d_gains)))
      throw Error(INVALID_ARGUMENT, "null argument");
    earb200_interp &h = *interp;
    use_device(h.device);
    cudaStream_t st = (cudaStream_t)stream;
    const size_t mat = (size_t)h.N * h.M;
    h.sample_idx.assign(sample_idx, sample_idx + n_points);
    h.same_as_prev.assign(n_points, 0);
    h.sorted = std::is_sorted(h.sample_idx.begin(), h.sample_idx.end());
    h.n_points = n_points;
    h.plans.clear();
    if (n_points == 0) return;
    h.d_gains.reserve(mat * n_points);
    EARB_CUDA(cudaMemcpyAsync(h.d_gains.ptr, d_gains, sizeof(float) * mat * n_points, cudaMemcpyDeviceToDevice, st));
    if (n_points > 1) {
      char *d_flags = h.d_same.reserve((size_t)n_points);
      launch_render_points_same((long)mat, (long)n_points, h.d_gains.ptr, d_flags, st);
      EARB_CUDA(cudaMemcpyAsync(h.same_as_prev.data() + 1, d_flags + 1, (size_t)(n_points - 1), cudaMemcpyDeviceToHost,
                                st));
      h.launches++;
    }
    if (h.use_tc) {
      h.d_images.reserve(render_tc_image_floats(h.N, h.M) * n_points);
      launch_render_prepare_points(h.N, h.M, (long)n_points, h.d_gains.ptr, h.d_images.ptr, st);
      h.launches++;
    }
    EARB_CUDA(cudaGetLastError());
    EARB_CUDA(cudaStreamSynchronize(st));
  });
}

int earb200_interp_process_device(earb200_interp *interp, int64_t block_start,
                                  int64_t nsamples, const float *d_in, int64_t in_stride,
                                  float *d_out, int64_t out_stride, void *stream) {
  return guarded([&] {
    if (!interp) throw Error(INVALID_ARGUMENT, "null argument");
    process_device(*interp, block_start, nsamples, d_in, in_stride, d_out, out_stride,
                   (cudaStream_t)stream);
  });
}

// process() whose output goes to the node-wide bus (earb200_bus_*): the partial result of this
// rank's objects for the bus samples [bus_t0, bus_t0 + nsamples)
int earb200_interp_process_device_bus(earb200_interp *interp, int64_t block_start, int64_t nsamples,
                                      const float *d_in, int64_t in_stride, float *d_partial,
                                      int64_t partial_stride, earb200_bus *bus, int64_t bus_t0, void *stream) {
  return guarded([&] {
    if (!interp || !bus) throw Error(INVALID_ARGUMENT, "null argument");
    process_device(*interp, block_start, nsamples, d_in, in_stride, d_partial, partial_stride, (cudaStream_t)stream,
                   bus, bus_t0);
  });
}

int earb200_interp_process_host(earb200_interp *interp, int64_t block_start,
                                int64_t nsamples, const float *in, int64_t in_stride,
                                float *out, int64_t out_stride) {
  return guarded([&] {
    if (!interp) throw Error(INVALID_ARGUMENT, "null argument");
    if (nsamples < 0) throw Error(INVALID_ARGUMENT, "negative sample count");
    if (nsamples == 0) return;
    if (!in || !out) throw Error(INVALID_ARGUMENT, "null buffer");
    earb200_interp &h = *interp;
    use_device(h.device);
    // device copies are packed with a 16-byte aligned channel stride
    const int64_t stride = (nsamples + 3) / 4 * 4;
    float *d_in = h.d_in.reserve((size_t)stride * h.N);
    float *d_out = h.d_out.reserve((size_t)stride * h.M);
    EARB_CUDA(cudaMemcpy2DAsync(d_in, sizeof(float) * stride, in, sizeof(float) * in_stride,
                                sizeof(float) * nsamples, h.N, cudaMemcpyHostToDevice, 0));
    if (h.accumulate)  // the bus being mixed into travels too
      EARB_CUDA(cudaMemcpy2DAsync(d_out, sizeof(float) * stride, out, sizeof(float) * out_stride,
                                  sizeof(float) * nsamples, h.M, cudaMemcpyHostToDevice, 0));
    process_device(h, block_start, nsamples, d_in, stride, d_out, stride, 0);
    EARB_CUDA(cudaMemcpy2DAsync(out, sizeof(float) * out_stride, d_out, sizeof(float) * stride,
                                sizeof(float) * nsamples, h.M, cudaMemcpyDeviceToHost, 0));
    EARB_CUDA(cudaStreamSynchronize(0));
  });
}

}  // extern "C"
