// C ABI: layouts, point-source panner (K1) and Objects gain calculator (K1+K2).
#include <cmath>
#include <cstdio>
#include <cstring>
#include <memory>
#include <vector>

#include "capi_common.hpp"
#include "kernels.hpp"

namespace earb {
  std::string &last_error() {
    static thread_local std::string err;
    return err;
  }
}  // namespace earb

using namespace earb;

// ===================================================================== panner
struct earb200_panner {
  int device = 0;
  PspTables tables;
  PspTables *d_tables = nullptr;
  DeviceBuffer<double> d_xyz, d_gains;
  DeviceBuffer<int32_t> d_region, d_inner;
  ~earb200_panner() {
    if (d_tables) cudaFree(d_tables);
  }
};

namespace {
  void panner_init(earb200_panner &p, const HostLayout &layout, int device) {
    use_device(device);
    p.device = device;
    buildPspTables(layout, p.tables);
    EARB_CUDA(cudaMalloc(&p.d_tables, sizeof(PspTables)));
    EARB_CUDA(cudaMemcpy(p.d_tables, &p.tables, sizeof(PspTables), cudaMemcpyHostToDevice));
  }

  void panner_handle_host(earb200_panner &p, int64_t n, const double *xyz, double *gains,
                          int32_t *region, int32_t *inner) {
    if (n < 0) throw Error(INVALID_ARGUMENT, "negative batch size");
    if (n == 0) return;
    use_device(p.device);
    const int S = p.tables.n_real;
    p.d_xyz.reserve(3 * n);
    p.d_gains.reserve((size_t)S * n);
    if (region) p.d_region.reserve(n);
    if (inner) p.d_inner.reserve(n);
    EARB_CUDA(cudaMemcpy(p.d_xyz.ptr, xyz, sizeof(double) * 3 * n, cudaMemcpyHostToDevice));
    launch_psp_batch(p.d_tables, n, p.d_xyz.ptr, p.d_gains.ptr, region ? p.d_region.ptr : nullptr,
                     inner ? p.d_inner.ptr : nullptr, 0);
    EARB_CUDA(cudaGetLastError());
    EARB_CUDA(cudaMemcpy(gains, p.d_gains.ptr, sizeof(double) * S * n, cudaMemcpyDeviceToHost));
    if (region)
      EARB_CUDA(cudaMemcpy(region, p.d_region.ptr, sizeof(int32_t) * n, cudaMemcpyDeviceToHost));
    if (inner)
      EARB_CUDA(cudaMemcpy(inner, p.d_inner.ptr, sizeof(int32_t) * n, cudaMemcpyDeviceToHost));
  }
}  // namespace

// ===================================================================== objects
struct earb200_objects {
  int device = 0;
  int n_sms = 0;
  HostLayout layout;
  earb200_panner panner;
  int n_out = 0;  // C_out incl. LFE
  int S = 0;
  ExtentGrid grid{};
  DeviceBuffer<int32_t> d_out_index, d_identity_index;
  // tcgen05 extent core: points as float4, gain table as the UMMA B image (TF32 hi | lo)
  DeviceBuffer<float4> d_grid_points;
  DeviceBuffer<float> d_grid_bimage;
  DeviceBuffer<unsigned int> d_n_active;
  // cell-culled tcgen05 extent core: per-cell stream [fp16 gain image | point coordinates],
  // cell centres + class thresholds
  DeviceBuffer<unsigned char> d_cell_stream, d_cell_sums;
  DeviceBuffer<float> d_cell_centres, d_cell_points;
  int n_cells = 0;
  // 0: the first-generation tcgen05 core (all weights evaluated, TF32 x 3): grids with more cells
  // than the cell core's tables hold, and EARB200_EXTENT_PATH=tf32 (cross-check, tested)
  int use_cells = 1;
  // per-chunk scratch
  // Largest chunk in objects (the scratch itself grows on demand).  Measured on 10^6 cfg-2 objects
  // per call: 262,144 -> 451 M vectors/s (4 launch sets), 524,288 -> 465 M (2), 1,048,576 -> 471 M
  // (1): every set of launches has ~30 us of fixed cost and a partly filled last tile per CTA.
  // A call is cut into chunks sized so that every CTA of the extent
  // core gets whole rounds of 512 passes (its time goes with the number of rounds, not with how
  // full the last one is): chunk = rounds * 512 * SMs / (passes per object), see pick_chunk.
  // EARB200_OBJECTS_CHUNK fixes the chunk size instead (tuning).
  static const int64_t kChunk;
  static const int64_t kChunkFixed;
  // passes per object: sampled from the depth column on the host path, fed back from the
  // previous call's pass counts on the device path (2 = every object has depth: never too small)
  double passes_per_object = 2.0;
  static constexpr int kFeedbackSlots = 64;
  unsigned int *h_active = nullptr;  // pinned: n_active of the chunks of the last device call
  int64_t fb_objects[kFeedbackSlots] = {};
  int fb_count = 0;
  bool fb_pending = false;
  cudaEvent_t ev_fb = nullptr;
  // two scratch sets: the FP64 epilogue of chunk k runs on a side stream while the extent core of
  // chunk k + 1 (a different pipe mix: FP32 issue + tensor) already occupies the SMs
  struct Scratch {
    DeviceBuffer<ExtentPass> d_passes;
    DeviceBuffer<ObjectAux> d_aux;
    DeviceBuffer<float> d_results;
    DeviceBuffer<double> d_point_gains;  // point source panner gains of the small-extent objects
    DeviceBuffer<int32_t> d_point_list;  // the objects that need them, compacted by the prelude
    DeviceBuffer<unsigned int> d_n_active;  // [0] active passes, [1] length of d_point_list
    cudaEvent_t ev_core = nullptr, ev_epi = nullptr;
  } scratch[2];
  cudaStream_t s_epi = nullptr;
  int64_t chunk_counter = 0;
  // staging for the *_host entry points
  DeviceBuffer<double> d_in, d_out_d;
  DeviceBuffer<float> d_out_f;
  DeviceBuffer<int32_t> d_region;
  int64_t launches = 0;
  // opt-in extensions beyond libear (EARB200_EXT_*): 0 = the reference's behaviour (throws)
  uint32_t extensions = 0;
  DeviceBuffer<double> d_spk;                       // [S][5]: unit-distance x y z, azimuth, elevation
  DeviceBuffer<double> d_ext_in, d_ext_g, d_ext_f;  // scratch of the extension path
  KernelTimer timer;
  int64_t timed_passes = 0;
  // streams / events of the pipelined *_host entry point
  cudaStream_t s_in = nullptr, s_compute = nullptr, s_out = nullptr;
  std::vector<cudaEvent_t> ev_in, ev_done;
  ~earb200_objects() {
    for (cudaEvent_t e : ev_in) cudaEventDestroy(e);
    for (cudaEvent_t e : ev_done) cudaEventDestroy(e);
    for (Scratch &sc : scratch) {
      if (sc.ev_core) cudaEventDestroy(sc.ev_core);
      if (sc.ev_epi) cudaEventDestroy(sc.ev_epi);
    }
    if (ev_fb) cudaEventDestroy(ev_fb);
    if (h_active) cudaFreeHost(h_active);
    if (s_epi) cudaStreamDestroy(s_epi);
    if (s_in) cudaStreamDestroy(s_in);
    if (s_compute) cudaStreamDestroy(s_compute);
    if (s_out) cudaStreamDestroy(s_out);
  }
};
const int64_t earb200_objects::kChunk = 1048576;
const int64_t earb200_objects::kChunkFixed = [] {
  const char *e = std::getenv("EARB200_OBJECTS_CHUNK");
  const long long v = e ? std::atoll(e) : 0;
  return (int64_t)(v >= 1024 && v <= earb200_objects::kChunk ? v : 0);
}();

namespace {
  void objects_init(earb200_objects &o, const HostLayout &layout, int device, int grid_kind = 0) {
    use_device(device);
    o.device = device;
    o.layout = layout;
    cudaDeviceProp prop;
    EARB_CUDA(cudaGetDeviceProperties(&prop, device));
    o.n_sms = prop.multiProcessorCount;
    panner_init(o.panner, layout, device);
    o.n_out = (int)layout.channels.size();
    o.S = o.panner.tables.n_real;
    if (o.n_out > EARB_MAX_OUT) throw Error(INVALID_ARGUMENT, "too many output channels");

    {  // loudspeaker positions (real, unit distance) for the opt-in channel lock
      std::vector<double> spk;
      for (const auto &ch : layout.channels) {
        if (ch.lfe) continue;
        const Vec3 p = cart(ch.pos.az, ch.pos.el, 1.0);
        spk.insert(spk.end(), {p.x, p.y, p.z, ch.pos.az, ch.pos.el});
      }
      if (!spk.empty())
        EARB_CUDA(cudaMemcpy(o.d_spk.reserve(spk.size()), spk.data(), sizeof(double) * spk.size(),
                             cudaMemcpyHostToDevice));
    }
    // LFE scatter map (ref gain_calculator_objects.cpp:51-52)
    std::vector<int32_t> out_index(EARB_MAX_SPK, 0), identity(EARB_MAX_SPK, 0);
    int s = 0;
    for (int c = 0; c < o.n_out; c++)
      if (!layout.channels[c].lfe) out_index[s++] = c;
    if (s != o.S) throw Error(INTERNAL_ERROR, "loudspeaker count mismatch");
    for (int i = 0; i < EARB_MAX_SPK; i++) identity[i] = i;
    EARB_CUDA(cudaMemcpy(o.d_out_index.reserve(EARB_MAX_SPK), out_index.data(),
                         sizeof(int32_t) * EARB_MAX_SPK, cudaMemcpyHostToDevice));
    EARB_CUDA(cudaMemcpy(o.d_identity_index.reserve(EARB_MAX_SPK), identity.data(),
                         sizeof(int32_t) * EARB_MAX_SPK, cudaMemcpyHostToDevice));

    // extent grid: positions, cached point-source gains (computed by K1 on the device),
    // per-block sums (ref polar_extent.cpp:98-165)
    // grid_kind 0: libear's generatePanningPositionsEven(37), 1,652 points (polar_extent.cpp:12-39,
    // 104) -- the grid all parity rests on.  1: the 5,200-point spherical t-design of
    // resources/Design_5200_100_random.dat (the grid BASELINE.json's north star names for the
    // spreading panner; libear itself only uses it for the HOA decoder): throughput shape only,
    // there is no reference output to compare with.
    if (grid_kind != 0 && grid_kind != 1) throw Error(INVALID_ARGUMENT, "unknown extent grid");
    std::vector<Vec3> pts = grid_kind == 0 ? extentGridPositions(37) : hoaDesignPoints();
    const int P = (int)pts.size();
    // padded to whole B stages of the tcgen05 core (a multiple of the FFMA core's 32-point blocks)
    static_assert(EXTENT_TC_STAGE_POINTS % EARB_EXT_BLOCK == 0, "padding");
    const int Pp = (P + EXTENT_TC_STAGE_POINTS - 1) / EXTENT_TC_STAGE_POINTS * EXTENT_TC_STAGE_POINTS;
    const int NB = Pp / EARB_EXT_BLOCK;
    o.grid.n_points = P;
    o.grid.n_points_pad = Pp;
    o.grid.n_blocks = NB;
    o.grid.n_spk = o.S;

    std::vector<double> xyz(3 * (size_t)P), gains((size_t)P * o.S);
    for (int p = 0; p < P; p++) {
      xyz[3 * p] = pts[p].x;
      xyz[3 * p + 1] = pts[p].y;
      xyz[3 * p + 2] = pts[p].z;
    }
    panner_handle_host(o.panner, P, xyz.data(), gains.data(), nullptr, nullptr);
    o.launches++;

    std::vector<float> fxyz(3 * (size_t)Pp), fg((size_t)o.S * Pp, 0.0f);
    for (int p = 0; p < Pp; p++) {
      const int q = p < P ? p : P - 1;  // padding repeats the last point, with zero gains
      fxyz[p] = (float)pts[q].x;
      fxyz[Pp + p] = (float)pts[q].y;
      fxyz[2 * Pp + p] = (float)pts[q].z;
    }
    for (int p = 0; p < P; p++)
      for (int sp = 0; sp < o.S; sp++) fg[(size_t)sp * Pp + p] = (float)gains[(size_t)p * o.S + sp];

    {
      std::vector<float4> pts4(Pp);
      for (int p = 0; p < Pp; p++) pts4[p] = make_float4(fxyz[p], fxyz[Pp + p], fxyz[2 * Pp + p], 0.0f);
      EARB_CUDA(cudaMemcpy(o.d_grid_points.reserve(pts4.size()), pts4.data(), sizeof(float4) * pts4.size(),
                           cudaMemcpyHostToDevice));
      std::vector<float> image(extent_tc_bimage_floats(o.grid));
      extent_tc_build_bimage(o.grid, fg.data(), image.data());
      EARB_CUDA(cudaMemcpy(o.d_grid_bimage.reserve(image.size()), image.data(), sizeof(float) * image.size(),
                           cudaMemcpyHostToDevice));
      const char *path = std::getenv("EARB200_EXTENT_PATH");
      ExtentCells cells;
      extent_cells_build(P, fxyz.data(), Pp, cells);
      o.n_cells = cells.n_cells;
      o.use_cells = extent_cell_supported(cells) && !(path && std::string(path) == "tf32");
      if (o.use_cells) {
      std::vector<unsigned char> cs(extent_cell_stream_bytes(cells)), sums(extent_cell_sum_bytes(cells));
      extent_cell_build_stream(cells, o.S, Pp, fg.data(), cs.data());
      extent_cell_build_sums(cells, o.S, Pp, fg.data(), sums.data());
      EARB_CUDA(cudaMemcpy(o.d_cell_stream.reserve(cs.size()), cs.data(), cs.size(), cudaMemcpyHostToDevice));
      EARB_CUDA(cudaMemcpy(o.d_cell_sums.reserve(sums.size()), sums.data(), sums.size(), cudaMemcpyHostToDevice));
      std::vector<float> cc(extent_cell_centre_floats(cells)), cpnt(extent_cell_point_floats(cells));
      extent_cell_build_centres(cells, cc.data());
      extent_cell_build_points(cells, Pp, fxyz.data(), cpnt.data());
      EARB_CUDA(cudaMemcpy(o.d_cell_centres.reserve(cc.size()), cc.data(), sizeof(float) * cc.size(),
                           cudaMemcpyHostToDevice));
      EARB_CUDA(cudaMemcpy(o.d_cell_points.reserve(cpnt.size()), cpnt.data(), sizeof(float) * cpnt.size(),
                           cudaMemcpyHostToDevice));
      }
    }
    for (earb200_objects::Scratch &sc : o.scratch) {
      // the per-chunk scratch grows with the chunks a handle actually sees (objects_run_chunk):
      // 1.3 KB per object and set -- a calculator used for a few thousand objects stays small, one
      // fed 10^6 objects at a time holds 1.3 GB
      sc.d_n_active.reserve(2);
      EARB_CUDA(cudaEventCreateWithFlags(&sc.ev_core, cudaEventDisableTiming));
      EARB_CUDA(cudaEventCreateWithFlags(&sc.ev_epi, cudaEventDisableTiming));
    }
    EARB_CUDA(cudaEventCreateWithFlags(&o.ev_fb, cudaEventDisableTiming));
    EARB_CUDA(cudaHostAlloc((void **)&o.h_active, sizeof(unsigned int) * earb200_objects::kFeedbackSlots, cudaHostAllocDefault));
    EARB_CUDA(cudaStreamCreateWithFlags(&o.s_epi, cudaStreamNonBlocking));
  }

  // features libear rejects (ref gain_calculator_objects.cpp:37-43), same order
  void check_flags(int64_t n, const int32_t *flags) {
    if (!flags) return;
    for (int64_t i = 0; i < n; i++) {
      const int32_t f = flags[i];
      if (!f) continue;
      if (f & EARB200_OBJ_CARTESIAN) throw Error(NOT_IMPLEMENTED, "not implemented: cartesian");
      if (f & EARB200_OBJ_CARTESIAN_POSITION)
        throw Error(NOT_IMPLEMENTED, "not implemented: cartesian");
      if (f & EARB200_OBJ_DIVERGENCE) throw Error(NOT_IMPLEMENTED, "not implemented: divergence");
      if (f & EARB200_OBJ_CHANNEL_LOCK)
        throw Error(NOT_IMPLEMENTED, "not implemented: channelLock");
      if (f & EARB200_OBJ_ZONE_EXCLUSION)
        throw Error(NOT_IMPLEMENTED, "not implemented: zoneExclusion");
      if (f & EARB200_OBJ_SCREEN_REF) throw Error(NOT_IMPLEMENTED, "not implemented: screenRef");
    }
  }

  // Objects per chunk for `rounds` whole rounds of 512 passes per CTA of the extent core (0: as
  // many as the scratch holds), with 2.5 % of slack: one pass too many costs a whole round.
  int64_t pick_chunk(const earb200_objects &o, double passes_per_object, int rounds) {
    if (earb200_objects::kChunkFixed) return earb200_objects::kChunkFixed;
    const double ppo = std::min(2.0, std::max(0.05, passes_per_object));
    const double per_round = 0.975 * 512.0 * (double)o.n_sms / ppo;
    if (rounds <= 0) rounds = std::max(1, (int)((double)earb200_objects::kChunk / per_round));
    int64_t chunk = (int64_t)(per_round * rounds);
    chunk = std::min<int64_t>(earb200_objects::kChunk, std::max<int64_t>(8192, chunk / 128 * 128));
    return chunk;
  }

  // one chunk (<= kChunk objects starting at `start`): prelude -> extent core on `stream`, the
  // epilogue on the side stream.  Returns the event that marks the chunk's outputs complete.
  template <typename OutT>
  cudaEvent_t objects_run_chunk(earb200_objects &o, int64_t start, int64_t m, const ObjectsSoA &in, int n_out,
                                const int32_t *d_out_index, OutT *d_direct, OutT *d_diffuse, int32_t *d_region,
                                cudaStream_t stream, int fb_slot = -1, bool epilogue_in_line = false) {
    ObjectsSoA c = in;
    if (c.xyz) {
      c.xyz += 3 * start;
    } else {
      c.azimuth += start;
      c.elevation += start;
      c.distance += start;
    }
    c.width += start;
    c.height += start;
    c.depth += start;
    c.gain += start;
    c.diffuse += start;
    earb200_objects::Scratch &sc = o.scratch[o.chunk_counter++ & 1];
    if ((size_t)m > sc.d_aux.cap) {
      // growing frees the old buffers, which synchronises the device: earlier chunks are done with them
      const size_t cap = std::max<size_t>((size_t)m, 16384);
      sc.d_passes.reserve(2 * cap);
      sc.d_aux.reserve(cap);
      sc.d_results.reserve(2 * cap * EARB_MAX_SPK);
      sc.d_point_gains.reserve(cap * EARB_MAX_SPK);
      sc.d_point_list.reserve(cap);
    }
    // the epilogue of the chunk that used this scratch set two chunks ago must have read it
    EARB_CUDA(cudaStreamWaitEvent(stream, sc.ev_epi, 0));
    EARB_CUDA(cudaMemsetAsync(sc.d_n_active.ptr, 0, 2 * sizeof(unsigned int), stream));
    launch_objects_prelude(m, c, sc.d_passes.ptr, sc.d_aux.ptr, sc.d_n_active.ptr, sc.d_point_list.ptr,
                           d_region != nullptr, stream);
    if (fb_slot >= 0 && fb_slot < earb200_objects::kFeedbackSlots) {
      EARB_CUDA(cudaMemcpyAsync(o.h_active + fb_slot, sc.d_n_active.ptr, sizeof(unsigned int), cudaMemcpyDeviceToHost,
                                stream));
      o.fb_objects[fb_slot] = m;
      o.fb_count = fb_slot + 1;
    }
    o.timer.start(stream);
    if (o.use_cells)
      launch_extent_cell(o.n_cells, o.grid.n_spk, o.d_cell_stream.ptr, o.d_cell_sums.ptr, o.d_cell_centres.ptr,
                         o.d_cell_points.ptr, sc.d_n_active.ptr, sc.d_passes.ptr, sc.d_results.ptr, o.n_sms,
                         stream);
    else
      launch_extent_core_tc(o.grid, o.d_grid_points.ptr, o.d_grid_bimage.ptr, sc.d_n_active.ptr, sc.d_passes.ptr,
                            sc.d_results.ptr, o.n_sms, stream);
    o.timer.stop(stream);
    if (o.timer.enabled) o.timed_passes += 2 * m;
    // the epilogue on the side stream (it runs beside the next chunk's prelude); in line, on the
    // core's own stream, was measured slower on the host path as well (e2e 220-225 against 234-236 M
    // vectors/s), although there the copy out of the chunk waits for it
    cudaStream_t s_epi = epilogue_in_line ? stream : o.s_epi;
    EARB_CUDA(cudaEventRecord(sc.ev_core, stream));
    if (!epilogue_in_line) EARB_CUDA(cudaStreamWaitEvent(o.s_epi, sc.ev_core, 0));
    launch_objects_epilogue<OutT>(o.panner.d_tables, o.grid.n_spk, m, c, sc.d_passes.ptr, sc.d_aux.ptr,
                                  sc.d_results.ptr, sc.d_point_gains.ptr, sc.d_n_active.ptr, sc.d_point_list.ptr,
                                  o.n_sms, n_out, d_out_index,
                                  d_direct + start * n_out, d_diffuse + start * n_out,
                                  d_region ? d_region + start : nullptr, s_epi);
    EARB_CUDA(cudaEventRecord(sc.ev_epi, s_epi));
    o.launches += 4;
    return sc.ev_epi;
  }

  // joins the side stream: everything launched so far is ordered before later work on `stream`
  void objects_join(earb200_objects &o, cudaStream_t stream) {
    for (earb200_objects::Scratch &sc : o.scratch) EARB_CUDA(cudaStreamWaitEvent(stream, sc.ev_epi, 0));
  }

  template <typename OutT>
  void objects_run(earb200_objects &o, int64_t n, const ObjectsSoA &in, int n_out,
                   const int32_t *d_out_index, OutT *d_direct, OutT *d_diffuse,
                   int32_t *d_region, cudaStream_t stream) {
    // passes per object as the previous call measured them (if its counters have arrived)
    if (o.fb_pending && cudaEventQuery(o.ev_fb) == cudaSuccess) {
      double active = 0.0, objects = 0.0;
      for (int k = 0; k < o.fb_count; k++) {
        active += o.h_active[k];
        objects += (double)o.fb_objects[k];
      }
      if (objects > 0) o.passes_per_object = active / objects;
      o.fb_pending = false;
    }
    int64_t chunk = pick_chunk(o, o.passes_per_object, 0);
    if (!earb200_objects::kChunkFixed && n > chunk) {
      // a short last chunk costs a whole set of launches: if the batch fits one chunk fewer without
      // another round of 512 passes per CTA (the 2.5 % of slack in pick_chunk usually allows it),
      // spread it evenly (10^6 objects: 4 x 250,112 instead of 4 x 246,272 + 14,912)
      const int64_t n_chunks = (n + chunk - 1) / chunk, rest = n - (n_chunks - 1) * chunk;
      if (rest < chunk / 4) {
        const int64_t even = ((n + n_chunks - 2) / (n_chunks - 1) + 127) / 128 * 128;
        const double ppo = std::min(2.0, std::max(0.05, o.passes_per_object));
        const double rounds = std::ceil((double)chunk * ppo / (512.0 * o.n_sms));
        if (even <= earb200_objects::kChunk && (double)even * ppo <= rounds * 512.0 * o.n_sms) chunk = even;
      }
    }
    const bool feed = !o.fb_pending && n > chunk;  // a call of one chunk has nothing to size
    if (feed) o.fb_count = 0;
    int slot = 0;
    for (int64_t start = 0; start < n; start += chunk, slot++)
      objects_run_chunk<OutT>(o, start, std::min<int64_t>(chunk, n - start), in, n_out, d_out_index, d_direct,
                              d_diffuse, d_region, stream, feed ? slot : -1);
    if (feed && o.fb_count > 0) {
      EARB_CUDA(cudaEventRecord(o.ev_fb, stream));
      o.fb_pending = true;
    }
    objects_join(o, stream);
    EARB_CUDA(cudaGetLastError());
  }

  // OPT-IN divergence / channel lock (objects_ext_kernels.cu; parity unpinned: libear throws).
  // Every object becomes three sources that go through the unchanged pipeline with unit gain; the
  // results are combined in the power domain.  Chunks of 65,536 objects bound the scratch.
  template <typename OutT>
  void objects_run_ext(earb200_objects &o, int64_t n, const ObjectsSoA &in, const double *d_div,
                       const double *d_range, const double *d_lock, OutT *d_direct, OutT *d_diffuse,
                       cudaStream_t stream) {
    const int64_t chunk = 65536;
    double *sc = o.d_ext_in.reserve((size_t)chunk * (9 + 3 * 5 + 3));
    double *xyz3 = sc, *w3 = xyz3 + 9 * chunk, *h3 = w3 + 3 * chunk, *d3 = h3 + 3 * chunk, *ones = d3 + 3 * chunk,
           *zeros = ones + 3 * chunk, *weights = zeros + 3 * chunk;
    double *g3 = o.d_ext_g.reserve((size_t)3 * chunk * o.n_out), *f3 = o.d_ext_f.reserve((size_t)3 * chunk * o.n_out);
    for (int64_t start = 0; start < n; start += chunk) {
      const int64_t m = std::min(chunk, n - start);
      ObjectsSoA c{in.azimuth + start, in.elevation + start, in.distance + start, in.width + start,
                   in.height + start, in.depth + start,     in.gain + start,     in.diffuse + start, nullptr};
      launch_objects_ext_expand(m, c, d_div ? d_div + start : nullptr, d_range ? d_range + start : nullptr,
                                d_lock ? d_lock + start : nullptr, o.S, o.d_spk.ptr, xyz3, w3, h3, d3, ones, zeros,
                                weights, stream);
      ObjectsSoA three{nullptr, nullptr, nullptr, w3, h3, d3, ones, zeros, xyz3};
      objects_run<double>(o, 3 * m, three, o.n_out, o.d_out_index.ptr, g3, f3, nullptr, stream);
      launch_objects_ext_combine<OutT>(m, o.n_out, g3, weights, c.gain, c.diffuse, d_direct + start * o.n_out,
                                       d_diffuse + start * o.n_out, stream);
      o.launches += 2;
    }
    EARB_CUDA(cudaGetLastError());
  }

  // Host buffers in, host buffers out, software-pipelined per chunk over three streams: while chunk
  // k is computed, chunk k + 1 is on its way in and chunk k - 1 on its way out (the copies overlap
  // the kernels when the caller's buffers are pinned; with pageable memory they still work, staged).
  template <typename OutT>
  void objects_host_pipeline(earb200_objects &o, int64_t n, const double *const src[8], OutT *direct,
                             OutT *diffuse, int32_t *region_idx) {
    if (!o.s_in) {
      EARB_CUDA(cudaStreamCreateWithFlags(&o.s_in, cudaStreamNonBlocking));
      EARB_CUDA(cudaStreamCreateWithFlags(&o.s_compute, cudaStreamNonBlocking));
      EARB_CUDA(cudaStreamCreateWithFlags(&o.s_out, cudaStreamNonBlocking));
    }
    // passes per object from a strided sample of the depth column (depth != 0: two passes)
    {
      const int64_t samples = std::min<int64_t>(n, 4096), stride = std::max<int64_t>(1, n / samples);
      int64_t deep = 0;
      for (int64_t k = 0; k < samples; k++) deep += src[5][k * stride] != 0.0;
      o.passes_per_object = 1.0 + (double)deep / (double)samples;
    }
    // Two rounds per chunk: small enough for the copies of neighbouring chunks to hide behind it.
    // The copies are the critical path (10^6 objects: 3.65 ms of PCIe traffic against 2.4 ms of
    // kernels), so what the schedule can still save is the fill (first chunk in + computed before
    // the copy out starts) and the drain (last chunk out after everything else has finished): the
    // first and the last two chunks are a quarter and a half of the others.
    const int64_t chunk = pick_chunk(o, o.passes_per_object, 2);
    std::vector<int64_t> sizes;
    {
      const bool ramp = !earb200_objects::kChunkFixed && n >= 5 * chunk;
      const int64_t q = chunk / 4 / 128 * 128, h = chunk / 2 / 128 * 128;
      int64_t left = n;
      auto take = [&](int64_t m) {
        m = std::min(m, left);
        if (m > 0) sizes.push_back(m);
        left -= m;
      };
      if (ramp) {
        take(q);
        take(h);
        while (left > chunk + h + q) take(chunk);
        // what remains (between h + q and chunk + h + q): its last quarter-chunk and half-chunk close
        // the schedule, the rest goes before them
        const int64_t rest = left - h - q;
        take(rest);
        take(h);
        take(q);
      } else {
        while (left > 0) take(chunk);
      }
    }
    const int64_t n_chunks = (int64_t)sizes.size();
    while ((int64_t)o.ev_in.size() < n_chunks) {
      cudaEvent_t a, b;
      EARB_CUDA(cudaEventCreateWithFlags(&a, cudaEventDisableTiming));
      EARB_CUDA(cudaEventCreateWithFlags(&b, cudaEventDisableTiming));
      o.ev_in.push_back(a);
      o.ev_done.push_back(b);
    }
    double *d = o.d_in.reserve(8 * (size_t)n);
    ObjectsSoA soa{d, d + n, d + 2 * n, d + 3 * n, d + 4 * n, d + 5 * n, d + 6 * n, d + 7 * n, nullptr};
    int32_t *d_reg = region_idx ? o.d_region.reserve(n) : nullptr;
    const size_t n_el = (size_t)n * o.n_out;
    OutT *d_out = reinterpret_cast<OutT *>(sizeof(OutT) == 8 ? (void *)o.d_out_d.reserve(2 * n_el)
                                                              : (void *)o.d_out_f.reserve(2 * n_el));
    OutT *d_direct = d_out, *d_diffuse = d_out + n_el;
    // EARB200_HOST_TRACE=1: per-chunk timeline of the three streams on stderr (diagnostics)
    static const bool trace = [] {
      const char *e = std::getenv("EARB200_HOST_TRACE");
      return e && *e == '1';
    }();
    std::vector<cudaEvent_t> tr;
    auto mark = [&](cudaStream_t st) {
      if (!trace) return;
      cudaEvent_t e;
      cudaEventCreate(&e);
      cudaEventRecord(e, st);
      tr.push_back(e);
    };
    mark(o.s_in);
    // All copies in are queued ahead of the chunk loop, one piece per chunk.  An event trace
    // (EARB200_HOST_TRACE=1) shows the copy out -- 192 B per object against 64 B in -- as the long
    // pole: 47 GB/s while copies in run beside it, 54 GB/s afterwards, and the copies in crawling
    // at 14 GB/s next to it.  Larger pieces (EARB200_HOST_PIECES=g: g chunks per copy) bring the
    // input in sooner but make the chunks at the head of a piece wait: measured on one box
    // g = 1 / 2 / 3 / 4: 233 / 230 / 226 / 221 M vectors/s, so g = 1 stays.
    // piece_of[k]: the piece whose event chunk k waits for.
    std::vector<int> piece_of(n_chunks, 0);
    {
      std::vector<int64_t> piece_end;  // chunk index one past the piece
      // the first two chunks one by one (the fill must not wait for a large copy), then groups of
      // kGroup chunks
      static const int64_t kGroup = [] {
        const char *e = std::getenv("EARB200_HOST_PIECES");
        const long v = e ? std::atol(e) : 1;
        return (int64_t)(v >= 1 && v <= 64 ? v : 1);
      }();
      const int64_t single = std::min<int64_t>(n_chunks, 2);
      for (int64_t k = 0; k < single; k++) piece_end.push_back(k + 1);
      for (int64_t k = single; k < n_chunks; k += kGroup) piece_end.push_back(std::min(n_chunks, k + kGroup));
      int64_t k0 = 0, obj0 = 0;
      for (size_t pc = 0; pc < piece_end.size(); pc++) {
        int64_t m = 0;
        for (int64_t k = k0; k < piece_end[pc]; k++) m += sizes[k], piece_of[k] = (int)pc;
        for (int f = 0; f < 8; f++)
          EARB_CUDA(cudaMemcpyAsync(d + (size_t)f * n + obj0, src[f] + obj0, sizeof(double) * m,
                                    cudaMemcpyHostToDevice, o.s_in));
        EARB_CUDA(cudaEventRecord(o.ev_in[pc], o.s_in));
        k0 = piece_end[pc];
        obj0 += m;
      }
    }
    int64_t start = 0;
    for (int64_t k = 0; k < n_chunks; start += sizes[k], k++) {
      const int64_t m = sizes[k];
      mark(o.s_in);  // [1 + 5k] (all copies in are queued ahead)
      EARB_CUDA(cudaStreamWaitEvent(o.s_compute, o.ev_in[piece_of[k]], 0));
      cudaEvent_t done = objects_run_chunk<OutT>(o, start, m, soa, o.n_out, o.d_out_index.ptr, d_direct,
                                                 d_diffuse, d_reg, o.s_compute);
      mark(o.s_compute);  // [2 + 5k] core done
      mark(o.s_epi);      // [3 + 5k] epilogue done
      EARB_CUDA(cudaStreamWaitEvent(o.s_out, done, 0));
      mark(o.s_out);      // [4 + 5k] copy out may start
      const size_t off = (size_t)start * o.n_out, cnt = (size_t)m * o.n_out;
      EARB_CUDA(cudaMemcpyAsync(direct + off, d_direct + off, sizeof(OutT) * cnt, cudaMemcpyDeviceToHost, o.s_out));
      EARB_CUDA(cudaMemcpyAsync(diffuse + off, d_diffuse + off, sizeof(OutT) * cnt, cudaMemcpyDeviceToHost, o.s_out));
      if (region_idx)
        EARB_CUDA(cudaMemcpyAsync(region_idx + start, d_reg + start, sizeof(int32_t) * m, cudaMemcpyDeviceToHost,
                                  o.s_out));
      mark(o.s_out);      // [5 + 5k] copy out done
    }
    EARB_CUDA(cudaGetLastError());
    EARB_CUDA(cudaStreamSynchronize(o.s_out));
    EARB_CUDA(cudaStreamSynchronize(o.s_compute));
    EARB_CUDA(cudaStreamSynchronize(o.s_epi));
    if (trace) {
      auto at = [&](size_t i) {
        float ms = 0.0f;
        cudaEventElapsedTime(&ms, tr[0], tr[i]);
        return ms;
      };
      for (int64_t k = 0; k < n_chunks; k++)
        fprintf(stderr, "host-trace chunk %2lld (%7lld obj): in %.3f  core %.3f  epi %.3f  out %.3f -> %.3f ms\n",
                (long long)k, (long long)sizes[k], at(1 + 5 * k), at(2 + 5 * k), at(3 + 5 * k), at(4 + 5 * k),
                at(5 + 5 * k));
      for (cudaEvent_t e : tr) cudaEventDestroy(e);
    }
  }
}  // namespace

extern "C" {

const char *earb200_last_error(void) { return last_error().c_str(); }
const char *earb200_version(void) { return "libear-b200 0.1.0 (libear 0.9.0 API)"; }

// ----------------------------------------------------------------- layouts
int earb200_num_layouts(void) { return (int)layoutNames().size(); }
const char *earb200_layout_name_at(int i) {
  const auto &n = layoutNames();
  return (i >= 0 && i < (int)n.size()) ? n[i].c_str() : nullptr;
}
int earb200_layout_create(const char *name, earb200_layout **out) {
  return guarded([&] {
    if (!name || !out) throw Error(INVALID_ARGUMENT, "null argument");
    auto l = std::make_unique<earb200_layout>();
    l->layout = getLayout(name);
    *out = l.release();
  });
}
int earb200_layout_create_with_positions(const char *name, int n_channels,
                                         const double *real_pos, earb200_layout **out) {
  return guarded([&] {
    if (!name || !out || !real_pos) throw Error(INVALID_ARGUMENT, "null argument");
    auto l = std::make_unique<earb200_layout>();
    l->layout = getLayout(name);
    if ((int)l->layout.channels.size() != n_channels)
      throw Error(INVALID_ARGUMENT, "number of positions does not match the layout");
    for (int i = 0; i < n_channels; i++)
      l->layout.channels[i].pos = {real_pos[3 * i], real_pos[3 * i + 1], real_pos[3 * i + 2]};
    *out = l.release();
  });
}
int earb200_layout_create_custom(const char *name, int n_channels, const char *const *channel_names,
                                 const double *pos, const double *nominal_pos, const double *az_range,
                                 const double *el_range, const int32_t *is_lfe, earb200_layout **out) {
  return guarded([&] {
    if (!name || !out || !pos || !channel_names) throw Error(INVALID_ARGUMENT, "null argument");
    if (n_channels < 0) throw Error(INVALID_ARGUMENT, "negative channel count");
    auto l = std::make_unique<earb200_layout>();
    l->layout.name = name;
    for (int i = 0; i < n_channels; i++) {
      if (!channel_names[i]) throw Error(INVALID_ARGUMENT, "null channel name");
      HostChannel c;
      c.name = channel_names[i];
      c.pos = {pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]};
      // defaults as ear::Channel (ref src/layout.cpp:27-45): nominal = real, ranges = the position
      c.nominal = nominal_pos ? Polar{nominal_pos[3 * i], nominal_pos[3 * i + 1], nominal_pos[3 * i + 2]} : c.pos;
      c.az_lo = az_range ? az_range[2 * i] : c.pos.az;
      c.az_hi = az_range ? az_range[2 * i + 1] : c.pos.az;
      c.el_lo = el_range ? el_range[2 * i] : c.pos.el;
      c.el_hi = el_range ? el_range[2 * i + 1] : c.pos.el;
      c.lfe = is_lfe && is_lfe[i];
      l->layout.channels.push_back(c);
    }
    *out = l.release();
  });
}
int earb200_convex_hull(int n_points, const double *xyz, int32_t *facets, int capacity, int *n_written) {
  return guarded([&] {
    if (!xyz || !facets || !n_written) throw Error(INVALID_ARGUMENT, "null argument");
    std::vector<Vec3> pts;
    for (int i = 0; i < n_points; i++) pts.push_back({xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]});
    int k = 0;
    for (const auto &f : convexHull(pts)) {
      if (k + 1 + (int)f.size() > capacity) throw Error(INVALID_ARGUMENT, "facet buffer too small");
      facets[k++] = (int32_t)f.size();
      for (int v : f) facets[k++] = v;
    }
    *n_written = k;
  });
}
int earb200_extent_cells(int n_rows, int capacity, int32_t *order, float *centre, double *radius,
                         double *grid_xyz, int *n_cells_out, int *n_points_out) {
  return guarded([&] {
    if (!n_cells_out || !n_points_out) throw Error(INVALID_ARGUMENT, "null argument");
    if (n_rows < 2) throw Error(INVALID_ARGUMENT, "the grid needs at least two rows");
    std::vector<Vec3> pts = extentGridPositions(n_rows);
    const int P = (int)pts.size();
    const int Pp = (P + EXTENT_TC_STAGE_POINTS - 1) / EXTENT_TC_STAGE_POINTS * EXTENT_TC_STAGE_POINTS;
    std::vector<float> fxyz(3 * (size_t)Pp, 0.0f);
    for (int p = 0; p < P; p++) {
      fxyz[p] = (float)pts[p].x;
      fxyz[Pp + p] = (float)pts[p].y;
      fxyz[2 * Pp + p] = (float)pts[p].z;
    }
    ExtentCells cells;
    extent_cells_build(P, fxyz.data(), Pp, cells);
    *n_cells_out = cells.n_cells;
    *n_points_out = P;
    if (cells.n_cells > capacity) throw Error(INVALID_ARGUMENT, "cell buffers too small");
    for (int c = 0; c < cells.n_cells; c++) {
      if (order)
        for (int i = 0; i < EXTENT_CELL_POINTS; i++)
          order[c * EXTENT_CELL_POINTS + i] = cells.order[(size_t)c * EXTENT_CELL_POINTS + i];
      if (centre)
        for (int i = 0; i < 3; i++) centre[3 * c + i] = cells.centre[3 * (size_t)c + i];
      if (radius) radius[c] = cells.radius[c];
    }
    if (grid_xyz)
      for (int p = 0; p < P; p++) {
        grid_xyz[3 * p] = pts[p].x;
        grid_xyz[3 * p + 1] = pts[p].y;
        grid_xyz[3 * p + 2] = pts[p].z;
      }
  });
}
void earb200_layout_destroy(earb200_layout *layout) { delete layout; }
int earb200_layout_num_channels(const earb200_layout *layout) {
  return layout ? (int)layout->layout.channels.size() : 0;
}
const char *earb200_layout_channel_name(const earb200_layout *layout, int i) {
  if (!layout || i < 0 || i >= (int)layout->layout.channels.size()) return nullptr;
  return layout->layout.channels[i].name.c_str();
}
int earb200_layout_channel_is_lfe(const earb200_layout *layout, int i) {
  if (!layout || i < 0 || i >= (int)layout->layout.channels.size()) return 0;
  return layout->layout.channels[i].lfe ? 1 : 0;
}
int earb200_layout_channel_position(const earb200_layout *layout, int i, double *pos,
                                    double *nominal) {
  return guarded([&] {
    if (!layout || i < 0 || i >= (int)layout->layout.channels.size())
      throw Error(INVALID_ARGUMENT, "channel index out of range");
    const HostChannel &c = layout->layout.channels[i];
    if (pos) {
      pos[0] = c.pos.az;
      pos[1] = c.pos.el;
      pos[2] = c.pos.dist;
    }
    if (nominal) {
      nominal[0] = c.nominal.az;
      nominal[1] = c.nominal.el;
      nominal[2] = c.nominal.dist;
    }
  });
}
int earb200_layout_channel_ranges(const earb200_layout *layout, int i, double *az_range,
                                  double *el_range) {
  return guarded([&] {
    if (!layout || i < 0 || i >= (int)layout->layout.channels.size())
      throw Error(INVALID_ARGUMENT, "channel index out of range");
    const HostChannel &c = layout->layout.channels[i];
    if (az_range) {
      az_range[0] = c.az_lo;
      az_range[1] = c.az_hi;
    }
    if (el_range) {
      el_range[0] = c.el_lo;
      el_range[1] = c.el_hi;
    }
  });
}

// ------------------------------------------------------------------ panner
int earb200_panner_create(const earb200_layout *layout, int device, earb200_panner **out) {
  return guarded([&] {
    if (!layout || !out) throw Error(INVALID_ARGUMENT, "null argument");
    auto p = std::make_unique<earb200_panner>();
    panner_init(*p, layout->layout, device);
    *out = p.release();
  });
}
void earb200_panner_destroy(earb200_panner *panner) { delete panner; }
int earb200_panner_num_channels(const earb200_panner *panner) {
  return panner ? panner->tables.n_real : 0;
}
int earb200_panner_num_regions(const earb200_panner *panner) {
  return panner ? panner->tables.n_regions : 0;
}
int earb200_panner_handle_host(earb200_panner *panner, int64_t n, const double *xyz,
                               double *gains, int32_t *region_idx, int32_t *inner_idx) {
  return guarded([&] {
    if (!panner || (n > 0 && (!xyz || !gains))) throw Error(INVALID_ARGUMENT, "null argument");
    panner_handle_host(*panner, n, xyz, gains, region_idx, inner_idx);
  });
}
int earb200_panner_handle_device(earb200_panner *panner, int64_t n, const double *d_xyz,
                                 double *d_gains, int32_t *d_region_idx,
                                 int32_t *d_inner_idx, void *stream) {
  return guarded([&] {
    if (!panner || (n > 0 && (!d_xyz || !d_gains))) throw Error(INVALID_ARGUMENT, "null argument");
    use_device(panner->device);
    launch_psp_batch(panner->d_tables, n, d_xyz, d_gains, d_region_idx, d_inner_idx,
                     (cudaStream_t)stream);
    EARB_CUDA(cudaGetLastError());
  });
}

// ----------------------------------------------------------------- objects
int earb200_objects_create(const earb200_layout *layout, int device, earb200_objects **out) {
  return guarded([&] {
    if (!layout || !out) throw Error(INVALID_ARGUMENT, "null argument");
    auto o = std::make_unique<earb200_objects>();
    objects_init(*o, layout->layout, device);
    *out = o.release();
  });
}
int earb200_objects_create_with_grid(const earb200_layout *layout, int device, int grid, earb200_objects **out) {
  return guarded([&] {
    if (!layout || !out) throw Error(INVALID_ARGUMENT, "null argument");
    auto o = std::make_unique<earb200_objects>();
    objects_init(*o, layout->layout, device, grid);
    *out = o.release();
  });
}
int earb200_objects_grid_points(const earb200_objects *calc) { return calc ? calc->grid.n_points : 0; }
void earb200_objects_destroy(earb200_objects *calc) { delete calc; }
int earb200_objects_num_outputs(const earb200_objects *calc) { return calc ? calc->n_out : 0; }
int64_t earb200_objects_launch_count(const earb200_objects *calc) {
  return calc ? calc->launches : 0;
}

int earb200_selftest_fast_sqrt(int device, int64_t n, const double *x, int64_t *n_mismatch) {
  return guarded([&] {
    if (n < 0 || (n > 0 && !x) || !n_mismatch) throw Error(INVALID_ARGUMENT, "invalid argument");
    use_device(device);
    DeviceBuffer<double> d_x;
    DeviceBuffer<unsigned long long> d_count;
    double *dx = d_x.reserve((size_t)std::max<int64_t>(n, 1));
    unsigned long long *dc = d_count.reserve(1);
    EARB_CUDA(cudaMemsetAsync(dc, 0, sizeof(unsigned long long), 0));
    if (n > 0) EARB_CUDA(cudaMemcpyAsync(dx, x, sizeof(double) * (size_t)n, cudaMemcpyHostToDevice, 0));
    launch_selftest_fast_sqrt(dx, (long)n, dc, 0);
    EARB_CUDA(cudaGetLastError());
    unsigned long long h = 0;
    EARB_CUDA(cudaMemcpyAsync(&h, dc, sizeof(h), cudaMemcpyDeviceToHost, 0));
    EARB_CUDA(cudaStreamSynchronize(0));
    *n_mismatch = (int64_t)h;
  });
}

int earb200_objects_set_profiling(earb200_objects *calc, int enable) {
  return guarded([&] {
    if (!calc) throw Error(INVALID_ARGUMENT, "null argument");
    calc->timer.enabled = enable != 0;
  });
}
int earb200_objects_kernel_time(earb200_objects *calc, double *ms, int64_t *launches,
                                int64_t *passes) {
  return guarded([&] {
    if (!calc || !ms) throw Error(INVALID_ARGUMENT, "null argument");
    use_device(calc->device);
    *ms = calc->timer.collect(launches);
    if (passes) *passes = calc->timed_passes;
    calc->timed_passes = 0;
  });
}

int earb200_objects_calculate_device(earb200_objects *calc, int64_t n,
                                     const earb200_objects_soa *d_in, int out_is_double,
                                     void *d_direct, void *d_diffuse, int32_t *d_region_idx,
                                     void *stream) {
  return guarded([&] {
    if (!calc || !d_in) throw Error(INVALID_ARGUMENT, "null argument");
    if (n < 0) throw Error(INVALID_ARGUMENT, "negative batch size");
    if (n == 0) return;
    if (!d_direct || !d_diffuse) throw Error(INVALID_ARGUMENT, "null output buffer");
    if (!d_in->azimuth || !d_in->elevation || !d_in->distance || !d_in->width || !d_in->height || !d_in->depth ||
        !d_in->gain || !d_in->diffuse)
      throw Error(INVALID_ARGUMENT, "null metadata field (all eight arrays are required)");
    check_flags(n, d_in->flags);
    use_device(calc->device);
    ObjectsSoA soa{d_in->azimuth, d_in->elevation, d_in->distance, d_in->width, d_in->height,
                   d_in->depth,   d_in->gain,      d_in->diffuse,  nullptr};
    if (out_is_double)
      objects_run<double>(*calc, n, soa, calc->n_out, calc->d_out_index.ptr, (double *)d_direct,
                          (double *)d_diffuse, d_region_idx, (cudaStream_t)stream);
    else
      objects_run<float>(*calc, n, soa, calc->n_out, calc->d_out_index.ptr, (float *)d_direct,
                         (float *)d_diffuse, d_region_idx, (cudaStream_t)stream);
  });
}

int earb200_objects_enable_extensions(earb200_objects *calc, uint32_t mask) {
  return guarded([&] {
    if (!calc) throw Error(INVALID_ARGUMENT, "null argument");
    if (mask & ~(uint32_t)(EARB200_EXT_DIVERGENCE | EARB200_EXT_CHANNEL_LOCK))
      throw Error(NOT_IMPLEMENTED, "not implemented: only divergence and channelLock exist as extensions");
    calc->extensions = mask;
  });
}

// device arrays; ext fields may be null (feature absent from the whole batch)
int earb200_objects_calculate_ext_device(earb200_objects *calc, int64_t n, const earb200_objects_soa *d_in,
                                         const earb200_objects_ext *d_ext, int out_is_double, void *d_direct,
                                         void *d_diffuse, void *stream) {
  return guarded([&] {
    if (!calc || !d_in || !d_ext) throw Error(INVALID_ARGUMENT, "null argument");
    if (n < 0) throw Error(INVALID_ARGUMENT, "negative batch size");
    if (n == 0) return;
    if (!d_direct || !d_diffuse) throw Error(INVALID_ARGUMENT, "null output buffer");
    if (!d_in->azimuth || !d_in->elevation || !d_in->distance || !d_in->width || !d_in->height || !d_in->depth ||
        !d_in->gain || !d_in->diffuse)
      throw Error(INVALID_ARGUMENT, "null metadata field (all eight arrays are required)");
    // without the opt-in the behaviour is libear's (ref gain_calculator_objects.cpp:37-43)
    if (d_ext->channel_lock_max_distance && !(calc->extensions & EARB200_EXT_CHANNEL_LOCK))
      throw Error(NOT_IMPLEMENTED, "not implemented: channelLock");
    if (d_ext->divergence && !(calc->extensions & EARB200_EXT_DIVERGENCE))
      throw Error(NOT_IMPLEMENTED, "not implemented: divergence");
    check_flags(n, d_in->flags);
    use_device(calc->device);
    ObjectsSoA soa{d_in->azimuth, d_in->elevation, d_in->distance, d_in->width, d_in->height,
                   d_in->depth,   d_in->gain,      d_in->diffuse,  nullptr};
    if (out_is_double)
      objects_run_ext<double>(*calc, n, soa, d_ext->divergence, d_ext->divergence_azimuth_range,
                              d_ext->channel_lock_max_distance, (double *)d_direct, (double *)d_diffuse,
                              (cudaStream_t)stream);
    else
      objects_run_ext<float>(*calc, n, soa, d_ext->divergence, d_ext->divergence_azimuth_range,
                             d_ext->channel_lock_max_distance, (float *)d_direct, (float *)d_diffuse,
                             (cudaStream_t)stream);
  });
}

// host arrays: staged, run through the device entry point, copied back (drop-in semantics)
int earb200_objects_calculate_ext_host(earb200_objects *calc, int64_t n, const earb200_objects_soa *in,
                                       const earb200_objects_ext *ext, int out_is_double, void *direct,
                                       void *diffuse) {
  return guarded([&] {
    if (!calc || !in || !ext) throw Error(INVALID_ARGUMENT, "null argument");
    if (n < 0) throw Error(INVALID_ARGUMENT, "negative batch size");
    if (n == 0) return;
    if (!direct || !diffuse) throw Error(INVALID_ARGUMENT, "null output buffer");
    if (!in->azimuth || !in->elevation || !in->distance || !in->width || !in->height || !in->depth || !in->gain ||
        !in->diffuse)
      throw Error(INVALID_ARGUMENT, "null metadata field (all eight arrays are required)");
    earb200_objects &o = *calc;
    use_device(o.device);
    const double *src[11] = {in->azimuth, in->elevation, in->distance, in->width, in->height, in->depth, in->gain,
                             in->diffuse, ext->divergence, ext->divergence_azimuth_range,
                             ext->channel_lock_max_distance};
    double *d = o.d_in.reserve(11 * (size_t)n);
    const double *dev[11];
    for (int f = 0; f < 11; f++) {
      dev[f] = nullptr;
      if (!src[f]) continue;
      EARB_CUDA(cudaMemcpyAsync(d + (size_t)f * n, src[f], sizeof(double) * n, cudaMemcpyHostToDevice, 0));
      dev[f] = d + (size_t)f * n;
    }
    earb200_objects_soa d_in{dev[0], dev[1], dev[2], dev[3], dev[4], dev[5], dev[6], dev[7], in->flags};
    earb200_objects_ext d_ext{dev[8], dev[9], dev[10]};
    const size_t n_el = (size_t)n * o.n_out, esz = out_is_double ? 8 : 4;
    void *d_out = out_is_double ? (void *)o.d_out_d.reserve(2 * n_el) : (void *)o.d_out_f.reserve(2 * n_el);
    void *d_dir = d_out, *d_dif = (char *)d_out + n_el * esz;
    const int rc = earb200_objects_calculate_ext_device(calc, n, &d_in, &d_ext, out_is_double, d_dir, d_dif, nullptr);
    if (rc != EARB200_OK) throw Error((Status)rc, last_error());
    EARB_CUDA(cudaMemcpyAsync(direct, d_dir, n_el * esz, cudaMemcpyDeviceToHost, 0));
    EARB_CUDA(cudaMemcpyAsync(diffuse, d_dif, n_el * esz, cudaMemcpyDeviceToHost, 0));
    EARB_CUDA(cudaStreamSynchronize(0));
  });
}

int earb200_objects_calculate_host(earb200_objects *calc, int64_t n,
                                   const earb200_objects_soa *in, int out_is_double,
                                   void *direct, void *diffuse, int32_t *region_idx) {
  return guarded([&] {
    if (!calc || !in) throw Error(INVALID_ARGUMENT, "null argument");
    if (n < 0) throw Error(INVALID_ARGUMENT, "negative batch size");
    if (n == 0) return;
    if (!direct || !diffuse) throw Error(INVALID_ARGUMENT, "null output buffer");
    check_flags(n, in->flags);
    earb200_objects &o = *calc;
    use_device(o.device);
    const double *src[8] = {in->azimuth, in->elevation, in->distance, in->width,
                            in->height,  in->depth,     in->gain,     in->diffuse};
    for (int k = 0; k < 8; k++)
      if (!src[k]) throw Error(INVALID_ARGUMENT, "null input array");
    if (out_is_double)
      objects_host_pipeline<double>(o, n, src, (double *)direct, (double *)diffuse, region_idx);
    else
      objects_host_pipeline<float>(o, n, src, (float *)direct, (float *)diffuse, region_idx);
  });
}

int earb200_objects_extent_host(earb200_objects *calc, int64_t n, const double *xyz,
                                const double *width, const double *height,
                                const double *depth, double *out) {
  return guarded([&] {
    if (!calc) throw Error(INVALID_ARGUMENT, "null argument");
    if (n < 0) throw Error(INVALID_ARGUMENT, "negative batch size");
    if (n == 0) return;
    if (!xyz || !width || !height || !depth || !out) throw Error(INVALID_ARGUMENT, "null argument");
    earb200_objects &o = *calc;
    use_device(o.device);
    // layout of the staging buffer: xyz[3n] width[n] height[n] depth[n] ones[n] zeros[n]
    double *d = o.d_in.reserve(8 * (size_t)n);
    EARB_CUDA(cudaMemcpyAsync(d, xyz, sizeof(double) * 3 * n, cudaMemcpyHostToDevice, 0));
    EARB_CUDA(cudaMemcpyAsync(d + 3 * n, width, sizeof(double) * n, cudaMemcpyHostToDevice, 0));
    EARB_CUDA(cudaMemcpyAsync(d + 4 * n, height, sizeof(double) * n, cudaMemcpyHostToDevice, 0));
    EARB_CUDA(cudaMemcpyAsync(d + 5 * n, depth, sizeof(double) * n, cudaMemcpyHostToDevice, 0));
    std::vector<double> ones((size_t)n, 1.0);
    EARB_CUDA(cudaMemcpyAsync(d + 6 * n, ones.data(), sizeof(double) * n, cudaMemcpyHostToDevice, 0));
    EARB_CUDA(cudaMemsetAsync(d + 7 * n, 0, sizeof(double) * n, 0));
    ObjectsSoA soa{nullptr, nullptr, nullptr, d + 3 * n, d + 4 * n, d + 5 * n, d + 6 * n, d + 7 * n, d};
    const size_t n_el = (size_t)n * o.S;
    double *dd = o.d_out_d.reserve(2 * n_el);
    // gain 1, diffuse 0, identity scatter: "direct" is the raw panning vector
    objects_run<double>(o, n, soa, o.S, o.d_identity_index.ptr, dd, dd + n_el, nullptr, 0);
    EARB_CUDA(cudaMemcpyAsync(out, dd, sizeof(double) * n_el, cudaMemcpyDeviceToHost, 0));
    EARB_CUDA(cudaStreamSynchronize(0));
  });
}

// No counterpart in libear: earb200_objects_calculate_device without any allocation, for callers
// that cannot call cudaMalloc.  The per-call scratch is the file-scope device array of
// objects_static.cu (shared by every call on the device), carved up here; the four kernels are the
// ones of objects_run_chunk, queued in line on the caller's stream.
int earb200_objects_calculate_device_static(earb200_objects *calc, int64_t n, const earb200_objects_soa *d_in,
                                            int out_is_double, void *d_direct, void *d_diffuse,
                                            int32_t *d_region_idx, void *stream) {
  return guarded([&] {
    if (!calc || !d_in) throw Error(INVALID_ARGUMENT, "null argument");
    if (n < 0) throw Error(INVALID_ARGUMENT, "negative batch size");
    if (n > EARB_OBJECTS_STATIC_MAX)
      throw Error(INVALID_ARGUMENT, "batch of " + std::to_string(n) + " objects exceeds the static scratch of " +
                                        "earb200_objects_calculate_device_static (EARB200_OBJECTS_STATIC_MAX = " +
                                        std::to_string(EARB_OBJECTS_STATIC_MAX) +
                                        " objects per call); split the batch or use earb200_objects_calculate_device");
    if (n == 0) return;
    if (!d_direct || !d_diffuse) throw Error(INVALID_ARGUMENT, "null output buffer");
    if (!d_in->azimuth || !d_in->elevation || !d_in->distance || !d_in->width || !d_in->height || !d_in->depth ||
        !d_in->gain || !d_in->diffuse)
      throw Error(INVALID_ARGUMENT, "null metadata field (all eight arrays are required)");
    check_flags(n, d_in->flags);
    earb200_objects &o = *calc;
    use_device(o.device);
    void *base = nullptr;
    size_t bytes = 0;
    EARB_CUDA(objects_static_scratch(o.device, &base, &bytes));
    constexpr size_t N = EARB_OBJECTS_STATIC_MAX;
    unsigned char *p = static_cast<unsigned char *>(base);
    unsigned int *d_counters = reinterpret_cast<unsigned int *>(p);
    p += 256;
    ExtentPass *d_passes = reinterpret_cast<ExtentPass *>(p);
    p += 2 * N * sizeof(ExtentPass);
    ObjectAux *d_aux = reinterpret_cast<ObjectAux *>(p);
    p += N * sizeof(ObjectAux);
    float *d_results = reinterpret_cast<float *>(p);
    p += 2 * N * EARB_MAX_SPK * sizeof(float);
    double *d_point_gains = reinterpret_cast<double *>(p);
    p += N * EARB_MAX_SPK * sizeof(double);
    int32_t *d_point_list = reinterpret_cast<int32_t *>(p);
    p += N * sizeof(int32_t);
    if ((size_t)(p - static_cast<unsigned char *>(base)) != bytes) throw Error(INTERNAL_ERROR, "static scratch layout");

    cudaStream_t s = (cudaStream_t)stream;
    ObjectsSoA c{d_in->azimuth, d_in->elevation, d_in->distance, d_in->width, d_in->height,
                 d_in->depth,   d_in->gain,      d_in->diffuse,  nullptr};
    // the array keeps the previous call's contents: the two counters must start at zero
    EARB_CUDA(cudaMemsetAsync(d_counters, 0, 2 * sizeof(unsigned int), s));
    launch_objects_prelude(n, c, d_passes, d_aux, d_counters, d_point_list, d_region_idx != nullptr, s);
    if (o.use_cells)
      launch_extent_cell(o.n_cells, o.grid.n_spk, o.d_cell_stream.ptr, o.d_cell_sums.ptr, o.d_cell_centres.ptr,
                         o.d_cell_points.ptr, d_counters, d_passes, d_results, o.n_sms, s);
    else
      launch_extent_core_tc(o.grid, o.d_grid_points.ptr, o.d_grid_bimage.ptr, d_counters, d_passes, d_results,
                            o.n_sms, s);
    if (out_is_double)
      launch_objects_epilogue<double>(o.panner.d_tables, o.grid.n_spk, n, c, d_passes, d_aux, d_results,
                                      d_point_gains, d_counters, d_point_list, o.n_sms, o.n_out, o.d_out_index.ptr,
                                      (double *)d_direct, (double *)d_diffuse, d_region_idx, s);
    else
      launch_objects_epilogue<float>(o.panner.d_tables, o.grid.n_spk, n, c, d_passes, d_aux, d_results,
                                     d_point_gains, d_counters, d_point_list, o.n_sms, o.n_out, o.d_out_index.ptr,
                                     (float *)d_direct, (float *)d_diffuse, d_region_idx, s);
    o.launches += 4;
    EARB_CUDA(cudaGetLastError());
  });
}

}  // extern "C"
