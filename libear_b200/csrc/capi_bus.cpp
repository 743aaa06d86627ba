// C ABI: the loudspeaker bus shared by the GPUs of one node -- the exchange step of
// object-sharded mixing (ref include/ear/gain_calculators.hpp:43-56: "summed with other
// objects into an n-channel bus") as part of the product, over peer memory, overlapped with the
// render that produces the partial buses (csrc/bus_kernels.cu, RenderBus in render_tc_kernels.cu).
//
// One handle per rank.  Every rank allocates ONE device region
//     [reduce flags | gather flags | slots, set 0 | slots, set 1 | full bus]
// and makes it reachable from its peers: across processes through a CUDA IPC handle the host
// program hands around however it likes (MPI, torch.distributed, a socket), inside one process
// through earb200_bus_connect_local.  Two slot sets alternate between epochs, so a rank may start
// storing the next epoch's tiles into a peer that is still summing the current one.
#include <cstring>
#include <memory>

#include "capi_common.hpp"
#include "kernels.hpp"

using namespace earb;

struct earb200_bus {
  int device = 0, rank = 0, n_ranks = 1, M = 0, n_sms = 148;
  int64_t T = 0, slice_len = 0;
  size_t slot_floats = 0, bytes = 0, off_slots[2] = {0, 0}, off_full = 0;
  unsigned char *mem = nullptr;
  unsigned char *peer[EARB_BUS_MAX_RANKS] = {};
  bool ipc_opened[EARB_BUS_MAX_RANKS] = {};
  bool connected = false;
  uint32_t epoch = 1;        // the epoch being rendered
  uint32_t published = 0;    // last epoch whose tiles this rank has announced
  uint32_t gathered = 0;     // last epoch reduced with gather
  int64_t launches = 0;
  ~earb200_bus() {
    for (int r = 0; r < EARB_BUS_MAX_RANKS; r++)
      if (ipc_opened[r] && peer[r]) cudaIpcCloseMemHandle(peer[r]);
    if (mem) cudaFree(mem);
  }
  uint32_t *flags(int r, int which) const { return reinterpret_cast<uint32_t *>(peer[r] + which * 256); }
  float *slots(int r, int set) const { return reinterpret_cast<float *>(peer[r] + off_slots[set]); }
  float *full(int r) const { return reinterpret_cast<float *>(peer[r] + off_full); }
};

namespace {
  constexpr size_t kFlagBytes = 512;  // 2 x 64 uint32: reduce flags, gather flags

  void require_connected(const earb200_bus &b) {
    if (!b.connected) throw Error(INVALID_ARGUMENT, "the bus is not connected to its peers yet");
  }
}  // namespace

// the map the render kernel stores through: this rank's slot at every owner, current epoch
RenderBus earb200_bus_map(const earb200_bus &b, int64_t t_origin) {
  RenderBus m{};
  m.n_ranks = b.n_ranks;
  m.rank = b.rank;
  m.slice_len = b.slice_len;
  m.t_origin = t_origin;
  for (int o = 0; o < b.n_ranks; o++) m.slot[o] = b.slots(o, (int)(b.epoch & 1)) + (size_t)b.rank * b.slot_floats;
  return m;
}

int64_t earb200_bus_capacity(const earb200_bus *bus) { return bus->slice_len * bus->n_ranks; }
int earb200_bus_num_channels(const earb200_bus *bus) { return bus->M; }
void earb200_bus_count_launch(earb200_bus *bus) { bus->launches++; }

extern "C" {

int earb200_bus_create(int device, int rank, int n_ranks, int n_channels, int64_t n_samples, earb200_bus **out) {
  return guarded([&] {
    if (!out) throw Error(INVALID_ARGUMENT, "null argument");
    if (n_ranks < 1 || n_ranks > EARB_BUS_MAX_RANKS || rank < 0 || rank >= n_ranks)
      throw Error(INVALID_ARGUMENT, "invalid rank / number of ranks (at most 8: the GPUs of one node)");
    if (n_channels < 1 || n_channels > EARB_MAX_OUT || n_samples < 1 || n_samples > (int64_t)1 << 30)
      throw Error(INVALID_ARGUMENT, "invalid bus shape (at most 2^30 samples)");
    use_device(device);
    auto b = std::make_unique<earb200_bus>();
    b->device = device, b->rank = rank, b->n_ranks = n_ranks, b->M = n_channels, b->T = n_samples;
    cudaDeviceProp prop;
    EARB_CUDA(cudaGetDeviceProperties(&prop, device));
    b->n_sms = prop.multiProcessorCount;
    // slices are whole render tiles (and whole float4s)
    const int64_t per = (n_samples + n_ranks - 1) / n_ranks;
    b->slice_len = (per + RENDER_TC_TILE - 1) / RENDER_TC_TILE * RENDER_TC_TILE;
    b->slot_floats = (size_t)n_channels * (size_t)b->slice_len;
    const size_t set_bytes = sizeof(float) * b->slot_floats * n_ranks;
    b->off_slots[0] = kFlagBytes;
    b->off_slots[1] = kFlagBytes + set_bytes;
    b->off_full = kFlagBytes + 2 * set_bytes;
    // the full bus covers whole slices, so that a gather never stores past its end
    b->bytes = b->off_full + sizeof(float) * b->slot_floats * n_ranks;
    EARB_CUDA(cudaMalloc(&b->mem, b->bytes));
    EARB_CUDA(cudaMemset(b->mem, 0, b->bytes));
    b->peer[rank] = b->mem;
    b->connected = n_ranks == 1;
    *out = b.release();
  });
}

void earb200_bus_destroy(earb200_bus *bus) { delete bus; }

int earb200_bus_ipc_handle_bytes(void) { return (int)sizeof(cudaIpcMemHandle_t); }

int earb200_bus_get_ipc_handle(earb200_bus *bus, void *handle_out) {
  return guarded([&] {
    if (!bus || !handle_out) throw Error(INVALID_ARGUMENT, "null argument");
    use_device(bus->device);
    cudaIpcMemHandle_t h;
    EARB_CUDA(cudaIpcGetMemHandle(&h, bus->mem));
    std::memcpy(handle_out, &h, sizeof(h));
  });
}

int earb200_bus_open_ipc(earb200_bus *bus, const void *handles) {
  return guarded([&] {
    if (!bus || !handles) throw Error(INVALID_ARGUMENT, "null argument");
    use_device(bus->device);
    for (int r = 0; r < bus->n_ranks; r++) {
      if (r == bus->rank || bus->peer[r]) continue;
      cudaIpcMemHandle_t h;
      std::memcpy(&h, (const unsigned char *)handles + (size_t)r * sizeof(h), sizeof(h));
      void *p = nullptr;
      EARB_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
      bus->peer[r] = (unsigned char *)p;
      bus->ipc_opened[r] = true;
    }
    bus->connected = true;
  });
}

// all ranks live in THIS process (one handle per device): direct peer access, no IPC
int earb200_bus_connect_local(earb200_bus *const *buses, int n) {
  return guarded([&] {
    if (!buses || n < 1) throw Error(INVALID_ARGUMENT, "null argument");
    for (int i = 0; i < n; i++)
      if (!buses[i] || buses[i]->n_ranks != n || buses[i]->rank != i || buses[i]->M != buses[0]->M ||
          buses[i]->T != buses[0]->T)
        throw Error(INVALID_ARGUMENT, "connect_local needs the handles of all ranks, in rank order, of one shape");
    for (int i = 0; i < n; i++) {
      use_device(buses[i]->device);
      for (int j = 0; j < n; j++) {
        if (i == j || buses[i]->device == buses[j]->device) continue;
        int can = 0;
        EARB_CUDA(cudaDeviceCanAccessPeer(&can, buses[i]->device, buses[j]->device));
        if (!can) throw Error(CUDA_ERROR, "the devices of the bus cannot access each other's memory");
        cudaError_t e = cudaDeviceEnablePeerAccess(buses[j]->device, 0);
        if (e == cudaErrorPeerAccessAlreadyEnabled)
          cudaGetLastError();
        else
          EARB_CUDA(e);
      }
      for (int j = 0; j < n; j++) buses[i]->peer[j] = buses[j]->mem;
      buses[i]->connected = true;
    }
  });
}

int earb200_bus_slice(const earb200_bus *bus, int rank, int64_t *t0, int64_t *t1) {
  return guarded([&] {
    if (!bus || rank < 0 || rank >= bus->n_ranks) throw Error(INVALID_ARGUMENT, "invalid argument");
    const int64_t a = std::min<int64_t>(bus->T, (int64_t)rank * bus->slice_len);
    if (t0) *t0 = a;
    if (t1) *t1 = std::min<int64_t>(bus->T, a + bus->slice_len);
  });
}

const float *earb200_bus_result(const earb200_bus *bus) { return bus ? bus->full(bus->rank) : nullptr; }
int64_t earb200_bus_result_stride(const earb200_bus *bus) { return bus ? bus->slice_len * bus->n_ranks : 0; }
int64_t earb200_bus_launch_count(const earb200_bus *bus) { return bus ? bus->launches : 0; }

// the samples [t0, t0 + n) of this rank's full bus into a caller's planar buffer (device to device)
int earb200_bus_copy_result(const earb200_bus *bus, float *d_dst, int64_t dst_stride, int64_t t0, int64_t n,
                            void *stream) {
  return guarded([&] {
    if (!bus || !d_dst) throw Error(INVALID_ARGUMENT, "null argument");
    if (t0 < 0 || n < 0 || t0 + n > bus->slice_len * bus->n_ranks) throw Error(INVALID_ARGUMENT, "range outside the bus");
    use_device(bus->device);
    EARB_CUDA(cudaMemcpy2DAsync(d_dst, sizeof(float) * dst_stride, bus->full(bus->rank) + t0,
                                sizeof(float) * bus->slice_len * bus->n_ranks, sizeof(float) * n, bus->M,
                                cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  });
}

// a local [M][n] partial bus (bus samples t0 ...) into the slots: for renders that did not go
// through earb200_interp_process_device_bus
int earb200_bus_push(earb200_bus *bus, const float *d_partial, int64_t partial_stride, int64_t t0, int64_t n,
                     void *stream) {
  return guarded([&] {
    if (!bus || !d_partial) throw Error(INVALID_ARGUMENT, "null argument");
    require_connected(*bus);
    if (t0 < 0 || n < 0 || t0 + n > bus->slice_len * bus->n_ranks) throw Error(INVALID_ARGUMENT, "range outside the bus");
    use_device(bus->device);
    launch_bus_push(d_partial, partial_stride, bus->M, t0, n, earb200_bus_map(*bus, t0), (cudaStream_t)stream);
    EARB_CUDA(cudaGetLastError());
    bus->launches++;
  });
}

// Tells every peer that this rank's tiles of the epoch are in place (in stream order: after the
// renders queued so far).  earb200_bus_reduce does it itself when it has not been done; the
// separate call is for hosts that drive several ranks through ONE stream or device, where every
// rank has to publish before any rank starts waiting.
int earb200_bus_publish(earb200_bus *bus, void *stream) {
  return guarded([&] {
    if (!bus) throw Error(INVALID_ARGUMENT, "null argument");
    require_connected(*bus);
    if (bus->published == bus->epoch) return;
    use_device(bus->device);
    uint32_t *pf[EARB_BUS_MAX_RANKS];
    for (int r = 0; r < bus->n_ranks; r++) pf[r] = bus->flags(r, 0);
    launch_bus_signal(pf, bus->n_ranks, bus->rank, bus->epoch, (cudaStream_t)stream);
    EARB_CUDA(cudaGetLastError());
    bus->published = bus->epoch;
    bus->launches++;
  });
}

// Ends the epoch: publishes (if not done), waits for the tiles of all ranks, sums the slots of this
// rank's slice in rank order.  d_out (nullable): [M][slice] with row pitch out_stride receives the
// slice (reduce-scatter).  gather != 0: every rank's full bus receives it as well (all-reduce);
// earb200_bus_wait_result then orders later work on the stream after the arrival of ALL slices.
int earb200_bus_reduce(earb200_bus *bus, float *d_out, int64_t out_stride, int gather, void *stream) {
  return guarded([&] {
    if (!bus) throw Error(INVALID_ARGUMENT, "null argument");
    require_connected(*bus);
    if (!d_out && !gather) throw Error(INVALID_ARGUMENT, "nothing to write: no output slice and no gather");
    use_device(bus->device);
    cudaStream_t st = (cudaStream_t)stream;
    uint32_t *pf[EARB_BUS_MAX_RANKS];
    if (bus->published != bus->epoch) {
      for (int r = 0; r < bus->n_ranks; r++) pf[r] = bus->flags(r, 0);
      launch_bus_signal(pf, bus->n_ranks, bus->rank, bus->epoch, st);
      bus->published = bus->epoch;
      bus->launches++;
    }
    float *full[EARB_BUS_MAX_RANKS];
    for (int r = 0; r < bus->n_ranks; r++) full[r] = bus->full(r);
    int64_t t0 = 0, t1 = 0;
    earb200_bus_slice(bus, bus->rank, &t0, &t1);
    launch_bus_reduce(bus->flags(bus->rank, 0), bus->n_ranks, bus->epoch, bus->slots(bus->rank, (int)(bus->epoch & 1)),
                      (long long)bus->slot_floats, bus->M, bus->slice_len, t1 - t0, d_out, out_stride,
                      gather ? full : nullptr, bus->slice_len * bus->n_ranks, (long long)bus->rank * bus->slice_len,
                      bus->n_sms, st);
    bus->launches++;
    if (gather) {
      for (int r = 0; r < bus->n_ranks; r++) pf[r] = bus->flags(r, 1);
      launch_bus_signal(pf, bus->n_ranks, bus->rank, bus->epoch, st);
      bus->gathered = bus->epoch;
      bus->launches++;
    }
    EARB_CUDA(cudaGetLastError());
    bus->epoch++;
  });
}

// after a reduce with gather: later work on `stream` runs once the slices of ALL ranks have landed
// in this rank's full bus (earb200_bus_result)
int earb200_bus_wait_result(earb200_bus *bus, void *stream) {
  return guarded([&] {
    if (!bus) throw Error(INVALID_ARGUMENT, "null argument");
    require_connected(*bus);
    if (bus->gathered == 0) throw Error(INVALID_ARGUMENT, "no reduce with gather has been queued");
    use_device(bus->device);
    launch_bus_wait(bus->flags(bus->rank, 1), bus->n_ranks, bus->gathered, (cudaStream_t)stream);
    EARB_CUDA(cudaGetLastError());
    bus->launches++;
  });
}

}  // extern "C"
