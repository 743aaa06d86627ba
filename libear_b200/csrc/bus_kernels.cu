// Exchange step of object-sharded mixing between the GPUs of ONE node, over peer memory
// (NVLink / NVSwitch), without a collective library.
//
// libear's contract for an Objects renderer is "gains are applied to this channel, and summed with
// other objects into an n-channel bus" (ref include/ear/gain_calculators.hpp:43-56,
// docs/dsp.rst:40-71).  With the objects sharded over GPUs every GPU holds a partial bus and the
// sum over GPUs is the one exchange step of the whole path.  Here the bus is cut into
// one time slice per rank; the render kernel's epilogue stores each partial tile straight into the
// slot the rendering rank has in the staging memory of the slice's owner (render_tc_kernels.cu,
// RenderBus) -- the transfer rides on the render, tile by tile -- and the kernels below finish the
// step: a flag per (owner, source) published with release semantics at system scope, and the
// owner's sum over the slots in rank order (deterministic: the same bits on every run), optionally
// stored into every rank's full bus (all-reduce semantics) instead of only the owner's slice
// (reduce-scatter semantics).
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>

#include "kernels.hpp"

namespace earb {
  namespace {
    struct PeerFlags {
      uint32_t *p[EARB_BUS_MAX_RANKS];
    };
    struct PeerBuses {
      float *p[EARB_BUS_MAX_RANKS];
    };

    __device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t *p) {
      uint32_t v;
      asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
      return v;
    }
    // spins until flags[r] >= epoch for every rank; a protocol error (a peer that never arrives)
    // traps after about two seconds instead of hanging the GPU
    __device__ __forceinline__ void wait_flags(const uint32_t *flags, int n_ranks, uint32_t epoch) {
      if (threadIdx.x < (unsigned)n_ranks) {
        const long long t0 = clock64();
        while ((int32_t)(ld_acquire_sys(flags + threadIdx.x) - epoch) < 0) {
          __nanosleep(200);
          if (clock64() - t0 > 4000000000ll) __trap();
        }
      }
      __syncthreads();
    }

    __global__ void bus_signal_kernel(PeerFlags peers, int n_ranks, int rank, uint32_t epoch) {
      // everything this rank stored before this kernel (the render's peer stores, stream order) is
      // made visible system-wide before the flags
      __threadfence_system();
      if (threadIdx.x < (unsigned)n_ranks)
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(peers.p[threadIdx.x] + rank), "r"(epoch) : "memory");
    }

    __global__ void bus_wait_kernel(const uint32_t *flags, int n_ranks, uint32_t epoch) {
      wait_flags(flags, n_ranks, epoch);
    }

    // one float4 of one channel per thread and step; the slots are streamed once
    __global__ void __launch_bounds__(256)
        bus_reduce_kernel(const uint32_t *flags, int n_ranks, uint32_t epoch, const float *__restrict__ slots,
                          long long slot_floats, int M, long long slice_len, long long valid_len,
                          float *__restrict__ out, long long out_stride, PeerBuses gather, int n_gather,
                          long long gather_stride, long long gather_offset) {
      wait_flags(flags, n_ranks, epoch);
      const long long quads = slice_len / 4, total = (long long)M * quads;
      for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
           i += (long long)gridDim.x * blockDim.x) {
        const long long m = i / quads, t = (i - m * quads) * 4;
        if (t >= valid_len) continue;
        const float *src = slots + m * slice_len + t;
        float4 acc = __ldcs(reinterpret_cast<const float4 *>(src));
        for (int r = 1; r < n_ranks; r++) {
          const float4 v = __ldcs(reinterpret_cast<const float4 *>(src + (long long)r * slot_floats));
          acc.x += v.x, acc.y += v.y, acc.z += v.z, acc.w += v.w;
        }
        if (t + 4 <= valid_len) {
          if (out) *reinterpret_cast<float4 *>(out + m * out_stride + t) = acc;
          for (int g = 0; g < n_gather; g++)
            __stcs(reinterpret_cast<float4 *>(gather.p[g] + m * gather_stride + gather_offset + t), acc);
        } else {
          const float a[4] = {acc.x, acc.y, acc.z, acc.w};
          for (int e = 0; t + e < valid_len; e++) {
            if (out) out[m * out_stride + t + e] = a[e];
            for (int g = 0; g < n_gather; g++) gather.p[g][m * gather_stride + gather_offset + t + e] = a[e];
          }
        }
      }
    }

    __global__ void __launch_bounds__(256)
        bus_push_kernel(const float *__restrict__ partial, long long partial_stride, int M, long long n,
                        const RenderBus bus) {
      const long long total = (long long)M * n;
      for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
           i += (long long)gridDim.x * blockDim.x) {
        const long long m = i / n, t = i - m * n, bt = t + bus.t_origin;
        const int owner = (int)(bt / bus.slice_len);
        bus.slot[owner][m * bus.slice_len + (bt - (long long)owner * bus.slice_len)] = partial[m * partial_stride + t];
      }
    }
    // the runs of a call that the fused epilogue did not take (ragged heads / tails of segments,
    // rendered into the partial bus by the FFMA kernels): one block per item
    __global__ void __launch_bounds__(256)
        bus_push_items_kernel(const RenderItem *__restrict__ items, const float *__restrict__ partial,
                              long long partial_stride, int M, const RenderBus bus) {
      const RenderItem it = items[blockIdx.x];
      const long long total = (long long)M * it.len;
      for (long long i = threadIdx.x; i < total; i += blockDim.x) {
        const long long m = i / it.len, t = it.t0 + (i - m * it.len), bt = t + bus.t_origin;
        const int owner = (int)(bt / bus.slice_len);
        bus.slot[owner][m * bus.slice_len + (bt - (long long)owner * bus.slice_len)] = partial[m * partial_stride + t];
      }
    }
  }  // namespace

  void launch_bus_push_items(const RenderItem *d_items, long n_items, const float *partial, long long partial_stride,
                             int M, const RenderBus &bus, cudaStream_t stream) {
    if (n_items <= 0) return;
    bus_push_items_kernel<<<(unsigned)n_items, 256, 0, stream>>>(d_items, partial, partial_stride, M, bus);
  }

  void launch_bus_signal(uint32_t *const *peer_flags, int n_ranks, int rank, uint32_t epoch, cudaStream_t stream) {
    PeerFlags pf{};
    for (int r = 0; r < n_ranks; r++) pf.p[r] = peer_flags[r];
    bus_signal_kernel<<<1, 32, 0, stream>>>(pf, n_ranks, rank, epoch);
  }

  void launch_bus_wait(const uint32_t *flags, int n_ranks, uint32_t epoch, cudaStream_t stream) {
    bus_wait_kernel<<<1, 32, 0, stream>>>(flags, n_ranks, epoch);
  }

  void launch_bus_reduce(const uint32_t *flags, int n_ranks, uint32_t epoch, const float *slots, long long slot_floats,
                         int M, long long slice_len, long long valid_len, float *out, long long out_stride,
                         float *const *gather, long long gather_stride, long long gather_offset, int n_sms,
                         cudaStream_t stream) {
    PeerBuses pb{};
    const int n_gather = gather ? n_ranks : 0;
    for (int r = 0; r < n_gather; r++) pb.p[r] = gather[r];
    const long long total = (long long)M * (slice_len / 4);
    const unsigned blocks = (unsigned)std::max<long long>(1, std::min<long long>((total + 255) / 256, (long long)n_sms * 8));
    bus_reduce_kernel<<<blocks, 256, 0, stream>>>(flags, n_ranks, epoch, slots, slot_floats, M, slice_len, valid_len, out,
                                                  out_stride, pb, n_gather, gather_stride, gather_offset);
  }

  void launch_bus_push(const float *partial, long long partial_stride, int M, long long t_origin, long long n,
                       const RenderBus &bus, cudaStream_t stream) {
    if (n <= 0) return;
    RenderBus b = bus;
    b.t_origin = t_origin;
    const long long total = (long long)M * n;
    const unsigned blocks = (unsigned)std::min<long long>((total + 255) / 256, 148 * 16);
    bus_push_kernel<<<blocks, 256, 0, stream>>>(partial, partial_stride, M, n, b);
  }
}  // namespace earb
