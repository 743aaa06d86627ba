// Scratch of earb200_objects_calculate_device_static: a file-scope device array instead of
// cudaMalloc'ed buffers, for callers that must not allocate (CUDA graph capture, serving loops,
// harnesses that forbid cudaMalloc).  No counterpart in
// libear.
//
// Every GPU that loads libearb200.so holds the array (EARB_OBJECTS_STATIC_BYTES, about 2.4 MB)
// until the process exits.  All calls on a device share it and it keeps the last call's
// contents: the entry point clears the two counters that must start at zero, everything else is
// written before it is read.  Calls that overlap on the GPU (two streams, two threads) race.
//
// In host code `g_objects_static_scratch` is only a stand-in for the symbol, not a device address:
// the address comes from cudaGetSymbolAddress, once per device with that device current, below.
#include <cuda_runtime.h>

#include <mutex>

#include "kernels.hpp"

namespace earb {

  static __device__ __align__(256) unsigned char g_objects_static_scratch[EARB_OBJECTS_STATIC_BYTES];

  namespace {
    constexpr int kMaxDevices = 64;
    std::mutex g_mutex;
    void *g_address[kMaxDevices] = {};
  }  // namespace

  // Device address of the array on `device`, which the caller has made current.  The look-up runs
  // once per device; later calls return the cached pointer.
  cudaError_t objects_static_scratch(int device, void **ptr, size_t *bytes) {
    *ptr = nullptr;
    *bytes = 0;
    if (device < 0 || device >= kMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_mutex);
    if (!g_address[device]) {
      int current = -1;
      cudaError_t e = cudaGetDevice(&current);
      if (e != cudaSuccess) return e;
      if (current != device) return cudaErrorInvalidDevice;
      void *p = nullptr;
      e = cudaGetSymbolAddress(&p, g_objects_static_scratch);
      if (e != cudaSuccess) return e;
      g_address[device] = p;
    }
    *ptr = g_address[device];
    *bytes = EARB_OBJECTS_STATIC_BYTES;
    return cudaSuccess;
  }

}  // namespace earb
