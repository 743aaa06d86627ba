// Shared plumbing of the C ABI implementation: error translation, CUDA checks and a
// grow-only device buffer.
#pragma once
#include <cuda_runtime.h>

#include <string>
#include <utility>
#include <vector>

#include "../../include/earb200.h"
#include "host_setup.hpp"

namespace earb {

  std::string &last_error();  // thread-local

  inline void cuda_check(cudaError_t e, const char *what) {
    if (e != cudaSuccess)
      throw Error(CUDA_ERROR, std::string(what) + ": " + cudaGetErrorString(e));
  }
#define EARB_CUDA(call) ::earb::cuda_check((call), #call)

  // selects the device for the calling thread; fails loudly when there is none
  inline void use_device(int device) {
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
      throw Error(CUDA_ERROR,
                  std::string("no CUDA device available (libear-b200 has no CPU fallback): ") +
                      cudaGetErrorString(e));
    if (device < 0 || device >= count) throw Error(INVALID_ARGUMENT, "invalid CUDA device index");
    EARB_CUDA(cudaSetDevice(device));
  }

  template <typename T>
  struct DeviceBuffer {
    T *ptr = nullptr;
    size_t cap = 0;  // elements
    DeviceBuffer() = default;
    DeviceBuffer(const DeviceBuffer &) = delete;
    DeviceBuffer &operator=(const DeviceBuffer &) = delete;
    ~DeviceBuffer() {
      if (ptr) cudaFree(ptr);
    }
    T *reserve(size_t n) {
      if (n > cap) {
        if (ptr) EARB_CUDA(cudaFree(ptr));
        ptr = nullptr;
        cap = 0;
        EARB_CUDA(cudaMalloc(&ptr, n * sizeof(T)));
        cap = n;
      }
      return ptr;
    }
  };

  // CUDA-event stopwatch around selected kernel launches (bench roofline measurement)
  struct KernelTimer {
    bool enabled = false;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> pending, pool;
    ~KernelTimer() {
      for (auto &p : pending) { cudaEventDestroy(p.first); cudaEventDestroy(p.second); }
      for (auto &p : pool) { cudaEventDestroy(p.first); cudaEventDestroy(p.second); }
    }
    void start(cudaStream_t s) {
      if (!enabled) return;
      std::pair<cudaEvent_t, cudaEvent_t> p;
      if (!pool.empty()) {
        p = pool.back();
        pool.pop_back();
      } else {
        EARB_CUDA(cudaEventCreate(&p.first));
        EARB_CUDA(cudaEventCreate(&p.second));
      }
      EARB_CUDA(cudaEventRecord(p.first, s));
      pending.push_back(p);
    }
    void stop(cudaStream_t s) {
      if (!enabled) return;
      EARB_CUDA(cudaEventRecord(pending.back().second, s));
    }
    // synchronises on the recorded events; returns total ms and the number of intervals
    double collect(int64_t *count) {
      double total = 0.0;
      for (auto &p : pending) {
        EARB_CUDA(cudaEventSynchronize(p.second));
        float ms = 0.0f;
        EARB_CUDA(cudaEventElapsedTime(&ms, p.first, p.second));
        total += ms;
        pool.push_back(p);
      }
      if (count) *count = (int64_t)pending.size();
      pending.clear();
      return total;
    }
  };

  // runs `body`, translating exceptions into a status code + earb200_last_error()
  template <typename F>
  int guarded(F &&body) {
    try {
      body();
      return EARB200_OK;
    } catch (const Error &e) {
      last_error() = e.what();
      return (int)e.status;
    } catch (const std::exception &e) {
      last_error() = std::string("internal error: ") + e.what();
      return EARB200_INTERNAL_ERROR;
    }
  }

}  // namespace earb

// opaque handle types of the C ABI
struct earb200_layout {
  earb::HostLayout layout;
};
