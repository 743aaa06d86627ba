// C ABI: multi-channel delay line (ear::dsp::DelayBuffer, ref include/ear/dsp/delay_buffer.hpp:12-29,
// src/dsp/delay_buffer_impl.cpp:12-47).  The delay memory stays on the device between calls.
#include <memory>

#include "capi_common.hpp"
#include "kernels.hpp"

using namespace earb;

struct earb200_delay {
  int device = 0, n_sms = 0;
  int C = 0;
  int64_t delay = 0;
  DeviceBuffer<float> d_mem[2];  // ping-pong: the kernel reads one and writes the other
  int cur = 0;
  DeviceBuffer<float> d_in, d_out;
  int64_t launches = 0;
};

namespace {
  void delay_process_device(earb200_delay &h, int64_t n, const float *d_in, int64_t in_stride, float *d_out,
                            int64_t out_stride, cudaStream_t stream) {
    if (n < 0) throw Error(INVALID_ARGUMENT, "negative sample count");
    if (n == 0 || h.C == 0) return;
    if (!d_out) throw Error(INVALID_ARGUMENT, "null output buffer");
    use_device(h.device);
    if (h.delay == 0) {  // [in] -> [out]
      if (d_in)
        EARB_CUDA(cudaMemcpy2DAsync(d_out, sizeof(float) * out_stride, d_in, sizeof(float) * in_stride,
                                    sizeof(float) * n, h.C, cudaMemcpyDeviceToDevice, stream));
      else
        EARB_CUDA(cudaMemset2DAsync(d_out, sizeof(float) * out_stride, 0, sizeof(float) * n, h.C, stream));
      return;
    }
    launch_delay(h.C, n, h.delay, d_in, in_stride, d_out, out_stride, h.d_mem[h.cur].ptr, h.d_mem[h.cur ^ 1].ptr,
                 h.n_sms, stream);
    EARB_CUDA(cudaGetLastError());
    h.cur ^= 1;
    h.launches++;
  }
}  // namespace

extern "C" {

int earb200_delay_create(int n_channels, int64_t delay, int device, earb200_delay **out) {
  return guarded([&] {
    if (!out) throw Error(INVALID_ARGUMENT, "null argument");
    if (n_channels < 0 || delay < 0) throw Error(INVALID_ARGUMENT, "negative channel count or delay");
    use_device(device);
    auto h = std::make_unique<earb200_delay>();
    h->device = device;
    h->C = n_channels;
    h->delay = delay;
    cudaDeviceProp prop;
    EARB_CUDA(cudaGetDeviceProperties(&prop, device));
    h->n_sms = prop.multiProcessorCount;
    const size_t m = std::max<size_t>(1, (size_t)n_channels * (size_t)delay);
    for (int k = 0; k < 2; k++) EARB_CUDA(cudaMemset(h->d_mem[k].reserve(m), 0, sizeof(float) * m));
    *out = h.release();
  });
}

void earb200_delay_destroy(earb200_delay *d) { delete d; }

int64_t earb200_delay_get_delay(const earb200_delay *d) { return d ? d->delay : 0; }

int64_t earb200_delay_launch_count(const earb200_delay *d) { return d ? d->launches : 0; }

int earb200_delay_process_device(earb200_delay *d, int64_t nsamples, const float *d_in, int64_t in_stride,
                                 float *d_out, int64_t out_stride, void *stream) {
  return guarded([&] {
    if (!d) throw Error(INVALID_ARGUMENT, "null handle");
    delay_process_device(*d, nsamples, d_in, in_stride, d_out, out_stride, (cudaStream_t)stream);
  });
}

int earb200_delay_process_host(earb200_delay *d, int64_t nsamples, const float *in, int64_t in_stride,
                               float *out, int64_t out_stride) {
  return guarded([&] {
    if (!d) throw Error(INVALID_ARGUMENT, "null handle");
    if (nsamples < 0) throw Error(INVALID_ARGUMENT, "negative sample count");
    if (nsamples == 0 || d->C == 0) return;
    if (!out) throw Error(INVALID_ARGUMENT, "null output buffer");
    use_device(d->device);
    const size_t n = (size_t)nsamples, C = (size_t)d->C;
    float *di = nullptr;
    if (in) {
      di = d->d_in.reserve(C * n);
      EARB_CUDA(cudaMemcpy2D(di, sizeof(float) * n, in, sizeof(float) * in_stride, sizeof(float) * n, C,
                             cudaMemcpyHostToDevice));
    }
    float *dout = d->d_out.reserve(C * n);
    delay_process_device(*d, nsamples, di, nsamples, dout, nsamples, 0);
    EARB_CUDA(cudaMemcpy2D(out, sizeof(float) * out_stride, dout, sizeof(float) * n, sizeof(float) * n, C,
                           cudaMemcpyDeviceToHost));
  });
}

}  // extern "C"
