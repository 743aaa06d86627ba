// C ABI: multi-channel delay line (ear::dsp::DelayBuffer, ref include/ear/dsp/delay_buffer.hpp:12-29,
// src/dsp/delay_buffer_impl.cpp:12-47).  The delay memory stays on the device between calls.
#include <cmath>
#include <memory>
#include <vector>

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

// FFT plug-in: batched 2B-point r2c / c2r transforms (ear::FFTImpl / FFTPlan,
// ref include/ear/fft.hpp:27-62; unpacked n/2+1 bins, unnormalised inverse like src/fft_kiss.cpp:61-88)
struct earb200_fft {
  int device = 0, n_sms = 0;
  int B = 0;  // n_fft / 2
  DeviceBuffer<float> d_tw2, d_in, d_out;
  int64_t launches = 0;
};

namespace {
  void delay_process_device(earb200_delay &h, int64_t n, const float *d_in, int64_t in_stride, float *d_out,
                            int64_t out_stride, cudaStream_t stream, const float *d_add = nullptr,
                            int64_t add_stride = 0) {
    if (n < 0) throw Error(INVALID_ARGUMENT, "negative sample count");
    if (n == 0 || h.C == 0) return;
    if (!d_out) throw Error(INVALID_ARGUMENT, "null output buffer");
    use_device(h.device);
    if (h.delay == 0 && !d_add) {  // [in] -> [out]
      if (d_in)
        EARB_CUDA(cudaMemcpy2DAsync(d_out, sizeof(float) * out_stride, d_in, sizeof(float) * in_stride,
                                    sizeof(float) * n, h.C, cudaMemcpyDeviceToDevice, stream));
      else
        EARB_CUDA(cudaMemset2DAsync(d_out, sizeof(float) * out_stride, 0, sizeof(float) * n, h.C, stream));
      return;
    }
    launch_delay(h.C, n, h.delay, d_in, in_stride, d_out, out_stride, h.d_mem[h.cur].ptr, h.d_mem[h.cur ^ 1].ptr,
                 d_add, add_stride, h.n_sms, stream);
    EARB_CUDA(cudaGetLastError());
    h.cur ^= 1;
    h.launches++;
  }
}  // namespace

struct earb200_convolver;
bool earb_conv_process_device(earb200_convolver &h, int64_t K, const float *d_in, int64_t in_stride,
                              float *d_out, int64_t out_stride, cudaStream_t stream, const float *add,
                              int64_t add_stride, int64_t add_delay, const float *add_mem);
int64_t earb200_convolver_block_size_internal(const earb200_convolver *conv);

extern "C" {

// out = BlockConvolver(in) + DelayBuffer(direct): the last two stages of the Objects signal flow
// (ref docs/dsp.rst:40-71) in one pass -- on the convolver's fast path the delayed direct bus joins
// the decorrelated bus in the convolver's store and the delay line's memory is advanced by a tiny
// kernel; on the generic path the two stages run one after the other.
int earb200_convolver_process_add_delayed_device(earb200_convolver *conv, earb200_delay *d, int64_t n_blocks,
                                                 const float *d_in, int64_t in_stride,
                                                 const float *d_direct, int64_t direct_stride, float *d_out,
                                                 int64_t out_stride, void *stream) {
  return guarded([&] {
    if (!conv || !d || !d_direct) throw Error(INVALID_ARGUMENT, "null argument");
    if (n_blocks < 0) throw Error(INVALID_ARGUMENT, "negative block count");
    if (n_blocks == 0) return;
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t n = n_blocks * earb200_convolver_block_size_internal(conv);
    use_device(d->device);
    const bool fused = d->delay > 0 && earb_conv_process_device(*conv, n_blocks, d_in, in_stride, d_out, out_stride, st,
                                                                d_direct, direct_stride, d->delay,
                                                                d->d_mem[d->cur].ptr);
    if (fused) {
      launch_delay_advance(d->C, n, d->delay, d_direct, direct_stride, d->d_mem[d->cur].ptr, d->d_mem[d->cur ^ 1].ptr,
                           st);
      EARB_CUDA(cudaGetLastError());
      d->cur ^= 1;
      d->launches++;
    } else {
      if (d->delay == 0)
        earb_conv_process_device(*conv, n_blocks, d_in, in_stride, d_out, out_stride, st, nullptr, 0, 0, nullptr);
      delay_process_device(*d, n, d_direct, direct_stride, d_out, out_stride, st, d_out, out_stride);
    }
  });
}

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

int earb200_delay_process_add_device(earb200_delay *d, int64_t nsamples, const float *d_in, int64_t in_stride,
                                     const float *d_add, int64_t add_stride, float *d_out, int64_t out_stride,
                                     void *stream) {
  return guarded([&] {
    if (!d) throw Error(INVALID_ARGUMENT, "null handle");
    delay_process_device(*d, nsamples, d_in, in_stride, d_out, out_stride, (cudaStream_t)stream, d_add, add_stride);
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

int earb200_fft_create(int64_t n_fft, int device, earb200_fft **out) {
  return guarded([&] {
    if (!out) throw Error(INVALID_ARGUMENT, "null argument");
    const int64_t B = n_fft / 2;
    if (n_fft < 32 || n_fft > 8192 || (n_fft & (n_fft - 1)) != 0)
      throw Error(INVALID_ARGUMENT, "n_fft must be a power of two in [32, 8192]");
    use_device(device);
    auto h = std::make_unique<earb200_fft>();
    h->device = device;
    h->B = (int)B;
    cudaDeviceProp prop;
    EARB_CUDA(cudaGetDeviceProperties(&prop, device));
    h->n_sms = prop.multiProcessorCount;
    std::vector<float> tw(4 * (size_t)B);
    const double pi = 3.141592653589793238462643383279502884;
    for (int64_t k = 0; k < 2 * B; k++) {
      const double ang = -2.0 * pi * (double)k / (double)(2 * B);
      tw[2 * k] = (float)std::cos(ang);
      tw[2 * k + 1] = (float)std::sin(ang);
    }
    EARB_CUDA(cudaMemcpy(h->d_tw2.reserve(tw.size()), tw.data(), sizeof(float) * tw.size(), cudaMemcpyHostToDevice));
    *out = h.release();
  });
}

void earb200_fft_destroy(earb200_fft *fft) { delete fft; }

int64_t earb200_fft_size(const earb200_fft *fft) { return fft ? 2 * (int64_t)fft->B : 0; }

int64_t earb200_fft_launch_count(const earb200_fft *fft) { return fft ? fft->launches : 0; }

int earb200_fft_forward_device(earb200_fft *fft, int64_t batch, const float *d_in, float *d_out, void *stream) {
  return guarded([&] {
    if (!fft || (batch > 0 && (!d_in || !d_out))) throw Error(INVALID_ARGUMENT, "null argument");
    if (batch < 0) throw Error(INVALID_ARGUMENT, "negative batch size");
    if (batch == 0) return;
    use_device(fft->device);
    launch_fft_r2c(fft->B, batch, d_in, d_out, fft->d_tw2.ptr, fft->n_sms, (cudaStream_t)stream);
    EARB_CUDA(cudaGetLastError());
    fft->launches++;
  });
}

int earb200_fft_reverse_device(earb200_fft *fft, int64_t batch, const float *d_in, float *d_out, void *stream) {
  return guarded([&] {
    if (!fft || (batch > 0 && (!d_in || !d_out))) throw Error(INVALID_ARGUMENT, "null argument");
    if (batch < 0) throw Error(INVALID_ARGUMENT, "negative batch size");
    if (batch == 0) return;
    use_device(fft->device);
    launch_fft_c2r(fft->B, batch, d_in, d_out, fft->d_tw2.ptr, fft->n_sms, (cudaStream_t)stream);
    EARB_CUDA(cudaGetLastError());
    fft->launches++;
  });
}

int earb200_fft_forward_host(earb200_fft *fft, int64_t batch, const float *in, float *out) {
  return guarded([&] {
    if (!fft || (batch > 0 && (!in || !out))) throw Error(INVALID_ARGUMENT, "null argument");
    if (batch <= 0) {
      if (batch < 0) throw Error(INVALID_ARGUMENT, "negative batch size");
      return;
    }
    use_device(fft->device);
    const size_t n_in = (size_t)batch * 2 * fft->B, n_out = (size_t)batch * (fft->B + 1) * 2;
    EARB_CUDA(cudaMemcpy(fft->d_in.reserve(n_in), in, sizeof(float) * n_in, cudaMemcpyHostToDevice));
    launch_fft_r2c(fft->B, batch, fft->d_in.ptr, fft->d_out.reserve(n_out), fft->d_tw2.ptr, fft->n_sms, 0);
    EARB_CUDA(cudaGetLastError());
    fft->launches++;
    EARB_CUDA(cudaMemcpy(out, fft->d_out.ptr, sizeof(float) * n_out, cudaMemcpyDeviceToHost));
  });
}

int earb200_fft_reverse_host(earb200_fft *fft, int64_t batch, const float *in, float *out) {
  return guarded([&] {
    if (!fft || (batch > 0 && (!in || !out))) throw Error(INVALID_ARGUMENT, "null argument");
    if (batch <= 0) {
      if (batch < 0) throw Error(INVALID_ARGUMENT, "negative batch size");
      return;
    }
    use_device(fft->device);
    const size_t n_in = (size_t)batch * (fft->B + 1) * 2, n_out = (size_t)batch * 2 * fft->B;
    EARB_CUDA(cudaMemcpy(fft->d_in.reserve(n_in), in, sizeof(float) * n_in, cudaMemcpyHostToDevice));
    launch_fft_c2r(fft->B, batch, fft->d_in.ptr, fft->d_out.reserve(n_out), fft->d_tw2.ptr, fft->n_sms, 0);
    EARB_CUDA(cudaGetLastError());
    fft->launches++;
    EARB_CUDA(cudaMemcpy(out, fft->d_out.ptr, sizeof(float) * n_out, cudaMemcpyDeviceToHost));
  });
}

}  // extern "C"
