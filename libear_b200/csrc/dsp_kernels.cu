// Streaming DSP helpers of the Objects signal flow: the compensation delay
// (ref src/dsp/delay_buffer_impl.cpp:19-40) and the delayed-direct + decorrelated-diffuse mix
// (ref docs/dsp.rst:40-71).  Both are pure HBM streaming: 8 B (delay) / 12 B (mix) per sample.
#include "kernels.hpp"

namespace earb {
  namespace {
    constexpr int kThreads = 256;
    constexpr int kPerThread = 4;

    // [mem; in] -> [out; mem_new]: sample i of the concatenation goes to out[i] when
    // i < n and to mem_new[i - n] otherwise.  One thread moves four consecutive samples.
    __device__ __forceinline__ float delay_fetch(const float *__restrict__ mem, const float *__restrict__ in,
                                                 long i, long delay) {
      return i < delay ? mem[i] : (in ? __ldcs(in + (i - delay)) : 0.0f);
    }

    __global__ void __launch_bounds__(kThreads)
        delay_kernel(int C, long n, long delay, const float *__restrict__ in, long in_stride,
                     float *__restrict__ out, long out_stride, const float *__restrict__ mem,
                     float *__restrict__ mem_new, const float *__restrict__ add, long add_stride) {
      const int c = blockIdx.y;
      const float *m = mem + (long)c * delay;
      const float *x = in ? in + (long)c * in_stride : nullptr;
      float *y = out + (long)c * out_stride;
      float *mn = mem_new + (long)c * delay;
      const float *ad = add ? add + (long)c * add_stride : nullptr;  // out = delayed + add
      const long total = n + delay;
      for (long base = ((long)blockIdx.x * kThreads + threadIdx.x) * kPerThread; base < total;
           base += (long)gridDim.x * kThreads * kPerThread) {
        if (base + kPerThread <= n && ((reinterpret_cast<size_t>(y + base) & 15) == 0)) {
          float4 v;
          v.x = delay_fetch(m, x, base, delay);
          v.y = delay_fetch(m, x, base + 1, delay);
          v.z = delay_fetch(m, x, base + 2, delay);
          v.w = delay_fetch(m, x, base + 3, delay);
          if (ad) {
            v.x += __ldcs(ad + base);
            v.y += __ldcs(ad + base + 1);
            v.z += __ldcs(ad + base + 2);
            v.w += __ldcs(ad + base + 3);
          }
          __stcs(reinterpret_cast<float4 *>(y + base), v);
        } else {
#pragma unroll
          for (int k = 0; k < kPerThread; k++) {
            const long i = base + k;
            if (i >= total) break;
            const float v = delay_fetch(m, x, i, delay);
            if (i < n)
              y[i] = ad ? v + ad[i] : v;
            else
              mn[i - n] = v;
          }
        }
      }
    }

    // out = a + b, planar channels (the last stage of the Objects flow)
    __global__ void __launch_bounds__(kThreads)
        add_kernel(long n, const float *__restrict__ a, long a_stride, const float *__restrict__ b,
                   long b_stride, float *__restrict__ out, long out_stride) {
      const int c = blockIdx.y;
      const float *pa = a + (long)c * a_stride, *pb = b + (long)c * b_stride;
      float *po = out + (long)c * out_stride;
      const bool vec = ((reinterpret_cast<size_t>(pa) | reinterpret_cast<size_t>(pb) |
                         reinterpret_cast<size_t>(po)) & 15) == 0;
      for (long base = ((long)blockIdx.x * kThreads + threadIdx.x) * kPerThread; base < n;
           base += (long)gridDim.x * kThreads * kPerThread) {
        if (vec && base + kPerThread <= n) {
          const float4 va = __ldcs(reinterpret_cast<const float4 *>(pa + base));
          const float4 vb = __ldcs(reinterpret_cast<const float4 *>(pb + base));
          __stcs(reinterpret_cast<float4 *>(po + base),
                 make_float4(va.x + vb.x, va.y + vb.y, va.z + vb.z, va.w + vb.w));
        } else {
          for (long i = base; i < n && i < base + kPerThread; i++) po[i] = pa[i] + pb[i];
        }
      }
    }

    __global__ void delay_advance_kernel(long n, long delay, const float *__restrict__ in, long in_stride,
                                         const float *__restrict__ mem, float *__restrict__ mem_new) {
      const int c = blockIdx.y;
      for (long j = (long)blockIdx.x * blockDim.x + threadIdx.x; j < delay; j += (long)gridDim.x * blockDim.x)
        mem_new[(long)c * delay + j] =
            delay_fetch(mem + (long)c * delay, in ? in + (long)c * in_stride : nullptr, n + j, delay);
    }

    int grid_x(long work_items, int rows, int n_sms) {
      long blocks = (work_items + (long)kThreads * kPerThread - 1) / ((long)kThreads * kPerThread);
      // a few waves of the machine spread over the channel rows
      const long cap = ((long)n_sms * 8 + rows - 1) / rows;
      if (blocks > cap) blocks = cap;
      return (int)(blocks < 1 ? 1 : blocks);
    }
  }  // namespace

  void launch_delay(int C, long n, long delay, const float *in, long in_stride, float *out,
                    long out_stride, const float *mem, float *mem_new, const float *add, long add_stride,
                    int n_sms, cudaStream_t stream) {
    if (C <= 0 || n + delay <= 0) return;
    dim3 grid(grid_x(n + delay, C, n_sms), C);
    delay_kernel<<<grid, kThreads, 0, stream>>>(C, n, delay, in, in_stride, out, out_stride, mem, mem_new, add,
                                                add_stride);
  }

  void launch_delay_advance(int C, long n, long delay, const float *in, long in_stride, const float *mem,
                            float *mem_new, cudaStream_t stream) {
    if (C <= 0 || delay <= 0) return;
    dim3 grid((unsigned)((delay + 255) / 256), C);
    delay_advance_kernel<<<grid, 256, 0, stream>>>(n, delay, in, in_stride, mem, mem_new);
  }

  void launch_add(int C, long n, const float *a, long a_stride, const float *b, long b_stride, float *out,
                  long out_stride, int n_sms, cudaStream_t stream) {
    if (C <= 0 || n <= 0) return;
    dim3 grid(grid_x(n, C, n_sms), C);
    add_kernel<<<grid, kThreads, 0, stream>>>(n, a, a_stride, b, b_stride, out, out_stride);
  }
}  // namespace earb
