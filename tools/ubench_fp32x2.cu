// Micro-benchmark: what a packed fp32 instruction (FFMA2) costs on sm_100 next to scalar FFMA and
// to ALU work (LOP3), per SM sub-partition.  One CTA per SM, W warps per sub-partition, every warp
// runs a long unrolled stream of independent instructions; the result is warp-instructions per
// cycle per sub-partition.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/build/ubench_fp32x2 tools/ubench_fp32x2.cu
#include <cuda_runtime.h>

#include <cstdio>

typedef unsigned long long u64;

__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) {
  u64 r;
  asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
__device__ __forceinline__ float fma1(float a, float b, float c) {
  float r;
  asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
  return r;
}
__device__ __forceinline__ unsigned lop(unsigned a, unsigned b, unsigned c) {
  unsigned r;
  asm volatile("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
  return r;
}

constexpr int kIters = 2048;

// MODE 0: 8 FFMA2 per step; 1: 8 FFMA; 2: 16 FFMA; 3: 8 FFMA2 + 8 LOP3; 4: 8 FFMA + 8 LOP3; 5: 8 LOP3;
// 6: 8 FFMA2 + 8 FFMA
template <int MODE>
__global__ void bench(float *out, long long *cycles, float seed) {
  u64 p[8];
  float f[16];
  unsigned u[8];
#pragma unroll
  for (int i = 0; i < 8; i++) {
    p[i] = ((u64)__float_as_uint(seed + i) << 32) | __float_as_uint(seed - i);
    u[i] = threadIdx.x * 17 + i;
  }
#pragma unroll
  for (int i = 0; i < 16; i++) f[i] = seed * i;
  const u64 m = ((u64)__float_as_uint(0.999f) << 32) | __float_as_uint(1.001f);
  const u64 a = ((u64)__float_as_uint(0.001f) << 32) | __float_as_uint(-0.001f);
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < kIters; it++) {
#pragma unroll
    for (int i = 0; i < 8; i++) {
      if (MODE == 0 || MODE == 3 || MODE == 6) p[i] = fma2(p[i], m, a);
      if (MODE == 1 || MODE == 2 || MODE == 4 || MODE == 6) f[i] = fma1(f[i], 0.999f, 0.001f);
      if (MODE == 2) f[i + 8] = fma1(f[i + 8], 0.999f, 0.001f);
      if (MODE == 3 || MODE == 4 || MODE == 5) u[i] = lop(u[i], u[(i + 1) & 7], 0x5a5a5a5au);
    }
  }
  const long long t1 = clock64();
  float s = 0.0f;
#pragma unroll
  for (int i = 0; i < 8; i++) s += __uint_as_float((unsigned)p[i]) + __uint_as_float((unsigned)(p[i] >> 32)) + u[i];
#pragma unroll
  for (int i = 0; i < 16; i++) s += f[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int MODE>
void run(const char *name, int per_step, int warps_per_sp, float *out, long long *cyc) {
  bench<MODE><<<148, 128 * warps_per_sp>>>(out, cyc, 1.0f);
  bench<MODE><<<148, 128 * warps_per_sp>>>(out, cyc, 1.0f);
  cudaDeviceSynchronize();
  long long h[148];
  cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
  double c = 0;
  for (int i = 0; i < 148; i++) c += (double)h[i];
  c /= 148;
  const double inst = (double)kIters * per_step * warps_per_sp;  // warp instructions per sub-partition
  printf("%-22s warps/sub-partition %d: %.3f warp-instructions/cycle/sub-partition (%.2f cycles per step of %d)\n",
         name, warps_per_sp, inst / c, c / kIters / warps_per_sp, per_step);
}

int main() {
  float *out;
  long long *cyc;
  cudaMalloc(&out, sizeof(float) * 148 * 1024);
  cudaMalloc(&cyc, sizeof(long long) * 148);
  for (int w = 1; w <= 4; w *= 2) {
    run<0>("8 FFMA2", 8, w, out, cyc);
    run<1>("8 FFMA", 8, w, out, cyc);
    run<2>("16 FFMA", 16, w, out, cyc);
    run<5>("8 LOP3", 8, w, out, cyc);
    run<3>("8 FFMA2 + 8 LOP3", 16, w, out, cyc);
    run<4>("8 FFMA + 8 LOP3", 16, w, out, cyc);
    run<6>("8 FFMA2 + 8 FFMA", 16, w, out, cyc);
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
