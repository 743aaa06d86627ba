// K5: HOA AllRAD decoder design (FP64).
//
// Replaces the 5,200-point contractions of GainCalculatorHOAImpl::calculate
// (ref src/hoa/gain_calculator_hoa.cpp:57-62) and the functions they call in
// src/hoa/hoa.hpp: sph_harm (:88-99), calc_Y_virt (:103-121), the product
// D = G_virt * (Y_virt^T / P) and the Frobenius norm of D * Y_virt used by
// normalize_decode_matrix (:140-143).  G_virt is produced by K1 on the t-design points.
// Compiled with --fmad=false (FP64 products and sums round separately, as on the CPU).
#include <cuda_runtime.h>

#include "kernels.hpp"

namespace earb {
  namespace {
    // associated Legendre function P_n^m(x) without the Condon-Shortley phase
    // (ref hoa.hpp:16-18 Alegendre), m >= 0, by upward recurrence in n
    __device__ double d_alegendre(int n, int m, double x) {
      double pmm = 1.0;
      if (m > 0) {
        const double somx2 = sqrt((1.0 - x) * (1.0 + x));
        double fact = 1.0;
        for (int i = 1; i <= m; i++) {
          pmm *= fact * somx2;
          fact += 2.0;
        }
      }
      if (n == m) return pmm;
      double pmmp1 = x * (2.0 * m + 1.0) * pmm;
      if (n == m + 1) return pmmp1;
      double pll = 0.0;
      for (int ll = m + 2; ll <= n; ll++) {
        pll = (x * (2.0 * ll - 1.0) * pmmp1 - (ll + m - 1.0) * pmm) / (double)(ll - m);
        pmm = pmmp1;
        pmmp1 = pll;
      }
      return pll;
    }

    // Y[c][p] = norm_c * P_n^|m|(sin el) * {sqrt2 cos(m az), 1, -sqrt2 sin(m az)}
    __global__ void hoa_sh_kernel(int P, int C, const double *__restrict__ xyz,
                                  const int *__restrict__ orders, const int *__restrict__ degrees,
                                  const double *__restrict__ norms, double *__restrict__ Y) {
      const int p = blockIdx.x * blockDim.x + threadIdx.x;
      if (p >= P) return;
      const double x = xyz[3 * p], y = xyz[3 * p + 1], z = xyz[3 * p + 2];
      const double az = -atan2(x, y);
      const double el = atan2(z, hypot(x, y));
      const double sin_el = sin(el);
      for (int c = 0; c < C; c++) {
        const int n = orders[c], m = degrees[c];
        double scale;
        if (m > 0)
          scale = sqrt(2.0) * cos(m * az);
        else if (m < 0)
          scale = -sqrt(2.0) * sin(m * az);
        else
          scale = 1.0;
        Y[(size_t)c * P + p] = norms[c] * d_alegendre(n, abs(m), sin_el) * scale;
      }
    }

    __device__ double block_sum(double v) {
      __shared__ double red[32];
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
      __syncthreads();
      if (l == 0) red[w] = v;
      __syncthreads();
      v = (threadIdx.x < (blockDim.x >> 5)) ? red[threadIdx.x] : 0.0;
      if (w == 0)
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      return v;  // valid in thread 0
    }

    // D[s][c] = sum_p G[p][s] * (Y[c][p] / P); one block per (s, c)
    __global__ void hoa_decode_kernel(int P, int S, int C, const double *__restrict__ G,
                                      const double *__restrict__ Y, double *__restrict__ D) {
      const int s = blockIdx.x, c = blockIdx.y;
      double acc = 0.0;
      for (int p = threadIdx.x; p < P; p += blockDim.x)
        acc += G[(size_t)p * S + s] * (Y[(size_t)c * P + p] / (double)P);
      acc = block_sum(acc);
      if (threadIdx.x == 0) D[s * C + c] = acc;
    }

    // partial[b] = sum over the block's points and all s of (sum_c D[s][c] Y[c][p])^2
    __global__ void hoa_power_kernel(int P, int S, int C, const double *__restrict__ D,
                                     const double *__restrict__ Y, double *__restrict__ partial) {
      const int p = blockIdx.x * blockDim.x + threadIdx.x;
      double acc = 0.0;
      if (p < P) {
        for (int s = 0; s < S; s++) {
          double v = 0.0;
          for (int c = 0; c < C; c++) v += D[s * C + c] * Y[(size_t)c * P + p];
          acc += v * v;
        }
      }
      acc = block_sum(acc);
      if (threadIdx.x == 0) partial[blockIdx.x] = acc;
    }
  }  // namespace

  int hoa_power_blocks(int P) { return (P + 255) / 256; }

  void launch_hoa_design(int P, int S, int C, const double *d_xyz, const double *d_G,
                         const int *d_orders, const int *d_degrees, const double *d_norms,
                         double *d_Y, double *d_D, double *d_partial, cudaStream_t stream) {
    hoa_sh_kernel<<<(P + 127) / 128, 128, 0, stream>>>(P, C, d_xyz, d_orders, d_degrees, d_norms, d_Y);
    hoa_decode_kernel<<<dim3(S, C), 256, 0, stream>>>(P, S, C, d_G, d_Y, d_D);
    hoa_power_kernel<<<hoa_power_blocks(P), 256, 0, stream>>>(P, S, C, d_D, d_Y, d_partial);
  }
}  // namespace earb
