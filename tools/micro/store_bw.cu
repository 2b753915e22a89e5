// Per-SM global store throughput on B200: one CTA per SM, T threads, coalesced 8-byte stores, P rows per step
// (as stage B writes P problems), U independent steps unrolled.   nvcc -arch=sm_100a -O3 -o store_bw store_bw.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int T, int P, int U>
__global__ void __launch_bounds__(T, 1) Store(double* out, size_t row_stride, int iters) {
  double* base = out + static_cast<size_t>(blockIdx.x) * P * row_stride + threadIdx.x;
  const double v = threadIdx.x * 1.0;
  for (int k = 0; k < iters; k += U) {
#pragma unroll
    for (int u = 0; u < U; ++u)
#pragma unroll
      for (int p = 0; p < P; ++p) base[p * row_stride + static_cast<size_t>(k + u) * T] = v + p + u;
  }
}

template <int T, int P, int U>
void Run(double* buf, int ctas, size_t per_cta_doubles) {
  const int iters = static_cast<int>(per_cta_doubles / P / T) / U * U;
  const size_t row = static_cast<size_t>(iters) * T;
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  Store<T, P, U><<<ctas, T>>>(buf, row, iters);
  cudaEventRecord(a);
  for (int r = 0; r < 5; ++r) Store<T, P, U><<<ctas, T>>>(buf, row, iters);
  cudaEventRecord(b);
  cudaEventSynchronize(b);
  float ms;
  cudaEventElapsedTime(&ms, a, b);
  const double bytes = 5.0 * ctas * P * row * 8;
  printf("T=%4d P=%d U=%d ctas=%3d: %7.1f GB/s total, %5.1f B/clk/SM at 1.9 GHz\n", T, P, U, ctas, bytes / ms / 1e6,
         bytes / ms / 1e6 / ctas / 1.9);
}

int main() {
  double* buf;
  const size_t n = size_t(3) << 27;   // 3 GiB
  if (cudaMalloc(&buf, n * 8) != cudaSuccess) { printf("alloc failed\n"); return 1; }
  const size_t per_cta = size_t(1) << 21;   // 16 MiB per CTA
  for (int ctas : {16, 148}) {
    Run<1024, 2, 1>(buf, ctas, per_cta);
    Run<512, 2, 1>(buf, ctas, per_cta);
    Run<256, 2, 1>(buf, ctas, per_cta);
    Run<256, 2, 4>(buf, ctas, per_cta);
    Run<128, 2, 4>(buf, ctas, per_cta);
    Run<256, 4, 4>(buf, ctas, per_cta);
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
