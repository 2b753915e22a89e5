// Per-SM global store throughput on B200 with 16-byte stores per lane (STG.E.128) against 8-byte stores (STG.E.64):
// one CTA per SM, T threads, each thread stores consecutive pairs of doubles, P rows per step as stage B writes P problems.
//   nvcc -arch=sm_100a -O3 -o store_bw128 store_bw128.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int T, int P, int W>   // W = doubles per store (1: STG.64, 2: STG.128)
__global__ void __launch_bounds__(T, 1) Store(double* out, size_t row_stride, int iters) {
  double* base = out + static_cast<size_t>(blockIdx.x) * P * row_stride + threadIdx.x * W;
  const double v = threadIdx.x * 1.0;
  for (int k = 0; k < iters; ++k) {
#pragma unroll
    for (int p = 0; p < P; ++p) {
      double* o = base + p * row_stride + static_cast<size_t>(k) * T * W;
      if (W == 1) __stcs(o, v + p);
      else __stcs(reinterpret_cast<double2*>(o), make_double2(v + p, v - p));
    }
  }
}

template <int T, int P, int W>
void Run(double* buf, int ctas, size_t per_cta_doubles) {
  const int iters = static_cast<int>(per_cta_doubles / P / T / W);
  const size_t row = static_cast<size_t>(iters) * T * W;
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  Store<T, P, W><<<ctas, T>>>(buf, row, iters);
  cudaEventRecord(a);
  for (int r = 0; r < 5; ++r) Store<T, P, W><<<ctas, T>>>(buf, row, iters);
  cudaEventRecord(b);
  cudaEventSynchronize(b);
  float ms;
  cudaEventElapsedTime(&ms, a, b);
  const double bytes = 5.0 * ctas * P * row * 8;
  printf("T=%4d P=%d bytes/lane=%2d ctas=%3d: %7.1f GB/s total, %5.1f B/clk/SM at 1.9 GHz\n", T, P, 8 * W, ctas, bytes / ms / 1e6,
         bytes / ms / 1e6 / ctas / 1.9);
}

int main() {
  double* buf;
  const size_t n = size_t(3) << 27;   // 3 GiB
  if (cudaMalloc(&buf, n * 8) != cudaSuccess) { printf("alloc failed\n"); return 1; }
  const size_t per_cta = size_t(1) << 21;   // 16 MiB per CTA
  for (int ctas : {16, 148}) {
    Run<512, 2, 1>(buf, ctas, per_cta);
    Run<512, 2, 2>(buf, ctas, per_cta);
    Run<256, 2, 2>(buf, ctas, per_cta);
    Run<256, 1, 2>(buf, ctas, per_cta);
    Run<128, 2, 2>(buf, ctas, per_cta);
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
