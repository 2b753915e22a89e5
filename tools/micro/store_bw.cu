// Per-SM global store throughput on B200: one CTA of 1024 threads per SM, coalesced stores of 8 / 16 bytes per thread,
// P rows written per iteration (like stage B writes P problems).   nvcc -arch=sm_100a -O3 -o store_bw store_bw.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int W, int P>
__global__ void __launch_bounds__(1024, 1) Store(double* out, size_t row_stride, int iters) {
  // each CTA owns a contiguous region: P rows of iters*1024*W doubles
  double* base = out + static_cast<size_t>(blockIdx.x) * P * row_stride + threadIdx.x * W;
  const double v = threadIdx.x * 1.0;
  for (int k = 0; k < iters; ++k) {
#pragma unroll
    for (int p = 0; p < P; ++p) {
      double* o = base + p * row_stride + static_cast<size_t>(k) * 1024 * W;
      if (W == 1) *o = v + p;
      else *reinterpret_cast<double2*>(o) = make_double2(v + p, v - p);
    }
  }
}

template <int W, int P>
void Run(double* buf, int ctas, int iters) {
  const size_t row = static_cast<size_t>(iters) * 1024 * W;
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  Store<W, P><<<ctas, 1024>>>(buf, row, iters);
  cudaEventRecord(a);
  for (int r = 0; r < 5; ++r) Store<W, P><<<ctas, 1024>>>(buf, row, iters);
  cudaEventRecord(b);
  cudaEventSynchronize(b);
  float ms;
  cudaEventElapsedTime(&ms, a, b);
  const double bytes = 5.0 * ctas * P * row * 8;
  printf("W=%d doubles/thread P=%d ctas=%d: %.1f GB/s total, %.1f B/clk/SM at 1.9 GHz (%.3f ms)\n", W, P, ctas,
         bytes / ms / 1e6, bytes / ms / 1e6 / ctas / 1.9, ms / 5);
}

int main() {
  double* buf;
  const size_t n = size_t(1) << 28;   // 2 GiB
  cudaMalloc(&buf, n * 8);
  for (int ctas : {1, 16, 148}) {
    const int iters = 2048 / 4;
    Run<1, 4>(buf, ctas, iters);
    Run<2, 4>(buf, ctas, iters / 2);
    Run<1, 1>(buf, ctas, iters * 4);
    Run<2, 1>(buf, ctas, iters * 2);
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
