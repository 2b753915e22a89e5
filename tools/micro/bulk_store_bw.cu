// Per-SM throughput of cp.async.bulk shared -> global stores on B200: one CTA per SM, NI issuing threads, each
// bulk copy CH bytes, D copies in flight per issuing thread.   nvcc -arch=sm_100a -O3 -o bulk_store_bw bulk_store_bw.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void BulkStore(void* g, const void* s, unsigned bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(g),
               "r"(static_cast<unsigned>(__cvta_generic_to_shared(s))), "r"(bytes) : "memory");
}

template <int D>
__global__ void __launch_bounds__(256, 1) Bulk(char* out, size_t per_cta, int ch, int ni) {
  extern __shared__ uint4 sm[];
  char* s = reinterpret_cast<char*>(sm);
  for (int i = threadIdx.x; i < ni * D * ch / 16; i += blockDim.x) sm[i] = make_uint4(i, i, i, i);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  if (threadIdx.x < ni) {
    char* dst = out + blockIdx.x * per_cta + static_cast<size_t>(threadIdx.x) * (per_cta / ni);
    const size_t n = per_cta / ni / ch;
    for (size_t k = 0; k < n; ++k) {
      BulkStore(dst + k * ch, s + (threadIdx.x * D + (k % D)) * ch, ch);
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(D - 1) : "memory");
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
}

template <int D>
void Run(char* buf, int ctas, int ch, int ni) {
  const size_t per_cta = size_t(16) << 20;
  const size_t smem = size_t(ni) * D * ch;
  if (smem > 200 * 1024) return;
  cudaFuncSetAttribute(Bulk<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  Bulk<D><<<ctas, 256, smem>>>(buf, per_cta, ch, ni);
  cudaEventRecord(a);
  for (int r = 0; r < 5; ++r) Bulk<D><<<ctas, 256, smem>>>(buf, per_cta, ch, ni);
  cudaEventRecord(b);
  cudaEventSynchronize(b);
  float ms;
  cudaEventElapsedTime(&ms, a, b);
  const double bytes = 5.0 * ctas * per_cta;
  printf("ctas=%3d chunk=%6d B issuers=%d depth=%d: %7.1f GB/s total, %5.1f B/clk/SM at 1.9 GHz  (%s)\n", ctas, ch, ni, D,
         bytes / ms / 1e6, bytes / ms / 1e6 / ctas / 1.9, cudaGetErrorString(cudaGetLastError()));
}

int main() {
  char* buf;
  if (cudaMalloc(&buf, size_t(148) * (16 << 20)) != cudaSuccess) { printf("alloc failed\n"); return 1; }
  for (int ctas : {16, 148})
    for (int ch : {512, 2048, 16384, 49152})
      for (int ni : {1, 2, 4}) {
        Run<1>(buf, ctas, ch, ni);
        Run<2>(buf, ctas, ch, ni);
      }
  return 0;
}
