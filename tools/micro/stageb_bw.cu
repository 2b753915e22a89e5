// Stage B in isolation: value = G[g] * B[b] from 4-byte descriptors, written in storage order.  Modes:
//   stg   : thread t takes entries t, t + NT, ...; coalesced 8-byte streaming stores (the lock-step kernel's stage B)
//   ring  : per-warp shared-memory ring + cp.async.bulk (the RING variants' stage B)
//   nostore: products only (accumulated), no global store: the shared-memory + issue cost of stage B
// Run on 16 CTAs (no HBM contention) and 148 CTAs: tells an SM-internal limit from the HBM limit.
//   nvcc -arch=sm_100a -O3 -o stageb_bw stageb_bw.cu
#include <cstdio>
#include <cstdint>
#include <vector>
#include <cuda_runtime.h>

constexpr int kEntries = 24000, kG = 6000, kB = 700;

__device__ __forceinline__ void FenceProxyAsync() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void BulkStore(void* g, const void* s, unsigned bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(g), "r"((unsigned)__cvta_generic_to_shared(s)), "r"(bytes) : "memory");
}

template <int NT, int P, int MODE, int D>   // MODE 0 stg, 1 ring, 2 nostore
__global__ void __launch_bounds__(NT, 1) StageB(const uint32_t* __restrict__ ent, double* out, int groups, double* sink) {
  extern __shared__ double sm[];
  double* G = sm;                 // kG * P, interleaved
  double* Bs = G + kG * P;        // kB
  double* ring = Bs + kB + 2;
  const int tid = threadIdx.x;
  for (int i = tid; i < kG * P; i += NT) G[i] = 1.0 + i * 1e-6;
  for (int i = tid; i < kB; i += NT) Bs[i] = 0.5 + i * 1e-3;
  __syncthreads();
  const char* Gb = (const char*)G; const char* Bb = (const char*)Bs;
  double acc = 0.0;
  unsigned ctr = 0;
  for (int grp = 0; grp < groups; ++grp) {
    double* o = out + ((size_t)blockIdx.x * groups + grp) * P * kEntries;
    if (MODE == 1) {
      constexpr int BE = 256, ROW = BE + 2, NW = NT / 32;
      const int warp = tid >> 5, lane = tid & 31;
      double* rw = ring + warp * (D * P * ROW);
      for (int blk = warp; blk * BE < kEntries; blk += NW) {
        double* sb = rw + (ctr % D) * (P * ROW);
        if (lane == 0) asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(D - 1) : "memory");
        __syncwarp();
        const int e0 = blk * BE;
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int e = e0 + lane + 32 * u;
          if (e < kEntries) {
            const uint32_t w = __ldg(ent + e);
            const double bv = *(const double*)(Bb + (w >> 16));
#pragma unroll
            for (int p = 0; p < P; ++p) sb[p * ROW + lane + 32 * u] = *(const double*)(Gb + (w & 0xFFFF) * P + 8 * p) * bv;
          }
        }
        FenceProxyAsync();
        __syncwarp();
        if (lane == 0) {
          const int nc = min(BE, kEntries - e0);
#pragma unroll
          for (int p = 0; p < P; ++p) BulkStore(o + (size_t)p * kEntries + e0, sb + p * ROW, nc * 8u);
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
        ++ctr;
      }
    } else {
      for (int e = tid; e < kEntries; e += 4 * NT) {
        uint32_t w[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) w[u] = e + u * NT < kEntries ? __ldg(ent + e + u * NT) : 0u;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          if (e + u * NT >= kEntries) break;
          const double bv = *(const double*)(Bb + (w[u] >> 16));
#pragma unroll
          for (int p = 0; p < P; ++p) {
            const double v = *(const double*)(Gb + (w[u] & 0xFFFF) * P + 8 * p) * bv;
            if (MODE == 0) __stcs(o + (size_t)p * kEntries + e + u * NT, v);
            else acc += v;
          }
        }
      }
    }
    __syncthreads();
  }
  if (MODE == 1 && (tid & 31) == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  if (MODE == 2 && acc == 12345.678) sink[0] = acc;
}

template <int NT, int P, int MODE, int D>
void Run(const uint32_t* ent, double* out, double* sink, int ctas, const char* name) {
  const int groups = 40;
  size_t smem = sizeof(double) * (kG * P + kB + 2 + (MODE == 1 ? (NT / 32) * D * P * 258 : 0)) + 16;
  cudaFuncSetAttribute(StageB<NT, P, MODE, D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  StageB<NT, P, MODE, D><<<ctas, NT, smem>>>(ent, out, groups, sink);
  cudaEventRecord(a);
  StageB<NT, P, MODE, D><<<ctas, NT, smem>>>(ent, out, groups, sink);
  cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  const double bytes = (double)ctas * groups * P * kEntries * 8;
  printf("%-8s NT=%3d P=%d D=%d ctas=%3d: %7.1f GB/s, %5.1f B/clk/SM, %6.0f clk per %d-problem group  (%s)\n", name, NT, P, D, ctas,
         bytes / ms / 1e6, bytes / ms / 1e6 / ctas / 1.9, ms * 1e-3 * 1.9e9 / groups, P, cudaGetErrorString(cudaGetLastError()));
}

int main(int argc, char** argv) {
  // descriptor pattern: "random" (worst case: every lane another bank), or "seq": consecutive entries read consecutive G
  // slots and runs of 12 entries share one B constant (close to the real class tables, conflict-free)
  const bool seq = argc > 1 && argv[1][0] == 's';
  printf("# descriptors: %s\n", seq ? "sequential G slots, B shared by runs of 12" : "random");
  std::vector<uint32_t> h(kEntries);
  uint64_t s = 88172645463325252ull;
  int e = 0;
  for (auto& w : h) {
    s ^= s << 13; s ^= s >> 7; s ^= s << 17;
    if (seq) w = (uint32_t)((e % kG) * 8) | ((uint32_t)(((e / 12) % kB) * 8) << 16);
    else w = (uint32_t)((s % kG) * 8) | ((uint32_t)(((s >> 20) % kB) * 8) << 16);
    ++e;
  }
  uint32_t* ent; double *out, *sink;
  cudaMalloc(&ent, kEntries * 4); cudaMemcpy(ent, h.data(), kEntries * 4, cudaMemcpyHostToDevice);
  cudaMalloc(&out, (size_t)148 * 40 * 2 * kEntries * 8); cudaMalloc(&sink, 8);
  for (int ctas : {16, 74, 148}) {
    Run<512, 2, 0, 2>(ent, out, sink, ctas, "stg");
    Run<256, 1, 0, 2>(ent, out, sink, ctas, "stg");
    Run<512, 2, 2, 2>(ent, out, sink, ctas, "nostore");
    Run<256, 1, 2, 2>(ent, out, sink, ctas, "nostore");
    Run<256, 1, 1, 2>(ent, out, sink, ctas, "ring");
    Run<256, 1, 1, 4>(ent, out, sink, ctas, "ring");
    Run<512, 1, 1, 2>(ent, out, sink, ctas, "ring");
    Run<256, 2, 1, 2>(ent, out, sink, ctas, "ring");
  }
  return 0;
}
