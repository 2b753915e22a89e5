// kernel_common.cuh — helpers shared by the evaluation kernel variants.
#pragma once
#include <cuda_runtime.h>

#include <cstddef>

#include "items.cuh"
#include "kernels.h"

namespace tb {

#ifndef TB_THREADS
#define TB_THREADS kEntryThreads
#endif
constexpr int kThreads = TB_THREADS;   // threads per problem lane

// Phase clocks of CTA 0 (timing build only: make timing; tools/phase_timing.py).  TB_TIMING_HERE is defined by the one
// translation unit per kernel family that owns the counters.
#if defined(TB_TIMING) && defined(TB_TIMING_HERE)
static __device__ unsigned long long tb_phase_clk[64];
#define TB_MARK(k) TB_MARK_T(k, 0)
#define TB_MARK_T(k, who)                                                      \
  if (blockIdx.x == 0 && threadIdx.x == (who) && (k) < 64) {                           \
    const long long now_ = clock64();                                          \
    tb_phase_clk[(k)] += static_cast<unsigned long long>(now_ - mark_);        \
    mark_ = now_;                                                              \
  }
#define TB_MARK_INIT long long mark_ = clock64();
#define TB_PHASE_CLOCK_READER(name)                                                              \
  extern "C" int name(unsigned long long* out, int n) {                                          \
    unsigned long long h[64];                                                                    \
    if (cudaMemcpyFromSymbol(h, tb::tb_phase_clk, sizeof(h)) != cudaSuccess) return -1;          \
    for (int i = 0; i < n && i < 64; ++i) out[i] = h[i];                                         \
    unsigned long long z[64] = {};                                                               \
    cudaMemcpyToSymbol(tb::tb_phase_clk, z, sizeof(z));                                          \
    return 0;                                                                                    \
  }
#else
#define TB_MARK(k)
#define TB_MARK_T(k, who)
#define TB_MARK_INIT
#define TB_PHASE_CLOCK_READER(name)
#endif

__device__ __forceinline__ bool NonFinite(double v) {
  return (__double2hiint(v) & 0x7ff00000) == 0x7ff00000;
}

// Stage B of one segment.  Thread t takes entries t, t + NT, ...: a warp's descriptor load and its P stores per
// entry are coalesced.  Descriptors are fetched kDescBatch at a time into a register ring of A = DescAhead<RT> batches: a batch
// is loaded A batches before its use (the first ones before the barrier that ends stage A), so the L2 latency
// of the descriptor stream - several hundred clocks while the value stores keep the memory pipe busy - stays off the
// critical path.
// Classes with optimised durations run one problem per CTA (8 warps): fewer warps to hide the latency, deeper ring.
template <bool RT> struct DescAhead { static constexpr int value = RT ? 8 : 4; };
constexpr int kDescBatch = 4;

template <int NT, int U>
__device__ __forceinline__ void LoadDescriptors(const uint32_t* __restrict__ ent, int left, uint32_t (&d)[U]) {
#pragma unroll
  for (int u = 0; u < U; ++u) d[u] = left > u * NT ? __ldg(ent + u * NT) : kNoEntry;
}
// the first A batches of a segment (thread's first entry at ent, `left` entries from there on)
template <int NT, int A, int U>
__device__ __forceinline__ void LoadDescriptorRing(const uint32_t* __restrict__ ent, int left, uint32_t (&d)[A][U]) {
#pragma unroll
  for (int a = 0; a < A; ++a) LoadDescriptors<NT, U>(ent + a * U * NT, left - a * U * NT, d[a]);
}

// Per-problem constants of the P problems of a group into shared memory, 8 bytes per thread.  The height table behind
// the fixed fields (program.h: ProblemConst::table, 264 of the record's 352 bytes) is only read for problems on a table
// terrain, so only those copy it.
// TABLES = false (kernels of classes on analytic terrains): the fixed fields only, no look at the terrain id.
template <int P, int NT, bool TABLES>
__device__ __forceinline__ void FetchProblemConsts(ProblemConst* dst, const ProblemConst* __restrict__ pcs, int base_n, int n_problems,
                                                   int tid) {
  constexpr int kFixed = static_cast<int>(offsetof(ProblemConst, table) / 8);
  constexpr int kWords = TABLES ? static_cast<int>(sizeof(ProblemConst) / 8) : kFixed;
  static_assert(sizeof(ProblemConst) % 8 == 0 && offsetof(ProblemConst, table) % 8 == 0, "ProblemConst is copied in 8-byte words");
  for (int i = tid; i < P * kWords; i += NT) {
    const int p = i / kWords, w = i - p * kWords;
    const ProblemConst* src = pcs + min(base_n + p, n_problems - 1);
    if (w < kFixed || src->terrain_id == kTerrainTableId)
      reinterpret_cast<double*>(dst + p)[w] = reinterpret_cast<const double*>(src)[w];
  }
}

// 8-byte asynchronous global -> shared copy; src_bytes = 0 writes zeros (cp.async zero-fill)
__device__ __forceinline__ void CpAsync8(void* smem_dst, const void* gsrc, int src_bytes) {
  const unsigned d = static_cast<unsigned>(__cvta_generic_to_shared(smem_dst));
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(d), "l"(gsrc), "r"(src_bytes) : "memory");
}

}  // namespace tb
