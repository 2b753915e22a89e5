// kernels.cu — the batched evaluation kernel (sm_100a) and its launcher.
//
// One CTA evaluates one problem at a time (grid-stride over the batch).  Shared memory holds the problem's
// spline node values, the Hermite coefficients of every polynomial and the G buffer of the current segment;
// x is read once (coalesced gather through node_src), g and the Jacobian values are written once, the
// values with fully coalesced 8-byte stores in storage order.  HBM traffic per eval = 8*(n_vars+n_rows+nnz)
// bytes plus L2-resident class tables (entries, B, items), which are shared by all problems of the batch.
#include <cuda_runtime.h>

#include "items.cuh"
#include "kernels.h"

namespace tb {

constexpr int kThreads = kEntryThreads;
#ifndef TB_MIN_CTAS
#define TB_MIN_CTAS 4
#endif
constexpr int kMinCtasPerSm = TB_MIN_CTAS;   // caps registers at 64: the long dynamic item spills to local memory, stage B does not

__device__ __forceinline__ bool NonFinite(double v) {
  return (__double2hiint(v) & 0x7ff00000) == 0x7ff00000;
}

// bar.sync on a named barrier: the 256 threads of one lane (lanes are whole warps, so this is warp-uniform)
__device__ __forceinline__ void LaneSync(int lane) {
  asm volatile("bar.sync %0, %1;" ::"r"(lane + 1), "n"(kThreads) : "memory");
}

// LANES problems are evaluated side by side by one CTA (256 threads each) so that the class tables are held once
// per SM; each lane has its own nodes/coef/G arrays and its own named barrier.
template <int WHAT, int LANES>
__global__ void __launch_bounds__(kThreads * LANES, LANES == 1 ? kMinCtasPerSm : 1)
EvalKernel(const __grid_constant__ DeviceProgram P, const ProblemConst* __restrict__ pcs, const double* __restrict__ x,
           double* __restrict__ g_out, double* __restrict__ vals_out, int* __restrict__ status, int n_problems) {
  extern __shared__ uint4 sm4[];
  unsigned char* blob = reinterpret_cast<unsigned char*>(sm4);
  const int lane = LANES == 1 ? 0 : threadIdx.x / kThreads;
  const int tid = LANES == 1 ? threadIdx.x : threadIdx.x % kThreads;
  const int lane_doubles = P.blob.n_node_values + 6 * P.blob.n_polys + P.g_capacity;
  double* nodes = reinterpret_cast<double*>(blob + P.blob.bytes) + lane * lane_doubles;
  double* coef = nodes + P.blob.n_node_values;
  double* G = coef + 6 * P.blob.n_polys;
  __shared__ ProblemConst pc_all[LANES];
  ProblemConst& pc_s = pc_all[lane];

  // class tables: one 16-byte-vector copy per CTA, reused for every problem this CTA evaluates
  {
    const uint4* src = reinterpret_cast<const uint4*>(P.blob_data);
    for (int i = threadIdx.x; i < P.blob.bytes / 16; i += kThreads * LANES) sm4[i] = __ldg(src + i);
  }
  __syncthreads();
  ItemCtx cx;
  cx.T = MakeTables(blob, P.blob);
  cx.nodes = nodes; cx.coef = coef; cx.G = G; cx.pc = &pc_s;
  const double* Bs = cx.T.B;

  for (int prob = blockIdx.x * LANES + lane; prob < n_problems; prob += gridDim.x * LANES) {
    const double* xp = x + static_cast<size_t>(prob) * P.n_vars;
    cx.g = (WHAT & TOWR_B200_EVAL_G) ? g_out + static_cast<size_t>(prob) * P.n_rows : nullptr;
    double* vp = (WHAT & TOWR_B200_EVAL_JAC) ? vals_out + static_cast<size_t>(prob) * P.nnz : nullptr;
    int bad = 0;

    // ---- stage N: node values from x (values that are not optimised are pinned to 0)
    if (tid == 0) pc_s = pcs[prob];
    for (int i = tid; i < P.blob.n_node_values; i += kThreads) {
      const unsigned src = cx.T.node_src[i];
      const double v = src != 0xFFFFu ? __ldg(xp + src) : 0.0;
      bad |= NonFinite(v);
      nodes[i] = v;
    }
    LaneSync(lane);

    // ---- stage C: c = -(3(p0-p1) + T(2 v0 + v1))/T^2, d = (2(p0-p1) + T(v0+v1))/T^3   (polynomial.cc:97-104)
    for (int i = tid; i < 3 * P.blob.n_polys; i += kThreads) {
      const int poly = i / 3, d = i - 3 * poly;
      const uint32_t pr = cx.T.polys[poly];
      const int n0 = pr & 0xFFFF;
      const double* du = cx.T.durs + 3 * (pr >> 16);
      const double T = du[0], T2 = du[1], T3 = du[2];
      const double p0 = nodes[n0 * 6 + d], v0 = nodes[n0 * 6 + 3 + d];
      const double p1 = nodes[n0 * 6 + 6 + d], v1 = nodes[n0 * 6 + 9 + d];
      coef[poly * 6 + d] = -(3.0 * (p0 - p1) + T * (2.0 * v0 + v1)) / T2;
      coef[poly * 6 + 3 + d] = (2.0 * (p0 - p1) + T * (v0 + v1)) / T3;
    }
    LaneSync(lane);

    for (int s = 0; s < P.n_segments; ++s) {
      const SegmentRec sg = P.segments[s];
      // ---- stage A
      if (tid == 0) { G[kGOne] = 1.0; G[kGMass] = pc_s.mass; }
      for (int i = tid; i < sg.n_items; i += kThreads) RunItem(cx, cx.T.items[sg.item0 + i]);
      LaneSync(lane);
      // ---- stage B: value = G[g] * B[b]; one 16-byte descriptor load feeds four coalesced 8-byte stores
      if (WHAT & TOWR_B200_EVAL_JAC) {
        const uint4* __restrict__ ent = P.entries + sg.desc0 + tid;
        double* __restrict__ out = vp + sg.entry0 + tid;
        const int n = sg.n_entries;
        const int chunks = (n + kEntryChunk - 1) / kEntryChunk;
        const char* Gb = reinterpret_cast<const char*>(G);
        const char* Bb = reinterpret_cast<const char*>(Bs);
        auto val = [&](uint32_t d) {     // descriptors hold byte offsets: G in the low half, B in the high half
          return *reinterpret_cast<const double*>(Gb + (d & 0xFFFFu)) * *reinterpret_cast<const double*>(Bb + (d >> 16));
        };
        uint4 w = __ldg(ent);
        for (int k = 0; k + 1 < chunks; ++k) {          // full chunks: no bounds checks
          const uint4 cur = w;
          w = __ldg(ent + (k + 1) * kEntryThreads);     // prefetch the next chunk's descriptors
          double* o = out + k * kEntryChunk;
          const double v0 = val(cur.x), v1 = val(cur.y), v2 = val(cur.z), v3 = val(cur.w);
          o[0] = v0;
          o[kEntryThreads] = v1;
          o[2 * kEntryThreads] = v2;
          o[3 * kEntryThreads] = v3;
        }
        if (chunks > 0) {                               // last chunk: may be partial
          double* o = out + (chunks - 1) * kEntryChunk;
          if (w.x != kNoEntry) o[0] = val(w.x);
          if (w.y != kNoEntry) o[kEntryThreads] = val(w.y);
          if (w.z != kNoEntry) o[2 * kEntryThreads] = val(w.z);
          if (w.w != kNoEntry) o[3 * kEntryThreads] = val(w.w);
        }
        // a non-finite Jacobian value needs a non-finite G slot (B is finite): check the slots, not the entries
        for (int i = tid; i < sg.g_size; i += kThreads) bad |= NonFinite(G[i]);
      }
      LaneSync(lane);
    }
    if (bad) atomicOr(status + prob, 2);
  }
}

size_t EvalSharedBytes(const DeviceProgram& P, int lanes) {
  return static_cast<size_t>(P.blob.bytes) + sizeof(double) * lanes *
         (static_cast<size_t>(P.blob.n_node_values) + 6 * static_cast<size_t>(P.blob.n_polys) + P.g_capacity);
}

template <int WHAT, int LANES>
static cudaError_t LaunchT(const DeviceProgram& P, const ProblemConst* pcs, const double* x, double* g, double* vals,
                           int* status, int n_problems, int max_ctas, cudaStream_t stream) {
  const size_t smem = EvalSharedBytes(P, LANES);
  cudaError_t err = cudaFuncSetAttribute(EvalKernel<WHAT, LANES>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
  if (err != cudaSuccess) return err;
  int grid = (n_problems + LANES - 1) / LANES;
  if (grid > max_ctas) grid = max_ctas;
  EvalKernel<WHAT, LANES><<<grid, kThreads * LANES, smem, stream>>>(P, pcs, x, g, vals, status, n_problems);
  return cudaGetLastError();
}

template <int LANES>
static cudaError_t LaunchL(const DeviceProgram& P, const ProblemConst* pcs, int what, const double* x, double* g, double* vals,
                           int* status, int n_problems, int max_ctas, cudaStream_t stream) {
  switch (what) {
    case TOWR_B200_EVAL_G: return LaunchT<1, LANES>(P, pcs, x, g, vals, status, n_problems, max_ctas, stream);
    case TOWR_B200_EVAL_JAC: return LaunchT<2, LANES>(P, pcs, x, g, vals, status, n_problems, max_ctas, stream);
    case TOWR_B200_EVAL_BOTH: return LaunchT<3, LANES>(P, pcs, x, g, vals, status, n_problems, max_ctas, stream);
  }
  return cudaErrorInvalidValue;
}

cudaError_t LaunchEval(const DeviceProgram& P, const ProblemConst* pcs, int what, const double* x, double* g,
                       double* vals, int* status, int n_problems, int lanes, int max_ctas, cudaStream_t stream) {
  if (n_problems <= 0) return cudaSuccess;
  switch (lanes) {
    case 1: return LaunchL<1>(P, pcs, what, x, g, vals, status, n_problems, max_ctas, stream);
    case 2: return LaunchL<2>(P, pcs, what, x, g, vals, status, n_problems, max_ctas, stream);
    case 4: return LaunchL<4>(P, pcs, what, x, g, vals, status, n_problems, max_ctas, stream);
  }
  return cudaErrorInvalidValue;
}

int EvalOccupancy(const DeviceProgram& P, int lanes) {
  int n = 0;
  const size_t smem = EvalSharedBytes(P, lanes);
  cudaError_t e = cudaErrorInvalidValue;
#define TB_OCC(L)                                                                                                   \
  if (cudaFuncSetAttribute(EvalKernel<3, L>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)) != \
      cudaSuccess) { cudaGetLastError(); return 0; }                                                                \
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, EvalKernel<3, L>, kThreads * L, smem);
  if (lanes == 1) { TB_OCC(1) } else if (lanes == 2) { TB_OCC(2) } else if (lanes == 4) { TB_OCC(4) }
#undef TB_OCC
  if (e != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

}  // namespace tb
