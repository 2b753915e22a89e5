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

constexpr int kThreads = 256;
constexpr int kMinCtasPerSm = 4;   // caps registers at 64: the long dynamic item spills to local memory, stage B does not

template <int WHAT>
__global__ void __launch_bounds__(kThreads, kMinCtasPerSm)
EvalKernel(const DeviceProgram P, const ProblemConst* __restrict__ pcs, const double* __restrict__ x,
           double* __restrict__ g_out, double* __restrict__ vals_out, int* __restrict__ status, int n_problems) {
  extern __shared__ double sm[];
  double* nodes = sm;
  double* coef = nodes + P.n_node_values;
  double* G = coef + 6 * P.n_polys;
  __shared__ ProblemConst pc_s;
  const int tid = threadIdx.x;

  for (int prob = blockIdx.x; prob < n_problems; prob += gridDim.x) {
    const double* xp = x + static_cast<size_t>(prob) * P.n_vars;
    double* gp = (WHAT & TOWR_B200_EVAL_G) ? g_out + static_cast<size_t>(prob) * P.n_rows : nullptr;
    double* vp = (WHAT & TOWR_B200_EVAL_JAC) ? vals_out + static_cast<size_t>(prob) * P.nnz : nullptr;
    int bad = 0;

    // ---- stage N: node values from x (values that are not optimised are pinned to 0)
    if (tid == 0) pc_s = pcs[prob];
    for (int i = tid; i < P.n_node_values; i += kThreads) {
      const int src = __ldg(P.node_src + i);
      nodes[i] = src >= 0 ? __ldg(xp + src) : 0.0;
    }
    __syncthreads();

    // ---- stage C: c = -(3(p0-p1) + T(2 v0 + v1))/T^2, d = (2(p0-p1) + T(v0+v1))/T^3   (polynomial.cc:97-104)
    for (int i = tid; i < 3 * P.n_polys; i += kThreads) {
      const int poly = i / 3, d = i - 3 * poly;
      const PolyRec* pr = P.polys + poly;
      const int n0 = __ldg(&pr->node0);
      const double T = __ldg(&pr->T), T2 = __ldg(&pr->T2), T3 = __ldg(&pr->T3);
      const double p0 = nodes[n0 * 6 + d], v0 = nodes[n0 * 6 + 3 + d];
      const double p1 = nodes[n0 * 6 + 6 + d], v1 = nodes[n0 * 6 + 9 + d];
      coef[poly * 6 + d] = -(3.0 * (p0 - p1) + T * (2.0 * v0 + v1)) / T2;
      coef[poly * 6 + 3 + d] = (2.0 * (p0 - p1) + T * (v0 + v1)) / T3;
    }
    __syncthreads();

    ItemCtx cx;
    cx.P = &P; cx.nodes = nodes; cx.coef = coef; cx.G = G; cx.g = gp; cx.pc = &pc_s;

    for (int s = 0; s < P.n_segments; ++s) {
      const SegmentRec sg = P.segments[s];
      // ---- stage A
      if (tid == 0) { G[kGOne] = 1.0; G[kGMass] = pc_s.mass; }
      for (int i = tid; i < sg.n_items; i += kThreads) RunItem(cx, P.items[sg.item0 + i]);
      __syncthreads();
      // ---- stage B
      if (WHAT & TOWR_B200_EVAL_JAC) {
        const EntryRec* __restrict__ ent = P.entries + sg.entry0;
        double* __restrict__ out = vp + sg.entry0;
        for (int e = tid; e < sg.n_entries; e += kThreads) {
          const uint2 w = __ldg(reinterpret_cast<const uint2*>(ent + e));
          const int gi = w.x & 0xFFFF, gs = (w.x >> 16) & 0xFFF, nt = ((w.x >> 28) & 3) + 1;
          const int bi = w.y & 0x3FFFFFFF, bs = w.y >> 31;
          double v = G[gi] * __ldg(P.B + bi);
          for (int t = 1; t < nt; ++t) v += G[gi + t * gs] * __ldg(P.B + bi + t * bs);
          bad |= !isfinite(v);
          out[e] = v;
        }
      }
      __syncthreads();
    }
    if (bad) atomicOr(status + prob, 2);
  }
}

size_t EvalSharedBytes(const DeviceProgram& P) {
  return sizeof(double) * (static_cast<size_t>(P.n_node_values) + 6 * static_cast<size_t>(P.n_polys) + P.g_capacity);
}

cudaError_t LaunchEval(const DeviceProgram& P, const ProblemConst* pcs, int what, const double* x, double* g,
                       double* vals, int* status, int n_problems, int max_ctas, cudaStream_t stream) {
  if (n_problems <= 0) return cudaSuccess;
  const size_t smem = EvalSharedBytes(P);
  const int grid = n_problems < max_ctas ? n_problems : max_ctas;
  cudaError_t err = cudaSuccess;
#define TB_LAUNCH(W)                                                                                         \
  err = cudaFuncSetAttribute(EvalKernel<W>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)); \
  if (err != cudaSuccess) return err;                                                                        \
  EvalKernel<W><<<grid, kThreads, smem, stream>>>(P, pcs, x, g, vals, status, n_problems);
  switch (what) {
    case TOWR_B200_EVAL_G: TB_LAUNCH(1) break;
    case TOWR_B200_EVAL_JAC: TB_LAUNCH(2) break;
    case TOWR_B200_EVAL_BOTH: TB_LAUNCH(3) break;
    default: return cudaErrorInvalidValue;
  }
#undef TB_LAUNCH
  return cudaGetLastError();
}

int EvalOccupancy(const DeviceProgram& P) {
  int n = 0;
  const size_t smem = EvalSharedBytes(P);
  cudaFuncSetAttribute(EvalKernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, EvalKernel<3>, kThreads, smem) != cudaSuccess) return 0;
  return n;
}

}  // namespace tb
