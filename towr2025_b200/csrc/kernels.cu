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
#define TB_MIN_CTAS 1
#endif
constexpr int kMinCtasPerSm = TB_MIN_CTAS;   // caps registers at 64: the long dynamic item spills to local memory, stage B does not

__device__ __forceinline__ bool NonFinite(double v) {
  return (__double2hiint(v) & 0x7ff00000) == 0x7ff00000;
}

// P problems are evaluated in lock step by one CTA of 256*P threads: the class tables are held once per SM, every
// stage spreads (item, problem) pairs over all threads, and in stage B one descriptor load and one B load feed the
// P problems.  All per-problem shared arrays are interleaved (element i of problem p at [i*P + p]).
template <int WHAT, int P>
__global__ void __launch_bounds__(kThreads * P, 1)
EvalKernel(const __grid_constant__ DeviceProgram Pg, const ProblemConst* __restrict__ pcs, const double* __restrict__ x,
           double* __restrict__ g_out, double* __restrict__ vals_out, int* __restrict__ status, int n_problems) {
  constexpr int NT = kThreads * P;
  extern __shared__ uint4 sm4[];
  unsigned char* blob = reinterpret_cast<unsigned char*>(sm4);
  const int nnv = Pg.blob.n_node_values, npoly = Pg.blob.n_polys;
  double* nodes = reinterpret_cast<double*>(blob + Pg.blob.bytes);
  double* coef = nodes + nnv * P;
  double* G = coef + 6 * npoly * P;
  __shared__ ProblemConst pc_all[P];
  __shared__ int bad_all[P];
  const int tid = threadIdx.x;

  // class tables: one 16-byte-vector copy per CTA, reused for every problem group this CTA evaluates
  {
    const uint4* src = reinterpret_cast<const uint4*>(Pg.blob_data);
    for (int i = tid; i < Pg.blob.bytes / 16; i += NT) sm4[i] = __ldg(src + i);
  }
  __syncthreads();
  ItemCtx<P> cx;
  cx.T = MakeTables(blob, Pg.blob);
  const double* Bs = cx.T.B;
  const int pl = tid % P;                               // problem lane of this thread in stages N, C, A

  for (int base = blockIdx.x * P; base < n_problems; base += gridDim.x * P) {
    // the last group may hold fewer than P problems: the spare lanes recompute the last problem (same values to
    // the same addresses) so that no stage needs a bounds test
    const int prob = min(base + pl, n_problems - 1);
    const double* xp = x + static_cast<size_t>(prob) * Pg.n_vars;
    cx.nodes = nodes + pl;
    cx.coef = coef + pl;
    cx.G = G + pl;
    cx.pc = &pc_all[pl];
    cx.g = (WHAT & TOWR_B200_EVAL_G) ? g_out + static_cast<size_t>(prob) * Pg.n_rows : nullptr;
    int bad = 0;

    // ---- stage N: node values from x (values that are not optimised are pinned to 0)
    if (tid < P) { pc_all[tid] = pcs[min(base + tid, n_problems - 1)]; bad_all[tid] = 0; }
    for (int i = tid; i < nnv * P; i += NT) {
      const unsigned src = cx.T.node_src[i / P];
      const double v = src != 0xFFFFu ? __ldg(xp + src) : 0.0;
      bad |= NonFinite(v);
      nodes[i] = v;
    }
    __syncthreads();

    // ---- stage C: c = -(3(p0-p1) + T(2 v0 + v1))/T^2, d = (2(p0-p1) + T(v0+v1))/T^3   (polynomial.cc:97-104)
    for (int i = tid; i < 3 * npoly * P; i += NT) {
      const int pd = i / P;                             // (poly, dim) pair; i % P == pl
      const int poly = pd / 3, d = pd - 3 * poly;
      const uint32_t pr = cx.T.polys[poly];
      const int n0 = pr & 0xFFFF;
      const double* du = cx.T.durs + 3 * (pr >> 16);
      const double T = du[0], T2 = du[1], T3 = du[2];
      const double* nd = nodes + pl;
      const double p0 = nd[(n0 * 6 + d) * P], v0 = nd[(n0 * 6 + 3 + d) * P];
      const double p1 = nd[(n0 * 6 + 6 + d) * P], v1 = nd[(n0 * 6 + 9 + d) * P];
      coef[(poly * 6 + d) * P + pl] = -(3.0 * (p0 - p1) + T * (2.0 * v0 + v1)) / T2;
      coef[(poly * 6 + 3 + d) * P + pl] = (2.0 * (p0 - p1) + T * (v0 + v1)) / T3;
    }
    __syncthreads();

    for (int s = 0; s < Pg.n_segments; ++s) {
      const SegmentRec sg = Pg.segments[s];
      // ---- stage A, level 0 then level 1: thread -> (item, problem lane)
      if (tid < P) { G[kGOne * P + tid] = 1.0; G[kGMass * P + tid] = pc_all[tid].mass; }
      for (int i = tid; i < sg.n_items0 * P; i += NT) RunItem<P>(cx, cx.T.items[sg.item0 + i / P]);
      __syncthreads();
      if (sg.n_items > sg.n_items0) {
        for (int i = tid; i < (sg.n_items - sg.n_items0) * P; i += NT)
          RunItem<P>(cx, cx.T.items[sg.item0 + sg.n_items0 + i / P]);
        __syncthreads();
      }
      // ---- stage B: value_p = G_p[g] * B[b] for the P problems; one 16-byte descriptor load feeds 4*P coalesced stores
      if (WHAT & TOWR_B200_EVAL_JAC) {
        const int tt = tid % kThreads, q = tid / kThreads;            // q: which chunks this 256-thread group takes
        const uint4* __restrict__ ent = Pg.entries + sg.desc0 + tt;
        const int chunks = (sg.n_entries + kEntryChunk - 1) / kEntryChunk;
        const char* Gb = reinterpret_cast<const char*>(G);
        const char* Bb = reinterpret_cast<const char*>(Bs);
        double* out0 = vals_out + static_cast<size_t>(base) * Pg.nnz + sg.entry0 + tt;
        const int np = min(P, n_problems - base);
        auto emit = [&](uint32_t d, double* o) {          // descriptor: G byte offset (low), B byte offset (high)
          const double bv = *reinterpret_cast<const double*>(Bb + (d >> 16));
          const double* gp = reinterpret_cast<const double*>(Gb + (d & 0xFFFFu) * P);
#pragma unroll
          for (int p = 0; p < P; ++p)
            if (p < np) o[static_cast<size_t>(p) * Pg.nnz] = gp[p] * bv;
        };
        if (q < chunks) {
          uint4 w = __ldg(ent + q * kEntryThreads);
          for (int k = q; k < chunks; k += P) {
            const uint4 cur = w;
            if (k + P < chunks) w = __ldg(ent + (k + P) * kEntryThreads);   // prefetch this thread's next chunk
            double* o = out0 + k * kEntryChunk;
            if (k + 1 < chunks) {                         // full chunk: no bounds checks
              emit(cur.x, o);
              emit(cur.y, o + kEntryThreads);
              emit(cur.z, o + 2 * kEntryThreads);
              emit(cur.w, o + 3 * kEntryThreads);
            } else {
              if (cur.x != kNoEntry) emit(cur.x, o);
              if (cur.y != kNoEntry) emit(cur.y, o + kEntryThreads);
              if (cur.z != kNoEntry) emit(cur.z, o + 2 * kEntryThreads);
              if (cur.w != kNoEntry) emit(cur.w, o + 3 * kEntryThreads);
            }
          }
        }
        // a non-finite Jacobian value needs a non-finite G slot (B is finite): check the slots, not the entries
        for (int i = tid; i < sg.g_size * P; i += NT)
          if (NonFinite(G[i])) atomicOr(&bad_all[i % P], 1);
      }
      __syncthreads();
    }
    if (bad) atomicOr(&bad_all[pl], 1);
    __syncthreads();
    if (tid < P && bad_all[tid] && base + tid < n_problems) atomicOr(status + base + tid, 2);
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
