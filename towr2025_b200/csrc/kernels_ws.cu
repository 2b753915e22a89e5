// kernels_ws.cu — warp-specialised variant of the evaluation kernel.
#if defined(TB_TIMING)
#define TB_TIMING_HERE
#endif
#include "eval_kernel.cuh"

namespace tb {

// ---------------------------------------------------------------------------------------------------------------
// Warp-specialised variant.  The SM's global store path moves 32 B per clock (tools/micro/store_bw.cu), only 1.37x
// the per-SM share of HBM bandwidth, so the value stores of stage B have to be in flight almost all the time.
// NA producer threads run stages N, C, A into a double-buffered G; NB consumer threads run stage B (and the G
// check) on the buffer the producers filled last: stage A of the next segment - also of the next problem group -
// overlaps the stores.  Hand-off through named barriers: full[b] (producers arrive, consumers wait), empty[b]
// (consumers arrive, producers wait).
__device__ __forceinline__ void BarSync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }
__device__ __forceinline__ void BarArrive(int id, int n) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(n) : "memory"); }

template <int WHAT, int P, int NA, int NB>
__global__ void __launch_bounds__(NA + NB, 1)
EvalKernelWS(const __grid_constant__ DeviceProgram Pg, const ProblemConst* __restrict__ pcs, const double* __restrict__ x,
             double* __restrict__ g_out, double* __restrict__ vals_out, int* __restrict__ status, int n_problems) {
  enum { kBarFull = 1, kBarEmpty = 3, kBarA = 5 };       // full[0..1] = 1,2 ; empty[0..1] = 3,4
  constexpr int NALL = NA + NB;
  extern __shared__ uint4 sm4[];
  unsigned char* blob = reinterpret_cast<unsigned char*>(sm4);
  const int nnv = Pg.blob.n_node_values, npoly = Pg.blob.n_polys;
  double* nodes = reinterpret_cast<double*>(blob + Pg.blob.bytes);
  double* coef = nodes + nnv * P;
  double* Gbuf = coef + 6 * npoly * P;                   // two buffers of g_capacity * P
  const int gstride = Pg.g_capacity * P;
  __shared__ ProblemConst pc_all[P];
  const int tid = threadIdx.x;
  {
    const uint4* src = reinterpret_cast<const uint4*>(Pg.blob_data);
    for (int i = tid; i < Pg.blob.bytes / 16; i += NALL) sm4[i] = __ldg(src + i);
  }
  __syncthreads();
  TB_MARK_INIT
  const Tables T = MakeTables(blob, Pg.blob);
  const int nseg = Pg.n_segments;
  const int n_groups = (n_problems + P - 1) / P;
  const int my_groups = (n_groups - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
  const int k_total = my_groups * nseg;                  // segments this CTA hands from producers to consumers

  if (tid < NA) {
    // ------------------------------------------------------------------------------------------ producers
    ItemCtx<Lockstep<P>> cx;
    cx.T = T;
    cx.n_ee = Pg.n_ee;
    cx.rtd = nullptr;
    cx.pdur = nullptr;
    const int pl = tid % P;
    auto fetch_group = [&](int base_n) {
      const double* xn = x + static_cast<size_t>(min(base_n + pl, n_problems - 1)) * Pg.n_vars;
      for (int i = tid; i < nnv * P; i += NA) {
        const unsigned src = T.node_src[i / P];
        CpAsync8(nodes + i, xn + (src != 0xFFFFu ? src : 0u), src != 0xFFFFu ? 8 : 0);
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
      if (tid < P) pc_all[tid] = pcs[min(base_n + tid, n_problems - 1)];
    };
    if (my_groups > 0) fetch_group(blockIdx.x * P);
    int k = 0;
    for (int base = blockIdx.x * P; base < n_problems; base += gridDim.x * P) {
      const int prob = min(base + pl, n_problems - 1);
      cx.nodes = nodes + pl;
      cx.coef = coef + pl;
      cx.pc = &pc_all[pl];
      cx.g = (WHAT & TOWR_B200_EVAL_G) ? g_out + static_cast<size_t>(prob) * Pg.n_rows : nullptr;
      double nfx = 0.0;
      asm volatile("cp.async.wait_all;" ::: "memory");
      BarSync(kBarA, NA);
      TB_MARK(0)
      for (int i = tid; i < 3 * npoly * P; i += NA) {    // stage C (polynomial.cc:97-104)
        const int pd = i / P;
        const int poly = pd / 3, d = pd - 3 * poly;
        const uint32_t pr = T.polys[poly];
        const int n0 = pr & 0xFFFF;
        const double* du = T.durs + 3 * (pr >> 16);
        const double* rd = T.rdurs + 2 * (pr >> 16);
        const double Td = du[0], T2 = du[1], T3 = du[2];
        const double* nd = nodes + pl;
        const double p0 = nd[(n0 * 6 + d) * P], v0 = nd[(n0 * 6 + 3 + d) * P];
        const double p1 = nd[(n0 * 6 + 6 + d) * P], v1 = nd[(n0 * 6 + 9 + d) * P];
        nfx = fma(p0, 0.0, fma(v0, 0.0, fma(p1, 0.0, fma(v1, 0.0, nfx))));
        coef[(poly * 6 + d) * P + pl] = ExactDiv(-(3.0 * (p0 - p1) + Td * (2.0 * v0 + v1)), T2, rd[0]);
        coef[(poly * 6 + 3 + d) * P + pl] = ExactDiv(2.0 * (p0 - p1) + Td * (v0 + v1), T3, rd[1]);
      }
      BarSync(kBarA, NA);
      TB_MARK(1)
      for (int s = 0; s < nseg; ++s, ++k) {
        const SegmentRec sg = T.segments[s];
        const int buf = k & 1;
        double* G = Gbuf + buf * gstride;
        if (k >= 2) BarSync(kBarEmpty + buf, NALL);       // the consumers are done with this buffer
        TB_MARK(2)
        cx.G = G + pl;
        if (tid < P) { G[kGOne * P + tid] = 1.0; G[kGMass * P + tid] = pc_all[tid].mass; G[kGMinusOne * P + tid] = -1.0; }
        for (int i = tid; i < sg.n_items0 * P; i += NA) RunItem<Lockstep<P>>(cx, T.items[sg.item0 + i / P]);
        TB_MARK(3)
        if (sg.n_items > sg.n_items0) {
          BarSync(kBarA, NA);
          TB_MARK(4)
          for (int i = tid; i < (sg.n_items - sg.n_items0) * P; i += NA)
            RunItem<Lockstep<P>>(cx, T.items[sg.item0 + sg.n_items0 + i / P]);
          TB_MARK(5)
        }
        __threadfence_block();
        BarArrive(kBarFull + buf, NALL);
      }
      // nodes[] and pc_all are free once every producer is past its last item of this group
      BarSync(kBarA, NA);
      TB_MARK(6)
      if (base + gridDim.x * P < n_problems) fetch_group(base + gridDim.x * P);
      if (nfx != nfx && base + pl < n_problems) atomicOr(status + base + pl, 2);
    }
  } else {
    // ------------------------------------------------------------------------------------------ consumers
    const int ct = tid - NA;
    int k = 0;
    for (int base = blockIdx.x * P; base < n_problems; base += gridDim.x * P) {
      double nf[P];
#pragma unroll
      for (int p = 0; p < P; ++p) nf[p] = 0.0;
      for (int s = 0; s < nseg; ++s, ++k) {
        const SegmentRec sg = T.segments[s];
        const int buf = k & 1;
        const double* G = Gbuf + buf * gstride;
        const uint32_t* ent = Pg.entries + sg.entry0 + ct;
        uint32_t d0[DescAhead<false>::value][kDescBatch];
        if (WHAT & TOWR_B200_EVAL_JAC) LoadDescriptorRing<NB>(ent, sg.n_entries - ct, d0);
        TB_MARK_T(16, NA)
        BarSync(kBarFull + buf, NALL);
        TB_MARK_T(17, NA)
        if (WHAT & TOWR_B200_EVAL_JAC) {
          double* o0 = vals_out + static_cast<size_t>(base) * Pg.nnz + sg.entry0 + ct;
          if (n_problems - base >= P)
            StageB<P, true, NB>(ent, sg.n_entries - ct, d0, G, T.B, o0, Pg.nnz, P, nf);
          else
            StageB<P, false, NB>(ent, sg.n_entries - ct, d0, G, T.B, o0, Pg.nnz, n_problems - base, nf);
        }
        TB_MARK_T(18, NA)
        if (k + 2 < k_total) BarArrive(kBarEmpty + buf, NALL);
      }
#pragma unroll
      for (int p = 0; p < P; ++p)
        if (nf[p] != nf[p] && base + p < n_problems) atomicOr(status + base + p, 2);
    }
  }
}

size_t EvalSharedBytesWS(const DeviceProgram& P, int lanes) {
  return static_cast<size_t>(P.blob.bytes) + sizeof(double) * lanes *
         (static_cast<size_t>(P.blob.n_node_values) + 6 * static_cast<size_t>(P.blob.n_polys) + 2 * static_cast<size_t>(P.g_capacity));
}

#ifndef TB_WS_NA
#define TB_WS_NA 256
#endif
#ifndef TB_WS_NB
#define TB_WS_NB 256
#endif

template <int WHAT, int LANES>
static cudaError_t LaunchWS(const DeviceProgram& P, const ProblemConst* pcs, const double* x, double* g, double* vals,
                            int* status, int n_problems, int max_ctas, cudaStream_t stream) {
  const size_t smem = EvalSharedBytesWS(P, LANES);
  auto kern = EvalKernelWS<WHAT, LANES, TB_WS_NA, TB_WS_NB>;
  cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
  if (err != cudaSuccess) return err;
  int grid = (n_problems + LANES - 1) / LANES;
  if (grid > max_ctas) grid = max_ctas;
  kern<<<grid, TB_WS_NA + TB_WS_NB, smem, stream>>>(P, pcs, x, g, vals, status, n_problems);
  return cudaGetLastError();
}

cudaError_t LaunchEvalWS(const DeviceProgram& P, const ProblemConst* pcs, int what, const double* x, double* g,
                         double* vals, int* status, int n_problems, int lanes, int max_ctas, cudaStream_t stream) {
  if (n_problems <= 0) return cudaSuccess;
  if (lanes != 2) return cudaErrorInvalidValue;
  switch (what) {
    case TOWR_B200_EVAL_G: return LaunchWS<1, 2>(P, pcs, x, g, vals, status, n_problems, max_ctas, stream);
    case TOWR_B200_EVAL_JAC: return LaunchWS<2, 2>(P, pcs, x, g, vals, status, n_problems, max_ctas, stream);
    case TOWR_B200_EVAL_BOTH: return LaunchWS<3, 2>(P, pcs, x, g, vals, status, n_problems, max_ctas, stream);
  }
  return cudaErrorInvalidValue;
}

bool EvalFitsWS(const DeviceProgram& P, int lanes, size_t smem_optin) {
  if (lanes != 2 || P.blob.rt) return false;   // optimised durations run on the lock-step kernel
  const size_t smem = EvalSharedBytesWS(P, lanes) + 1024;   // static shared memory of the kernel
  return smem <= smem_optin;
}

}  // namespace tb

TB_PHASE_CLOCK_READER(towr_b200_debug_phase_clocks_ws)
