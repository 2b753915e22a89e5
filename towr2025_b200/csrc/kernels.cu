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

#ifndef TB_THREADS
#define TB_THREADS kEntryThreads
#endif
constexpr int kThreads = TB_THREADS;   // threads per problem lane
#ifndef TB_MIN_CTAS
#define TB_MIN_CTAS 1
#endif
constexpr int kMinCtasPerSm = TB_MIN_CTAS;   // caps registers at 64: the long dynamic item spills to local memory, stage B does not

// Phase clocks of CTA 0 (timing build only: make timing; tools/phase_timing.py)
#ifdef TB_TIMING
__device__ unsigned long long tb_phase_clk[64];
#define TB_MARK(k) TB_MARK_T(k, 0)
#define TB_MARK_T(k, who)                                                      \
  if (blockIdx.x == 0 && threadIdx.x == (who)) {                               \
    const long long now_ = clock64();                                          \
    tb_phase_clk[(k)] += static_cast<unsigned long long>(now_ - mark_);        \
    mark_ = now_;                                                              \
  }
#define TB_MARK_INIT long long mark_ = clock64();
#else
#define TB_MARK(k)
#define TB_MARK_T(k, who)
#define TB_MARK_INIT
#endif

__device__ __forceinline__ bool NonFinite(double v) {
  return (__double2hiint(v) & 0x7ff00000) == 0x7ff00000;
}

// Stage B of one segment.  Thread t takes entries t, t + NT, ...: a warp's descriptor load and its P stores per
// entry are coalesced.  Descriptors are fetched kDescBatch at a time, one batch ahead of their use (the first batch
// before the barrier that ends stage A), so the L2 latency of the descriptor stream is off the critical path.
constexpr int kDescBatch = 4;

template <int NT>
__device__ __forceinline__ void LoadDescriptors(const uint32_t* __restrict__ ent, int left, uint32_t (&d)[kDescBatch]) {
#pragma unroll
  for (int u = 0; u < kDescBatch; ++u) d[u] = left > u * NT ? __ldg(ent + u * NT) : kNoEntry;
}

// FULL: all P problem lanes of the group exist (no per-store predicate).
// nf[p] accumulates value * 0 per problem lane: NaN exactly when a stored value is not finite (status bit 2).
template <int P, bool FULL, int NT, bool GG = false>
__device__ __forceinline__ void StageB(const uint32_t* __restrict__ ent, int left, uint32_t (&d)[kDescBatch],
                                       const double* G, const double* Bs, double* __restrict__ o, int nnz, int np,
                                       double (&nf)[P]) {
  constexpr int U = kDescBatch;
  const char* Gb = reinterpret_cast<const char*>(G);
  const char* Bb = reinterpret_cast<const char*>(Bs);
  auto emit = [&](uint32_t w, double* op) {              // descriptor: G byte offset (low 16), B byte offset (high 16)
    double bv = 0.0;
    double bw[P];                                         // GG entries: the second factor is a per-problem G slot
    bool gg = false;
    if constexpr (GG) gg = (w & 0x10000u) != 0;
    if (!gg) {
      bv = *reinterpret_cast<const double*>(Bb + (w >> 16));
    } else {
      const double* wa = reinterpret_cast<const double*>(Gb + ((w >> 16) & 0xFFF8u) * P);
#pragma unroll
      for (int p = 0; p < P; ++p) bw[p] = wa[p];
    }
    const char* ga = Gb + (w & 0xFFFFu) * P;
    double gv[P];
    if constexpr (P == 1) {
      gv[0] = *reinterpret_cast<const double*>(ga);
    } else {
#pragma unroll
      for (int p = 0; p < P; p += 2) {
        const double2 t = *reinterpret_cast<const double2*>(ga + 8 * p);
        gv[p] = t.x;
        gv[p + 1] = t.y;
      }
    }
#pragma unroll
    for (int p = 0; p < P; ++p) {
      const double v = gv[p] * ((GG && gg) ? bw[p] : bv);
      nf[p] = fma(v, 0.0, nf[p]);
      if (FULL || p < np) op[static_cast<size_t>(p) * nnz] = v;
    }
  };
  while (left > 0) {
    const int left_next = left - U * NT;
    uint32_t dn[U];
    if (left_next > 0) LoadDescriptors<NT>(ent + U * NT, left_next, dn);
    if (left > (U - 1) * NT) {
#pragma unroll
      for (int u = 0; u < U; ++u) emit(d[u], o + u * NT);
    } else {
#pragma unroll
      for (int u = 0; u < U - 1; ++u)
        if (left > u * NT) emit(d[u], o + u * NT);
    }
    if (left_next <= 0) break;
#pragma unroll
    for (int u = 0; u < U; ++u) d[u] = dn[u];
    left = left_next;
    ent += U * NT;
    o += U * NT;
  }
}

// 8-byte asynchronous global -> shared copy; src_bytes = 0 writes zeros (cp.async zero-fill)
__device__ __forceinline__ void CpAsync8(void* smem_dst, const void* gsrc, int src_bytes) {
  const unsigned d = static_cast<unsigned>(__cvta_generic_to_shared(smem_dst));
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(d), "l"(gsrc), "r"(src_bytes) : "memory");
}

// P problems are evaluated in lock step by one CTA of 256*P threads: the class tables are held once per SM, every
// stage spreads (item, problem) pairs over all threads, and in stage B one descriptor load and one B load feed the
// P problems.  All per-problem shared arrays are interleaved (element i of problem p at [i*P + p]).
// RT: structure classes with optimised phase durations (stage D below; items.cuh Lockstep<P, true>).
template <int WHAT, int P, bool RT>
__global__ void __launch_bounds__(kThreads * P, 1)
EvalKernel(const __grid_constant__ DeviceProgram Pg, const ProblemConst* __restrict__ pcs, const double* __restrict__ x,
           double* __restrict__ g_out, double* __restrict__ vals_out, int* __restrict__ status, int n_problems) {
  constexpr int NT = kThreads * P;
  extern __shared__ uint4 sm4[];
  unsigned char* blob = reinterpret_cast<unsigned char*>(sm4);
  const int nnv = Pg.blob.n_node_values, npoly = Pg.blob.n_polys;
  double* nodes = reinterpret_cast<double*>(blob + Pg.blob.bytes);
  double* coef = nodes + nnv * P;
  double* rtd = coef + 6 * npoly * P;                   // RT: 5 per polynomial, then the phase durations
  double* pdur = rtd + (RT ? 5 * npoly * P : 0);
  double* G = pdur + (RT ? Pg.blob.n_pdur * P : 0);
  __shared__ ProblemConst pc_all[P];
  __shared__ int bad_all[P];
  const int tid = threadIdx.x;

  // class tables: one 16-byte-vector copy per CTA, reused for every problem group this CTA evaluates
  {
    const uint4* src = reinterpret_cast<const uint4*>(Pg.blob_data);
    for (int i = tid; i < Pg.blob.bytes / 16; i += NT) sm4[i] = __ldg(src + i);
  }
  __syncthreads();
  TB_MARK_INIT
  ItemCtx<Lockstep<P, RT>> cx;
  cx.n_ee = Pg.n_ee;
  cx.T = MakeTables(blob, Pg.blob);
  const double* Bs = cx.T.B;
  const int pl = tid % P;                               // problem lane of this thread in stages N, C, A

  // stage N: node values from x (values that are not optimised are pinned to 0), fetched asynchronously: for the first
  // group here, for every later group while the previous group's last stage B runs (nodes[] is free by then)
  auto fetch_group = [&](int base_n) {
    const double* xn = x + static_cast<size_t>(min(base_n + pl, n_problems - 1)) * Pg.n_vars;
    for (int i = tid; i < nnv * P; i += NT) {
      const unsigned src = cx.T.node_src[i / P];
      CpAsync8(nodes + i, xn + (src != 0xFFFFu ? src : 0u), src != 0xFFFFu ? 8 : 0);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    if (tid < P) pc_all[tid] = pcs[min(base_n + tid, n_problems - 1)];
  };
  if (blockIdx.x * P < n_problems) fetch_group(blockIdx.x * P);

  for (int base = blockIdx.x * P; base < n_problems; base += gridDim.x * P) {
    // the last group may hold fewer than P problems: the spare lanes recompute the last problem (same values to
    // the same addresses) so that no stage needs a bounds test
    const int prob = min(base + pl, n_problems - 1);
    cx.nodes = nodes + pl;
    cx.coef = coef + pl;
    cx.G = G + pl;
    cx.rtd = rtd + pl;
    cx.pdur = pdur + pl;
    cx.pc = &pc_all[pl];
    cx.g = (WHAT & TOWR_B200_EVAL_G) ? g_out + static_cast<size_t>(prob) * Pg.n_rows : nullptr;
    double nf[P];
#pragma unroll
    for (int p = 0; p < P; ++p) nf[p] = 0.0;
    double nfx = 0.0;                                     // non-finite x of problem lane pl
    if (tid < P) bad_all[tid] = 0;
    asm volatile("cp.async.wait_all;" ::: "memory");
    TB_MARK(0)
    __syncthreads();
    TB_MARK(1)

    // ---- stage D (optimised durations): phase durations from x (phase_durations.cc:78-100: the last one is
    // T - sum of the others), then every polynomial's duration T, T^2, T^3, 1/T^2, 1/T^3
    // (ConvertPhaseToPolyDurations, nodes_variables_phase_based.cc:75-86; base splines keep their class constants)
    if constexpr (RT) {
      const double* xg = x + static_cast<size_t>(prob) * Pg.n_vars;
      if (tid < Pg.blob.n_rt_feet * P) {
        const RtFoot ft = cx.T.rt_feet[tid / P];
        double sum = 0.0;
        for (int i = 0; i < ft.n_phases - 1; ++i) {
          const double d = __ldg(xg + ft.col0 + i);
          nfx = fma(d, 0.0, nfx);
          pdur[(ft.phase0 + i) * P + pl] = d;
          sum += d;
        }
        pdur[(ft.phase0 + ft.n_phases - 1) * P + pl] = Pg.total_time - sum;
      }
      __syncthreads();
      for (int i = tid; i < npoly * P; i += NT) {
        const int poly = i / P;
        const uint32_t pm = cx.T.poly_meta[poly];
        double* o = rtd + poly * 5 * P + pl;
        if (pm == kPolyFixed) {
          const int di = cx.T.polys[poly] >> 16;
          o[0] = cx.T.durs[3 * di]; o[P] = cx.T.durs[3 * di + 1]; o[2 * P] = cx.T.durs[3 * di + 2];
          o[3 * P] = cx.T.rdurs[2 * di]; o[4 * P] = cx.T.rdurs[2 * di + 1];
        } else {
          const RtFoot ft = cx.T.rt_feet[pm >> 24];
          const double Td = pdur[(ft.phase0 + (pm & 0xFF)) * P + pl] / static_cast<double>((pm >> 8) & 0xFF);
          const double T2 = Td * Td, T3 = Cube(Td);
          o[0] = Td; o[P] = T2; o[2 * P] = T3; o[3 * P] = 1.0 / T2; o[4 * P] = 1.0 / T3;
        }
      }
      __syncthreads();
    }

    // ---- stage C: c = -(3(p0-p1) + T(2 v0 + v1))/T^2, d = (2(p0-p1) + T(v0+v1))/T^3   (polynomial.cc:97-104)
    for (int i = tid; i < 3 * npoly * P; i += NT) {
      const int pd = i / P;                             // (poly, dim) pair; i % P == pl
      const int poly = pd / 3, d = pd - 3 * poly;
      const uint32_t pr = cx.T.polys[poly];
      const int n0 = pr & 0xFFFF;
      double T, T2, T3, r2, r3;
      if constexpr (RT) {
        const double* o = rtd + poly * 5 * P + pl;
        T = o[0]; T2 = o[P]; T3 = o[2 * P]; r2 = o[3 * P]; r3 = o[4 * P];
      } else {
        const double* du = cx.T.durs + 3 * (pr >> 16);
        const double* rd = cx.T.rdurs + 2 * (pr >> 16);
        T = du[0]; T2 = du[1]; T3 = du[2]; r2 = rd[0]; r3 = rd[1];
      }
      const double* nd = nodes + pl;
      const double p0 = nd[(n0 * 6 + d) * P], v0 = nd[(n0 * 6 + 3 + d) * P];
      const double p1 = nd[(n0 * 6 + 6 + d) * P], v1 = nd[(n0 * 6 + 9 + d) * P];
      nfx = fma(p0, 0.0, fma(v0, 0.0, fma(p1, 0.0, fma(v1, 0.0, nfx))));   // NaN iff a node value is not finite
      coef[(poly * 6 + d) * P + pl] = ExactDiv(-(3.0 * (p0 - p1) + T * (2.0 * v0 + v1)), T2, r2);
      coef[(poly * 6 + 3 + d) * P + pl] = ExactDiv(2.0 * (p0 - p1) + T * (v0 + v1), T3, r3);
    }
    TB_MARK(2)
    __syncthreads();
    TB_MARK(3)

    for (int s = 0; s < Pg.n_segments; ++s) {
      const SegmentRec sg = cx.T.segments[s];
      // ---- stage A, level 0 then level 1: thread -> (item, problem lane)
      if (tid < P) { G[kGOne * P + tid] = 1.0; G[kGMass * P + tid] = pc_all[tid].mass; G[kGMinusOne * P + tid] = -1.0; }
      for (int i = tid; i < sg.n_items0 * P; i += NT) RunItem<Lockstep<P, RT>>(cx, cx.T.items[sg.item0 + i / P]);
      TB_MARK(8 + 6 * s)
      if (sg.n_items > sg.n_items0) {
        __syncthreads();
        TB_MARK(9 + 6 * s)
        for (int i = tid; i < (sg.n_items - sg.n_items0) * P; i += NT)
          RunItem<Lockstep<P, RT>>(cx, cx.T.items[sg.item0 + sg.n_items0 + i / P]);
        TB_MARK(10 + 6 * s)
      }
      // first descriptor batch of stage B: issued before the barrier, consumed after it
      const uint32_t* ent = Pg.entries + sg.entry0 + tid;
      uint32_t d0[kDescBatch];
      if (WHAT & TOWR_B200_EVAL_JAC) LoadDescriptors<NT>(ent, sg.n_entries - tid, d0);
      __syncthreads();
      TB_MARK(11 + 6 * s)
      // all items of this group are done with nodes[] / pc_all after the last segment's stage A: fetch the next group
      if (s == Pg.n_segments - 1 && base + gridDim.x * P < n_problems) fetch_group(base + gridDim.x * P);
      // ---- stage B: value_p = G_p[g] * B[b] for the P problems: one descriptor and one B load feed P coalesced stores
      if (WHAT & TOWR_B200_EVAL_JAC) {
        double* o0 = vals_out + static_cast<size_t>(base) * Pg.nnz + sg.entry0 + tid;
        if (n_problems - base >= P)
          StageB<P, true, NT, RT>(ent, sg.n_entries - tid, d0, G, Bs, o0, Pg.nnz, P, nf);
        else
          StageB<P, false, NT, RT>(ent, sg.n_entries - tid, d0, G, Bs, o0, Pg.nnz, n_problems - base, nf);
      }
      TB_MARK(12 + 6 * s)
      __syncthreads();
      TB_MARK(13 + 6 * s)
    }
    if (nfx != nfx) atomicOr(&bad_all[pl], 1);
#pragma unroll
    for (int p = 0; p < P; ++p)
      if (nf[p] != nf[p]) atomicOr(&bad_all[p], 1);
    __syncthreads();
    if (tid < P && bad_all[tid] && base + tid < n_problems) atomicOr(status + base + tid, 2);
  }
}

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
        uint32_t d0[kDescBatch];
        if (WHAT & TOWR_B200_EVAL_JAC) LoadDescriptors<NB>(ent, sg.n_entries - ct, d0);
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

size_t EvalSharedBytes(const DeviceProgram& P, int lanes) {
  const size_t rt = P.blob.rt ? 5 * static_cast<size_t>(P.blob.n_polys) + P.blob.n_pdur : 0;
  return static_cast<size_t>(P.blob.bytes) + sizeof(double) * lanes *
         (static_cast<size_t>(P.blob.n_node_values) + 6 * static_cast<size_t>(P.blob.n_polys) + rt + P.g_capacity);
}

template <int WHAT, int LANES>
static cudaError_t LaunchT(const DeviceProgram& P, const ProblemConst* pcs, const double* x, double* g, double* vals,
                           int* status, int n_problems, int max_ctas, cudaStream_t stream) {
  const size_t smem = EvalSharedBytes(P, LANES);
  auto kern = P.blob.rt ? EvalKernel<WHAT, LANES, true> : EvalKernel<WHAT, LANES, false>;
  cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
  if (err != cudaSuccess) return err;
  int grid = (n_problems + LANES - 1) / LANES;
  if (grid > max_ctas) grid = max_ctas;
  kern<<<grid, kThreads * LANES, smem, stream>>>(P, pcs, x, g, vals, status, n_problems);
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
  {                                                                                                                 \
    auto kern = P.blob.rt ? EvalKernel<3, L, true> : EvalKernel<3, L, false>;                                       \
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)) !=          \
        cudaSuccess) { cudaGetLastError(); return 0; }                                                              \
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kern, kThreads * L, smem);                                \
  }
  if (lanes == 1) { TB_OCC(1) } else if (lanes == 2) { TB_OCC(2) } else if (lanes == 4) { TB_OCC(4) }
#undef TB_OCC
  if (e != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

}  // namespace tb

#ifdef TB_TIMING
extern "C" int towr_b200_debug_phase_clocks(unsigned long long* out, int n) {
  unsigned long long h[64];
  if (cudaMemcpyFromSymbol(h, tb::tb_phase_clk, sizeof(h)) != cudaSuccess) return -1;
  for (int i = 0; i < n && i < 64; ++i) out[i] = h[i];
  unsigned long long z[64] = {};
  cudaMemcpyToSymbol(tb::tb_phase_clk, z, sizeof(z));
  return 0;
}
#endif
