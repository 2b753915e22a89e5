// eval_kernel_split.cuh — the evaluation kernel with producer / consumer warp groups (sm_100a).
//
// The lock-step kernel (eval_kernel.cuh) alternates stage A (FP64 / shared-memory latency bound, no HBM traffic) and
// stage B (bound by the SM's 32 B/clk path to L2), so each unit idles while the other works.  Here one CTA runs both at
// once on different warps:
//
//   producers   256 * P threads (as many as the lock-step kernel, so a level still runs in as few rounds): stages N, D, C
//               and stage A of segment q into G buffer q & 1
//   consumers   128 * P threads: stage B of segment q - 1 from the other G buffer
//
// 768 threads (P = 2) leave 85 registers per thread, fewer than the items need; the kernel is therefore launched at 80
// registers per thread and each group resizes its share with `setmaxnreg` (producers kProdRegs, consumers kConsRegs).
// The pool is what the CTA itself releases (USETMAXREG ... CTAPOOL), not the SM's unallocated registers: the consumers'
// 256 * (80 - 48) = 8192 registers are exactly the producers' 512 * (96 - 80); asking for more never returns.  Hand-off through named barriers: full[b] (producers arrive, consumers wait) and
// empty[b] (consumers arrive, producers wait); the producers' own level barrier is a third named barrier.  The value
// stores, the descriptor stream and the G reads belong to the consumers alone, the x gather, the items and the g stores
// to the producers.
#pragma once
#include "eval_kernel.cuh"

namespace tb {

constexpr int kSplitConsumersPerLane = 128;
constexpr int kProdRegs = 96, kConsRegs = 48;
static_assert(kThreads * (kProdRegs - 80) == kSplitConsumersPerLane * (80 - kConsRegs), "setmaxnreg: the producers can only take what the consumers release");
constexpr int kBarProducers = 1, kBarFull = 2, kBarEmpty = 4;     // named barriers (0 is __syncthreads); full / empty + buffer
constexpr int kSplitAhead = 3;                                    // consumer descriptor batches in flight

template <int N>
__device__ __forceinline__ void RegInc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N)); }
template <int N>
__device__ __forceinline__ void RegDec() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N)); }
__device__ __forceinline__ void BarSync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }
__device__ __forceinline__ void BarArrive(int id, int n) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(n) : "memory"); }

template <int WHAT, int P, bool RT, bool CURV>
__global__ void __launch_bounds__((kThreads + kSplitConsumersPerLane) * P, 1)
EvalKernelSplit(const __grid_constant__ DeviceProgram Pg, const ProblemConst* __restrict__ pcs, const double* __restrict__ x,
                double* __restrict__ g_out, double* __restrict__ vals_out, int* __restrict__ status, int n_problems) {
  constexpr int NTP = kThreads * P, NTC = kSplitConsumersPerLane * P, NTALL = NTP + NTC;
  constexpr bool JAC = (WHAT & TOWR_B200_EVAL_JAC) != 0;
  extern __shared__ uint4 sm4[];
  unsigned char* blob = reinterpret_cast<unsigned char*>(sm4);
  const int nnv = Pg.blob.n_node_values, npoly = Pg.blob.n_polys;
  double* nodes = reinterpret_cast<double*>(blob + Pg.blob.bytes);
  double* coef = nodes + nnv * P;
  double* rtd = coef + 6 * npoly * P;
  double* pdur = rtd + (RT ? kRtd * npoly * P : 0);
  double* G0 = pdur + (RT ? Pg.blob.n_pdur * P : 0);      // two G buffers of g_capacity * P doubles
  __shared__ ProblemConst pc_s[P];

  {
    const uint4* src = reinterpret_cast<const uint4*>(Pg.blob_data);
    for (int i = threadIdx.x; i < Pg.blob.bytes / 16; i += NTALL) sm4[i] = __ldg(src + i);
  }
  __syncthreads();
  const int first = static_cast<int>(blockIdx.x) * P, stride = static_cast<int>(gridDim.x) * P;
  const int n_groups = first < n_problems ? (n_problems - first + stride - 1) / stride : 0;
  const int total_seq = n_groups * Pg.n_segments;         // segments this CTA evaluates

  if (threadIdx.x >= NTP) {
    // ------------------------------------------------------------------ consumers: stage B
    RegDec<kConsRegs>();
    if constexpr (JAC) {
      const int tid = static_cast<int>(threadIdx.x) - NTP;
      const double* Bs = MakeTables(blob, Pg.blob).B;
      const SegmentRec* segs = MakeTables(blob, Pg.blob).segments;
      int seq = 0;
      for (int base = first; base < n_problems; base += stride) {
        double nf[P];
#pragma unroll
        for (int p = 0; p < P; ++p) nf[p] = 0.0;
        for (int s = 0; s < Pg.n_segments; ++s, ++seq) {
          const SegmentRec sg = segs[s];
          const int b = seq & 1;
          const uint32_t* ent = Pg.entries + sg.entry0 + tid;
          uint32_t d0[kSplitAhead][kDescBatch];
          LoadDescriptorRing<NTC, kSplitAhead, kDescBatch>(ent, sg.n_entries - tid, d0);
          BarSync(kBarFull + b, NTALL);
          const double* G = G0 + static_cast<size_t>(b) * Pg.g_capacity * P;
          double* o0 = vals_out + static_cast<size_t>(base) * Pg.nnz + sg.entry0 + tid;
          if (n_problems - base >= P)
            StageB<P, true, NTC, RT, kSplitAhead>(ent, sg.n_entries - tid, d0, G, Bs, o0, Pg.nnz, P, nf);
          else
            StageB<P, false, NTC, RT, kSplitAhead>(ent, sg.n_entries - tid, d0, G, Bs, o0, Pg.nnz, n_problems - base, nf);
          if (seq + 2 < total_seq) BarArrive(kBarEmpty + b, NTALL);     // a later segment reuses this buffer
        }
#pragma unroll
        for (int p = 0; p < P; ++p)
          if (nf[p] != nf[p] && base + p < n_problems) atomicOr(status + base + p, 2);
      }
    }
    return;
  }

  // -------------------------------------------------------------------- producers: stages N, D, C, A
  RegInc<kProdRegs>();
  const int tid = static_cast<int>(threadIdx.x);
  ItemCtx<Lockstep<P, RT, CURV>> cx;
  cx.n_ee = Pg.n_ee;
  cx.T = MakeTables(blob, Pg.blob);
  const int pl = tid % P;
  auto fetch_group = [&](int base_n) {
    const double* xn = x + static_cast<size_t>(min(base_n + pl, n_problems - 1)) * Pg.n_vars;
    for (int i = tid; i < nnv * P; i += NTP) {
      const unsigned src = cx.T.node_src[i / P];
      CpAsync8(nodes + i, xn + (src != 0xFFFFu ? src : 0u), src != 0xFFFFu ? 8 : 0);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    FetchProblemConsts<P, NTP, CURV>(pc_s, pcs, base_n, n_problems, tid);
  };
  if (first < n_problems) fetch_group(first);
  int seq = 0;
  for (int base = first; base < n_problems; base += stride) {
    const int prob = min(base + pl, n_problems - 1);
    cx.nodes = nodes + pl;
    cx.coef = coef + pl;
    cx.rtd = rtd + pl;
    cx.pdur = pdur + pl;
    cx.pc = &pc_s[pl];
    cx.g = (WHAT & TOWR_B200_EVAL_G) ? g_out + static_cast<size_t>(prob) * Pg.n_rows : nullptr;
    double nfx = 0.0;
    asm volatile("cp.async.wait_all;" ::: "memory");
    BarSync(kBarProducers, NTP);

    if constexpr (RT) {                                   // stage D, as in eval_kernel.cuh
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
        pdur[(ft.phase0 + ft.n_phases - 1) * P + pl] = ft.total - sum;
      }
      BarSync(kBarProducers, NTP);
      for (int i = tid; i < npoly * P; i += NTP) {
        const int poly = i / P;
        const uint32_t pm = cx.T.poly_meta[poly];
        double* o = rtd + poly * kRtd * P + pl;
        if (pm == kPolyFixed) {
          const int di = cx.T.polys[poly] >> 16;
          o[kRtT * P] = cx.T.durs[3 * di]; o[kRtT2 * P] = cx.T.durs[3 * di + 1]; o[kRtT3 * P] = cx.T.durs[3 * di + 2];
          o[kRtR2 * P] = cx.T.rdurs[2 * di]; o[kRtR3 * P] = cx.T.rdurs[2 * di + 1];
          o[kRtR1 * P] = 0.0; o[kRtR4 * P] = 0.0; o[kRtEnd * P] = 0.0;
        } else {
          const RtFoot ft = cx.T.rt_feet[pm >> 24];
          const double Td = pdur[(ft.phase0 + (pm & 0xFF)) * P + pl] / static_cast<double>((pm >> 8) & 0xFF);
          const double T2 = Td * Td, T3 = Cube(Td);
          o[kRtT * P] = Td; o[kRtT2 * P] = T2; o[kRtT3 * P] = T3; o[kRtR2 * P] = 1.0 / T2; o[kRtR3 * P] = 1.0 / T3;
          o[kRtR1 * P] = 1.0 / Td; o[kRtR4 * P] = 1.0 / Pow4(Td);
        }
      }
      BarSync(kBarProducers, NTP);
      for (int i = tid; i < Pg.blob.n_rt_splines * P; i += NTP) {
        const RtSpline sp = cx.T.rt_splines[i / P];
        double acc = 0.0;
        for (int j = 0; j < sp.n_polys; ++j) {
          double* o = rtd + (sp.poly0 + j) * kRtd * P + pl;
          acc += o[kRtT * P];
          o[kRtEnd * P] = acc;
        }
      }
      BarSync(kBarProducers, NTP);
    }

    // stage C
    for (int i = tid; i < 3 * npoly * P; i += NTP) {
      const int pd = i / P;
      const int poly = pd / 3, d = pd - 3 * poly;
      const uint32_t pr = cx.T.polys[poly];
      const int n0 = pr & 0xFFFF;
      double T, T2, T3, r2, r3;
      if constexpr (RT) {
        const double* o = rtd + poly * kRtd * P + pl;
        T = o[kRtT * P]; T2 = o[kRtT2 * P]; T3 = o[kRtT3 * P]; r2 = o[kRtR2 * P]; r3 = o[kRtR3 * P];
      } else {
        const double* du = cx.T.durs + 3 * (pr >> 16);
        const double* rd = cx.T.rdurs + 2 * (pr >> 16);
        T = du[0]; T2 = du[1]; T3 = du[2]; r2 = rd[0]; r3 = rd[1];
      }
      const double* nd = nodes + pl;
      const double p0 = nd[(n0 * 6 + d) * P], v0 = nd[(n0 * 6 + 3 + d) * P];
      const double p1 = nd[(n0 * 6 + 6 + d) * P], v1 = nd[(n0 * 6 + 9 + d) * P];
      nfx = fma(p0, 0.0, fma(v0, 0.0, fma(p1, 0.0, fma(v1, 0.0, nfx))));
      coef[(poly * 6 + d) * P + pl] = ExactDiv(-(3.0 * (p0 - p1) + T * (2.0 * v0 + v1)), T2, r2);
      coef[(poly * 6 + 3 + d) * P + pl] = ExactDiv(2.0 * (p0 - p1) + T * (v0 + v1), T3, r3);
    }
    BarSync(kBarProducers, NTP);

    for (int s = 0; s < Pg.n_segments; ++s, ++seq) {
      const SegmentRec sg = cx.T.segments[s];
      const int b = seq & 1;
      double* G = G0 + static_cast<size_t>(b) * Pg.g_capacity * P;
      if (JAC && seq >= 2) BarSync(kBarEmpty + b, NTALL);   // the consumers have drained this buffer's previous segment
      cx.G = G + pl;
      if constexpr (RT && TB_PREZERO_G) {
        for (int i = tid; i < sg.g_size * P; i += NTP) G[i] = 0.0;
        BarSync(kBarProducers, NTP);
      }
      if (tid < P) { G[kGOne * P + tid] = 1.0; G[kGMass * P + tid] = pc_s[tid].mass; G[kGMinusOne * P + tid] = -1.0; }
      auto run_level = [&](int first_item, int count) {
        constexpr int kPerRound = NTP / P;
        int r = 0;
        for (int base_i = 0; base_i < count; base_i += kPerRound, ++r) {
          const int idx = base_i + ((r & 1) ? kPerRound - 1 - tid / P : tid / P);
          if (idx < count) RunItem<Lockstep<P, RT, CURV>>(cx, cx.T.items[first_item + idx]);
        }
      };
      run_level(sg.item0, sg.n_items0);
      if (sg.n_items > sg.n_items0) {
        BarSync(kBarProducers, NTP);
        run_level(sg.item0 + sg.n_items0, sg.n_items - sg.n_items0);
      }
      if (s == Pg.n_segments - 1) {
        // every producer is done with nodes[] / coef[] / pc_s of this group: gather the next group's x
        BarSync(kBarProducers, NTP);
        if (base + stride < n_problems) fetch_group(base + stride);
      }
      if (JAC) BarArrive(kBarFull + b, NTALL);
    }
    if (nfx != nfx) atomicOr(status + prob, 2);
  }
}

}  // namespace tb
