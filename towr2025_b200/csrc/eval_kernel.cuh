// eval_kernel.cuh — the evaluation kernel (sm_100a).
//
// One CTA runs S independent *streams* of NT = 256 * P threads; a stream evaluates P problems at a time in lock step
// (grid-stride over the batch).  Shared memory holds the class tables once per CTA and, per stream, the problems'
// spline node values, the Hermite coefficients of every polynomial and the G buffer of the current segment; x is
// read once (coalesced gather through node_src), g and the Jacobian values are written once, in storage order.
// HBM traffic per eval = 8*(n_vars+n_rows+nnz) bytes plus L2-resident class tables (entries, B, items), which are
// shared by all problems of the batch.
//
//   S = 1            the lock-step kernel of round 1: all threads run stage A, then all run stage B
//   S = 2, P = 1     two streams in anti-phase (the second starts `skew` clocks late): one stream's stage A overlaps
//                    the other's value stores; the streams only share the class tables and synchronise on named
//                    barriers of their own
//   RING             stage B writes its products into small per-warp shared-memory rings in storage order and lane 0 of
//                    each warp hands every 2 KB block to the TMA engine (cp.async.bulk shared -> global).  The value stores then
//                    never enter the SM's LSU pipe, which serialises global stores with the shared-memory traffic of
//                    stage A (profiles/r01_ncu_summary.md: mio_throttle / lg_throttle), and drain while the next
//                    chunk - or the next segment's stage A - runs.
#pragma once
#include "kernel_common.cuh"

namespace tb {

// value stores of stage B: written once and never read back by the kernel, so they go out with the streaming
// (evict-first) policy and leave the class tables and the descriptor stream in L2 (+11 % on the classes with optimised
// durations, whose descriptor stream is as long as the output; neutral on ANYmal trot)
#define TB_STORE_VALUE(ptr, v) __stcs((ptr), (v))

// one product: descriptor w = G byte offset (low 16) | B byte offset (high 16; bit 16 set: second factor is a G slot)
template <int P, bool GG>
__device__ __forceinline__ void Products(uint32_t w, const char* Gb, const char* Bb, double (&v)[P]) {
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
  for (int p = 0; p < P; ++p) v[p] = gv[p] * ((GG && gg) ? bw[p] : bv);
}

// FULL: all P problem lanes of the group exist (no per-store predicate).
// nf[p] accumulates value * 0 per problem lane: NaN exactly when a stored value is not finite (status bit 2).
template <int P, bool FULL, int NT, bool GG = false, int A = DescAhead<GG>::value>
__device__ __forceinline__ void StageB(const uint32_t* __restrict__ ent, int left, uint32_t (&d)[A][kDescBatch],
                                       const double* G, const double* Bs, double* __restrict__ o, int nnz, int np,
                                       double (&nf)[P]) {
  constexpr int U = kDescBatch;
  const char* Gb = reinterpret_cast<const char*>(G);
  const char* Bb = reinterpret_cast<const char*>(Bs);
  auto emit = [&](uint32_t w, double* op) {
    double v[P];
    Products<P, GG>(w, Gb, Bb, v);
#pragma unroll
    for (int p = 0; p < P; ++p) {
      nf[p] = fma(v[p], 0.0, nf[p]);
      if (FULL || p < np) TB_STORE_VALUE(op + static_cast<size_t>(p) * nnz, v[p]);
    }
  };
  while (true) {
#pragma unroll
    for (int a = 0; a < A; ++a) {                         // batch in d[a]; refilled with the batch A further on
      if (left <= 0) return;
      if (left > (U - 1) * NT) {
#pragma unroll
        for (int u = 0; u < U; ++u) emit(d[a][u], o + u * NT);
      } else {
#pragma unroll
        for (int u = 0; u < U - 1; ++u)
          if (left > u * NT) emit(d[a][u], o + u * NT);
      }
      const int left_a = left - A * U * NT;
      if (left_a > 0) LoadDescriptors<NT, U>(ent + A * U * NT, left_a, d[a]);
      left -= U * NT;
      ent += U * NT;
      o += U * NT;
    }
  }
}

// ---- RING: cp.async.bulk (TMA engine) stores from a shared-memory ring
__device__ __forceinline__ void FenceProxyAsync() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void BulkStore(void* gdst, const void* ssrc, unsigned bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst),
               "r"(static_cast<unsigned>(__cvta_generic_to_shared(ssrc))), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void BulkCommit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void BulkWaitRead() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void BulkWaitAll() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// all NT threads of stream `sidx`
template <int S, int NT>
__device__ __forceinline__ void StreamSync(int sidx) {
  if constexpr (S == 1) __syncthreads();
  else asm volatile("bar.sync %0, %1;" ::"r"(sidx + 1), "n"(NT) : "memory");
}

// Ring geometry (program.h): every WARP of a stream owns kRingDepth slots of kRingBlock = 32 * kRingUB consecutive entries
// per problem lane.  A lane's row has two spare elements: the row is shifted by the 8-byte parity of its destination so
// that the body of every copy is 16-byte aligned on both sides.
//
// Stage B of one segment through the ring.  Warp w takes blocks w, w + NW, ... of the segment (a block = kRingBlock
// consecutive entries; lane l takes entries l, l + 32, ... of it, so descriptor loads stay coalesced).  Per block: lane 0
// waits until the engine has READ the copy that used this slot kRingDepth blocks ago, the warp writes its products into
// the slot and fences them for the async proxy, __syncwarp, lane 0 issues one cp.async.bulk per problem lane (2 KB).  No
// CTA-wide barrier: the warps of a stream run independently and up to NW * kRingDepth copies are in flight.
// ctr: blocks this warp has staged so far (selects the slot).  dst[p]: global address of the segment's first entry of
// problem lane p; np lanes exist.  d: descriptors of this warp's first AH blocks (loaded before the barrier).
template <int P, int NT, int AH, bool GG>
__device__ __forceinline__ void StageBRing(const uint32_t* __restrict__ ent_seg, int n_entries, int tid,
                                           uint32_t (&d)[AH][kRingUB], const double* G, const double* Bs, double* ring_warp,
                                           unsigned& ctr, double* const (&dst)[P], int np, double (&nf)[P]) {
  constexpr int UB = kRingUB, BE = kRingBlock, ROW = kRingBlock + 2, NW = NT / 32;
  const char* Gb = reinterpret_cast<const char*>(G);
  const char* Bb = reinterpret_cast<const char*>(Bs);
  const int warp = tid >> 5, lane = tid & 31;
  int q[P];
#pragma unroll
  for (int p = 0; p < P; ++p) q[p] = static_cast<int>((reinterpret_cast<uintptr_t>(dst[p]) >> 3) & 1);
  const int n_blocks = (n_entries + BE - 1) / BE;
  for (int blk0 = warp; blk0 < n_blocks; blk0 += NW * AH) {
#pragma unroll
    for (int a = 0; a < AH; ++a) {
      const int blk = blk0 + a * NW;
      if (blk >= n_blocks) break;
      const int e0 = blk * BE;
      double* sb = ring_warp + (ctr % kRingDepth) * (P * ROW);
      if (lane == 0) BulkWaitRead<kRingDepth - 1>();
      __syncwarp();
#pragma unroll
      for (int u = 0; u < UB; ++u) {
        if (e0 + lane + 32 * u < n_entries) {
          double v[P];
          Products<P, GG>(d[a][u], Gb, Bb, v);
#pragma unroll
          for (int p = 0; p < P; ++p) {
            nf[p] = fma(v[p], 0.0, nf[p]);
            sb[p * ROW + q[p] + lane + 32 * u] = v[p];
          }
        }
      }
      // this warp's block AH further on
      const int e_next = (blk + AH * NW) * BE;
#pragma unroll
      for (int u = 0; u < UB; ++u) d[a][u] = e_next + lane + 32 * u < n_entries ? __ldg(ent_seg + e_next + lane + 32 * u) : kNoEntry;
      FenceProxyAsync();
      __syncwarp();
      if (lane == 0) {
        const int nc = min(BE, n_entries - e0);
#pragma unroll
        for (int p = 0; p < P; ++p) {
          if (p < np) {
            double* D0 = dst[p] + e0;
            const double* src = sb + p * ROW + q[p];          // entry i of the block at src[i]
            if (q[p]) D0[0] = src[0];                         // odd first element: plain store
            const int body = (nc - q[p]) & ~1;
            if (body > 0) BulkStore(D0 + q[p], src + q[p], static_cast<unsigned>(body) * 8u);
            if ((nc - q[p]) & 1) D0[nc - 1] = src[nc - 1];    // odd last element
          }
        }
        BulkCommit();
      }
      ++ctr;
    }
  }
}

// RT: structure classes with optimised phase durations (stage D below; items.cuh Lockstep<P, true>).
template <int WHAT, int P, bool RT, int S = 1, bool RING = false, bool CURV = false>
__global__ void __launch_bounds__(kThreads * P * S, (P == 1 && S == 1 && !RT) ? 2 : 1)   // one lane, one stream: two CTAs per SM (128 registers each)
EvalKernel(const __grid_constant__ DeviceProgram Pg, const ProblemConst* __restrict__ pcs, const double* __restrict__ x,
           double* __restrict__ g_out, double* __restrict__ vals_out, int* __restrict__ status, int n_problems, int skew) {
  constexpr int NT = kThreads * P;
  constexpr int UB = RING ? kRingUB : kDescBatch;                   // descriptors per batch
  constexpr int AH = RING ? 2 : DescAhead<RT>::value;               // batches fetched ahead
  extern __shared__ uint4 sm4[];
  unsigned char* blob = reinterpret_cast<unsigned char*>(sm4);
  const int nnv = Pg.blob.n_node_values, npoly = Pg.blob.n_polys;
  const int sidx = S == 1 ? 0 : static_cast<int>(threadIdx.x) / NT;
  const int tid = S == 1 ? static_cast<int>(threadIdx.x) : static_cast<int>(threadIdx.x) % NT;
  // per-stream arrays (StreamDoubles, program.h)
  double* nodes = reinterpret_cast<double*>(blob + Pg.blob.bytes) + static_cast<size_t>(sidx) * StreamDoubles(Pg, P, RING);
  double* coef = nodes + nnv * P;
  double* rtd = coef + 6 * npoly * P;                   // RT: kRtd per polynomial, then the phase durations
  double* pdur = rtd + (RT ? kRtd * npoly * P : 0);
  double* G = pdur + (RT ? Pg.blob.n_pdur * P : 0);
  // RING: this warp's slots (16-byte aligned: the stream's arrays hold an even number of doubles, a slot as well)
  double* const ring = reinterpret_cast<double*>((reinterpret_cast<uintptr_t>(G + Pg.g_capacity * P) + 15) & ~static_cast<uintptr_t>(15)) +
                       (tid >> 5) * (kRingDepth * P * (kRingBlock + 2));
  __shared__ ProblemConst pc_all[P * S];
  __shared__ int bad_all[P * S];
  ProblemConst* const pc_s = pc_all + sidx * P;
  int* const bad_s = bad_all + sidx * P;

  // class tables: one 16-byte-vector copy per CTA, reused for every problem group this CTA evaluates
  {
    const uint4* src = reinterpret_cast<const uint4*>(Pg.blob_data);
    for (int i = threadIdx.x; i < Pg.blob.bytes / 16; i += NT * S) sm4[i] = __ldg(src + i);
  }
  __syncthreads();
  TB_MARK_INIT
  ItemCtx<Lockstep<P, RT, CURV>> cx;
  cx.n_ee = Pg.n_ee;
  cx.T = MakeTables(blob, Pg.blob);
  const double* Bs = cx.T.B;
  const int pl = tid % P;                               // problem lane of this thread in stages N, C, A
  const int first = (static_cast<int>(blockIdx.x) * S + sidx) * P, stride = static_cast<int>(gridDim.x) * S * P;
  unsigned ring_ctr = 0;

  // stage N: node values from x (values that are not optimised are pinned to 0), fetched asynchronously: for the first
  // group here, for every later group while the previous group's last stage B runs (nodes[] is free by then)
  auto fetch_group = [&](int base_n) {
    const double* xn = x + static_cast<size_t>(min(base_n + pl, n_problems - 1)) * Pg.n_vars;
    for (int i = tid; i < nnv * P; i += NT) {
      const unsigned src = cx.T.node_src[i / P];
      CpAsync8(nodes + i, xn + (src != 0xFFFFu ? src : 0u), src != 0xFFFFu ? 8 : 0);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    FetchProblemConsts<P, NT, CURV>(pc_s, pcs, base_n, n_problems, tid);
  };
  if (first < n_problems) fetch_group(first);
  // skew > 0: later streams (S > 1) or odd CTAs (S = 1) start `skew` clocks late, so that not every SM of the chip is in
  // its HBM-idle stage A - or its HBM-bound stage B - at the same time
  {
    const int late = S > 1 ? sidx : static_cast<int>(blockIdx.x & 1);
    if (late > 0 && skew > 0) {
      const long long t0 = clock64();
      while (clock64() - t0 < static_cast<long long>(skew) * late) __nanosleep(200);
    }
  }

  for (int base = first; base < n_problems; base += stride) {
    // the last group may hold fewer than P problems: the spare lanes recompute the last problem (same values to
    // the same addresses) so that no stage needs a bounds test
    const int prob = min(base + pl, n_problems - 1);
    cx.nodes = nodes + pl;
    cx.coef = coef + pl;
    cx.G = G + pl;
    cx.rtd = rtd + pl;
    cx.pdur = pdur + pl;
    cx.pc = &pc_s[pl];
    cx.g = (WHAT & TOWR_B200_EVAL_G) ? g_out + static_cast<size_t>(prob) * Pg.n_rows : nullptr;
    double nf[P];
#pragma unroll
    for (int p = 0; p < P; ++p) nf[p] = 0.0;
    double nfx = 0.0;                                     // non-finite x of problem lane pl
    if (tid < P) bad_s[tid] = 0;
    asm volatile("cp.async.wait_all;" ::: "memory");
    TB_MARK(0)
    StreamSync<S, NT>(sidx);
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
        pdur[(ft.phase0 + ft.n_phases - 1) * P + pl] = ft.total - sum;
      }
      StreamSync<S, NT>(sidx);
      for (int i = tid; i < npoly * P; i += NT) {
        const int poly = i / P;
        const uint32_t pm = cx.T.poly_meta[poly];
        double* o = rtd + poly * kRtd * P + pl;
        if (pm == kPolyFixed) {
          const int di = cx.T.polys[poly] >> 16;
          o[kRtT * P] = cx.T.durs[3 * di]; o[kRtT2 * P] = cx.T.durs[3 * di + 1]; o[kRtT3 * P] = cx.T.durs[3 * di + 2];
          o[kRtR2 * P] = cx.T.rdurs[2 * di]; o[kRtR3 * P] = cx.T.rdurs[2 * di + 1];
          o[kRtR1 * P] = 0.0; o[kRtR4 * P] = 0.0; o[kRtEnd * P] = 0.0;   // base splines are never searched at run time
        } else {
          const RtFoot ft = cx.T.rt_feet[pm >> 24];
          const double Td = pdur[(ft.phase0 + (pm & 0xFF)) * P + pl] / static_cast<double>((pm >> 8) & 0xFF);
          const double T2 = Td * Td, T3 = Cube(Td);
          o[kRtT * P] = Td; o[kRtT2 * P] = T2; o[kRtT3 * P] = T3; o[kRtR2 * P] = 1.0 / T2; o[kRtR3 * P] = 1.0 / T3;
          o[kRtR1 * P] = 1.0 / Td; o[kRtR4 * P] = 1.0 / Pow4(Td);
        }
      }
      StreamSync<S, NT>(sidx);
      // end time of every polynomial within its PhaseSpline: t += d from 0, as Spline::GetSegmentID (spline.cc:48-66)
      for (int i = tid; i < Pg.blob.n_rt_splines * P; i += NT) {
        const RtSpline sp = cx.T.rt_splines[i / P];
        double acc = 0.0;
        for (int j = 0; j < sp.n_polys; ++j) {
          double* o = rtd + (sp.poly0 + j) * kRtd * P + pl;
          acc += o[kRtT * P];
          o[kRtEnd * P] = acc;
        }
      }
      StreamSync<S, NT>(sidx);
    }

    // ---- stage C: c = -(3(p0-p1) + T(2 v0 + v1))/T^2, d = (2(p0-p1) + T(v0+v1))/T^3   (polynomial.cc:97-104)
    for (int i = tid; i < 3 * npoly * P; i += NT) {
      const int pd = i / P;                             // (poly, dim) pair; i % P == pl
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
      nfx = fma(p0, 0.0, fma(v0, 0.0, fma(p1, 0.0, fma(v1, 0.0, nfx))));   // NaN iff a node value is not finite
      coef[(poly * 6 + d) * P + pl] = ExactDiv(-(3.0 * (p0 - p1) + T * (2.0 * v0 + v1)), T2, r2);
      coef[(poly * 6 + 3 + d) * P + pl] = ExactDiv(2.0 * (p0 - p1) + T * (v0 + v1), T3, r3);
    }
    TB_MARK(2)
    StreamSync<S, NT>(sidx);
    TB_MARK(3)


    for (int s = 0; s < Pg.n_segments; ++s) {
      const SegmentRec sg = cx.T.segments[s];
      if constexpr (RT && TB_PREZERO_G) {                // the all-column weight vectors start from 0 (device_math.cuh)
        for (int i = tid; i < sg.g_size * P; i += NT) G[i] = 0.0;
        StreamSync<S, NT>(sidx);
      }
      // ---- stage A, level 0 then level 1: thread -> (item, problem lane)
      if (tid < P) { G[kGOne * P + tid] = 1.0; G[kGMass * P + tid] = pc_s[tid].mass; G[kGMinusOne * P + tid] = -1.0; }
      // a level with more (item, lane) pairs than threads runs in rounds of NT / P items.  The kinds are ordered longest
      // first (pack.cc) and odd rounds map the items to the threads in reverse, so the short items of a second round go to
      // the threads whose first item was short
      auto run_level = [&](int first_item, int count) {
        constexpr int kPerRound = NT / P;
        int r = 0;
        for (int base_i = 0; base_i < count; base_i += kPerRound, ++r) {
          const int idx = base_i + ((r & 1) ? kPerRound - 1 - tid / P : tid / P);
          if (idx < count) RunItem<Lockstep<P, RT, CURV>>(cx, cx.T.items[first_item + idx]);
        }
      };
      run_level(sg.item0, sg.n_items0);
      TB_MARK(8 + 6 * s)
      if (sg.n_items > sg.n_items0) {
        StreamSync<S, NT>(sidx);
        TB_MARK(9 + 6 * s)
        run_level(sg.item0 + sg.n_items0, sg.n_items - sg.n_items0);
        TB_MARK(10 + 6 * s)
      }
      // first descriptor batches of stage B: issued before the barrier, consumed after it
      const uint32_t* ent = Pg.entries + sg.entry0 + tid;
      uint32_t d0[AH][UB];
      if (WHAT & TOWR_B200_EVAL_JAC) {
        if constexpr (RING) {                           // this warp's first AH blocks (StageBRing's entry order)
#pragma unroll
          for (int a = 0; a < AH; ++a)
#pragma unroll
            for (int u = 0; u < UB; ++u) {
              const int e = ((tid >> 5) + a * (NT / 32)) * kRingBlock + (tid & 31) + 32 * u;
              d0[a][u] = e < sg.n_entries ? __ldg(Pg.entries + sg.entry0 + e) : kNoEntry;
            }
        } else {
          LoadDescriptorRing<NT, AH, UB>(ent, sg.n_entries - tid, d0);
        }
      }
      StreamSync<S, NT>(sidx);
      TB_MARK(11 + 6 * s)
      // all items of this group are done with nodes[] / pc_s after the last segment's stage A: fetch the next group
      if (s == Pg.n_segments - 1 && base + stride < n_problems) fetch_group(base + stride);
      // ---- stage B: value_p = G_p[g] * B[b] for the P problems: one descriptor and one B load feed P products
      if (WHAT & TOWR_B200_EVAL_JAC) {
        if constexpr (RING) {
          double* dst[P];
#pragma unroll
          for (int p = 0; p < P; ++p) dst[p] = vals_out + static_cast<size_t>(min(base + p, n_problems - 1)) * Pg.nnz + sg.entry0;
          StageBRing<P, NT, AH, RT>(Pg.entries + sg.entry0, sg.n_entries, tid, d0, G, Bs, ring, ring_ctr, dst, min(P, n_problems - base), nf);
        } else {
          double* o0 = vals_out + static_cast<size_t>(base) * Pg.nnz + sg.entry0 + tid;
          if (n_problems - base >= P)
            StageB<P, true, NT, RT>(ent, sg.n_entries - tid, d0, G, Bs, o0, Pg.nnz, P, nf);
          else
            StageB<P, false, NT, RT>(ent, sg.n_entries - tid, d0, G, Bs, o0, Pg.nnz, n_problems - base, nf);
        }
      }
      TB_MARK(12 + 6 * s)
      StreamSync<S, NT>(sidx);                          // every G read of this segment before the next stage A
      TB_MARK(13 + 6 * s)
    }
    if (nfx != nfx) atomicOr(&bad_s[pl], 1);
#pragma unroll
    for (int p = 0; p < P; ++p)
      if (nf[p] != nf[p]) atomicOr(&bad_s[p], 1);
    StreamSync<S, NT>(sidx);
    if (tid < P && bad_s[tid] && base + tid < n_problems) atomicOr(status + base + tid, 2);
  }
  // the ring must outlive the bulk copies that read it
  if (RING && (WHAT & TOWR_B200_EVAL_JAC) && (tid & 31) == 0) BulkWaitAll();
}

}  // namespace tb
