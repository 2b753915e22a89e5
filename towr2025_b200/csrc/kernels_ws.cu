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

// BULK: the consumer warps do not store to global memory themselves.  Each warp multiplies a sub-block of kSubBlock
// consecutive entries into a private ring slot in shared memory (storage order, one row per problem lane) and one lane per
// problem hands the row to the TMA engine (cp.async.bulk shared -> global, >= 2 KB per copy: the engine retires about one
// copy per 64 clocks, tools/micro/bulk_store_bw.cu).  The value stores then never enter the LSU pipe, which the producers
// need for stage A's shared-memory traffic, and the consumers are done with a G buffer as soon as its products are staged.
constexpr int kSubBlock = 256;          // entries per sub-block: 8 per lane
constexpr int kRingSlots = 2;           // sub-blocks in flight per consumer warp
constexpr int kRingRow = kSubBlock + 2; // doubles per problem lane in a slot (+ parity slack, even)

template <int WHAT, int P, int NA, int NB, bool BULK = false>
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
  // BULK: ring of staging slots, one set per consumer warp, 16-byte aligned
  double* const ring_base = reinterpret_cast<double*>((reinterpret_cast<uintptr_t>(Gbuf + 2 * gstride) + 15) & ~static_cast<uintptr_t>(15));
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
    int slot = 0;                                          // BULK: ring slot counter of this warp
    for (int base = blockIdx.x * P; base < n_problems; base += gridDim.x * P) {
      double nf[P];
#pragma unroll
      for (int p = 0; p < P; ++p) nf[p] = 0.0;
      for (int s = 0; s < nseg; ++s, ++k) {
        const SegmentRec sg = T.segments[s];
        const int buf = k & 1;
        const double* G = Gbuf + buf * gstride;
        if constexpr (BULK) {
          constexpr int NBW = NB / 32, U = kSubBlock / 32;
          const int cw = ct >> 5, lane = ct & 31;
          double* const ring = ring_base + cw * (kRingSlots * P * kRingRow);
          const uint32_t* ent = Pg.entries + sg.entry0;
          const int ne = sg.n_entries;
          const int n_sub = (ne + kSubBlock - 1) / kSubBlock;
          uint32_t dn[U];
          auto load_desc = [&](int j, uint32_t (&d)[U]) {
            const int e0 = j * kSubBlock + lane;
#pragma unroll
            for (int u = 0; u < U; ++u) d[u] = (j < n_sub && e0 + 32 * u < ne) ? __ldg(ent + e0 + 32 * u) : kNoEntry;
          };
          if (WHAT & TOWR_B200_EVAL_JAC) load_desc(cw, dn);            // before the wait: off the critical path
          TB_MARK_T(16, NA)
          BarSync(kBarFull + buf, NALL);
          TB_MARK_T(17, NA)
          if (WHAT & TOWR_B200_EVAL_JAC) {
            // 8-byte parity of each lane's first destination element (sub-blocks start at even offsets)
            int q[P];
#pragma unroll
            for (int p = 0; p < P; ++p) {
              const size_t e0 = static_cast<size_t>(min(base + p, n_problems - 1)) * Pg.nnz + sg.entry0;
              q[p] = static_cast<int>(((reinterpret_cast<uintptr_t>(vals_out) >> 3) + e0) & 1);
            }
            const char* Gb = reinterpret_cast<const char*>(G);
            const char* Bb = reinterpret_cast<const char*>(T.B);
            for (int j = cw; j < n_sub; j += NBW, ++slot) {
              uint32_t d[U];
#pragma unroll
              for (int u = 0; u < U; ++u) d[u] = dn[u];
              load_desc(j + NBW, dn);                                   // next sub-block of this warp
              // the slot's previous copy must have been read out
              if (lane < P) asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(kRingSlots - 1) : "memory");
              __syncwarp();
              TB_MARK_T(19, NA)
              double* const sl = ring + (slot % kRingSlots) * (P * kRingRow);
              auto emit = [&](uint32_t w, int u) {
                const double bv = *reinterpret_cast<const double*>(Bb + (w >> 16));
                const char* ga = Gb + (w & 0xFFFFu) * P;
#pragma unroll
                for (int p = 0; p < P; p += 2) {
                  const double2 t = *reinterpret_cast<const double2*>(ga + 8 * p);
                  const double v0 = t.x * bv, v1 = t.y * bv;
                  nf[p] = fma(v0, 0.0, nf[p]);
                  nf[p + 1] = fma(v1, 0.0, nf[p + 1]);
                  sl[p * kRingRow + q[p] + 32 * u + lane] = v0;
                  sl[(p + 1) * kRingRow + q[p + 1] + 32 * u + lane] = v1;
                }
              };
              if (ne - j * kSubBlock >= kSubBlock) {
                // full sub-block: all operands first, then the products (the staging stores could alias G and B for the
                // compiler, which would otherwise keep every load behind the previous entry's stores)
                double bv[U];
                double2 gv[U][P / 2];
#pragma unroll
                for (int u = 0; u < U; ++u) {
                  bv[u] = *reinterpret_cast<const double*>(Bb + (d[u] >> 16));
                  const char* ga = Gb + (d[u] & 0xFFFFu) * P;
#pragma unroll
                  for (int p = 0; p < P; p += 2) gv[u][p / 2] = *reinterpret_cast<const double2*>(ga + 8 * p);
                }
#pragma unroll
                for (int u = 0; u < U; ++u)
#pragma unroll
                  for (int p = 0; p < P; p += 2) {
                    const double v0 = gv[u][p / 2].x * bv[u], v1 = gv[u][p / 2].y * bv[u];
                    nf[p] = fma(v0, 0.0, nf[p]);
                    nf[p + 1] = fma(v1, 0.0, nf[p + 1]);
                    sl[p * kRingRow + q[p] + 32 * u + lane] = v0;
                    sl[(p + 1) * kRingRow + q[p + 1] + 32 * u + lane] = v1;
                  }
              } else {
#pragma unroll
                for (int u = 0; u < U; ++u)
                  if (d[u] != kNoEntry) emit(d[u], u);
              }
              TB_MARK_T(20, NA)
#if defined(TB_WSB_FENCE_ONE)
              __syncwarp();
              if (lane < P) FenceProxyAsync();
#elif !defined(TB_WSB_NOFENCE)
              FenceProxyAsync();
              __syncwarp();
#else
              __syncwarp();
#endif
              if (lane < P && base + lane < n_problems) {
                const int e0 = j * kSubBlock, cnt = min(kSubBlock, ne - e0);
                double* dst = vals_out + static_cast<size_t>(base + lane) * Pg.nnz + sg.entry0 + e0;
                int ql = q[0];
#pragma unroll
                for (int p = 1; p < P; ++p)
                  if (lane == p) ql = q[p];
                const double* src = sl + lane * kRingRow + ql;         // entry i of the sub-block at src[i]
                if (ql) dst[0] = src[0];
                const int body = (cnt - ql) & ~1;
                if (body > 0) BulkStore(dst + ql, src + ql, static_cast<unsigned>(body) * 8u);
                if ((cnt - ql) & 1) dst[cnt - 1] = src[cnt - 1];
              }
              if (lane < P) BulkCommit();
              TB_MARK_T(21, NA)
            }
          }
        } else {
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
        }
        TB_MARK_T(18, NA)
        if (k + 2 < k_total) BarArrive(kBarEmpty + buf, NALL);
      }
#pragma unroll
      for (int p = 0; p < P; ++p)
        if (nf[p] != nf[p] && base + p < n_problems) atomicOr(status + base + p, 2);
    }
    // the ring must outlive the bulk copies that read it
    if (BULK && (WHAT & TOWR_B200_EVAL_JAC) && (ct & 31) < P) BulkWaitAll();
  }
}

size_t EvalSharedBytesWS(const DeviceProgram& P, int lanes) {
  return static_cast<size_t>(P.blob.bytes) + sizeof(double) * lanes *
         (static_cast<size_t>(P.blob.n_node_values) + 6 * static_cast<size_t>(P.blob.n_polys) + 2 * static_cast<size_t>(P.g_capacity));
}

#ifndef TB_WS_NA
#define TB_WS_NA 320
#endif
#ifndef TB_WS_NB
#define TB_WS_NB 192
#endif
#ifndef TB_WSB_NA
#define TB_WSB_NA 384
#endif
#ifndef TB_WSB_NB
#define TB_WSB_NB 128
#endif

size_t EvalSharedBytesWSBulk(const DeviceProgram& P, int lanes) {
  return EvalSharedBytesWS(P, lanes) + 16 + sizeof(double) * (TB_WSB_NB / 32) * kRingSlots * lanes * kRingRow;
}

template <int WHAT, int LANES, bool BULK>
static cudaError_t LaunchWS(const DeviceProgram& P, const ProblemConst* pcs, const double* x, double* g, double* vals,
                            int* status, int n_problems, int max_ctas, cudaStream_t stream) {
  constexpr int NA = BULK ? TB_WSB_NA : TB_WS_NA, NB = BULK ? TB_WSB_NB : TB_WS_NB;
  const size_t smem = BULK ? EvalSharedBytesWSBulk(P, LANES) : EvalSharedBytesWS(P, LANES);
  auto kern = EvalKernelWS<WHAT, LANES, NA, NB, BULK>;
  cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
  if (err != cudaSuccess) return err;
  int grid = (n_problems + LANES - 1) / LANES;
  if (grid > max_ctas) grid = max_ctas;
  kern<<<grid, NA + NB, smem, stream>>>(P, pcs, x, g, vals, status, n_problems);
  return cudaGetLastError();
}

cudaError_t LaunchEvalWS(const DeviceProgram& P, const ProblemConst* pcs, int what, const double* x, double* g,
                         double* vals, int* status, int n_problems, int lanes, int max_ctas, cudaStream_t stream) {
  if (n_problems <= 0) return cudaSuccess;
  if (lanes != 2) return cudaErrorInvalidValue;
  switch (what) {
    case TOWR_B200_EVAL_G: return LaunchWS<1, 2, false>(P, pcs, x, g, vals, status, n_problems, max_ctas, stream);
    case TOWR_B200_EVAL_JAC: return LaunchWS<2, 2, false>(P, pcs, x, g, vals, status, n_problems, max_ctas, stream);
    case TOWR_B200_EVAL_BOTH: return LaunchWS<3, 2, false>(P, pcs, x, g, vals, status, n_problems, max_ctas, stream);
  }
  return cudaErrorInvalidValue;
}

cudaError_t LaunchEvalWSBulk(const DeviceProgram& P, const ProblemConst* pcs, int what, const double* x, double* g,
                             double* vals, int* status, int n_problems, int lanes, int max_ctas, cudaStream_t stream) {
  if (n_problems <= 0) return cudaSuccess;
  if (lanes != 2) return cudaErrorInvalidValue;
  switch (what) {
    case TOWR_B200_EVAL_G: return LaunchWS<1, 2, true>(P, pcs, x, g, vals, status, n_problems, max_ctas, stream);
    case TOWR_B200_EVAL_JAC: return LaunchWS<2, 2, true>(P, pcs, x, g, vals, status, n_problems, max_ctas, stream);
    case TOWR_B200_EVAL_BOTH: return LaunchWS<3, 2, true>(P, pcs, x, g, vals, status, n_problems, max_ctas, stream);
  }
  return cudaErrorInvalidValue;
}

bool EvalFitsWSBulk(const DeviceProgram& P, int lanes, size_t smem_optin) {
  if (lanes != 2 || P.blob.rt) return false;
  return EvalSharedBytesWSBulk(P, lanes) + 1024 <= smem_optin;
}

bool EvalFitsWS(const DeviceProgram& P, int lanes, size_t smem_optin) {
  if (lanes != 2 || P.blob.rt) return false;   // optimised durations run on the lock-step kernel
  const size_t smem = EvalSharedBytesWS(P, lanes) + 1024;   // static shared memory of the kernel
  return smem <= smem_optin;
}

}  // namespace tb

TB_PHASE_CLOCK_READER(towr_b200_debug_phase_clocks_ws)
