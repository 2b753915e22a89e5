// TEST INFRASTRUCTURE ONLY — host emulation of the evaluation kernel's per-problem arithmetic.
//
// Compiles the *device* headers (items.cuh, device_math.cuh) with g++ by defining the CUDA qualifiers away and
// replays the kernel's stages serially for one problem.  It exists so the per-sample math and the table
// interpretation can be checked against the oracle on the GPU-less build box; it is never linked into
// libtowr_b200.so and nothing in the product imports it.  The GPU parity tests (-m gpu) are the real gate.
#define __device__
#define __host__
#define __forceinline__ inline
#define __noinline__
#define __restrict__
#include <math.h>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>
template <class T> static inline T __ldg(const T* p) { return *p; }

#include "../../towr2025_b200/csrc/items.cuh"
#include "../../towr2025_b200/csrc/structure.h"

using namespace tb;

static std::string g_err;

extern "C" const char* emu_last_error() { return g_err.c_str(); }

// x: n_vars; g: n_rows; vals: nnz.  Mirrors the kernel's stages for one problem (P = 1).
template <bool RT, bool CURV>
static void EmuPart(const HostProgram& hp, const ProgramPart& part, const towr_b200_spec* s, const double* x, double* g, double* vals);

// every part of the class's program (pack.cc): one kernel launch each on the device, disjoint rows and entries
template <bool RT, bool CURV>
static void EmuEval(const HostProgram& hp, const towr_b200_spec* s, const double* x, double* g, double* vals) {
  for (const ProgramPart& part : hp.parts) EmuPart<RT, CURV>(hp, part, s, x, g, vals);
}

template <bool RT, bool CURV>
static void EmuPart(const HostProgram& hp, const ProgramPart& part, const towr_b200_spec* s, const double* x, double* g, double* vals) {
  const BlobLayout& L = part.blob;
  ProblemConst pc;
  FillProblemConst(*s, pc);
  ItemCtx<Lockstep<1, RT, CURV>> cx;
  cx.T = MakeTables(part.blob_data.data(), L);
  cx.n_ee = hp.n_ee;
  std::vector<double> nodes(L.n_node_values), coef(6 * L.n_polys), G(hp.g_capacity), rtd(kRtd * L.n_polys), pdur(L.n_pdur + 1);
  for (int i = 0; i < L.n_node_values; ++i) nodes[i] = cx.T.node_src[i] != 0xFFFF ? x[cx.T.node_src[i]] : 0.0;
  if (RT) {   // stage D
    for (int f = 0; f < L.n_rt_feet; ++f) {
      const RtFoot ft = cx.T.rt_feet[f];
      double sum = 0.0;
      for (int i = 0; i < ft.n_phases - 1; ++i) { pdur[ft.phase0 + i] = x[ft.col0 + i]; sum += x[ft.col0 + i]; }
      pdur[ft.phase0 + ft.n_phases - 1] = ft.total - sum;
    }
    for (int poly = 0; poly < L.n_polys; ++poly) {
      const uint32_t pm = cx.T.poly_meta[poly];
      double* o = rtd.data() + poly * kRtd;
      if (pm == kPolyFixed) {
        const int di = cx.T.polys[poly] >> 16;
        o[kRtT] = cx.T.durs[3 * di]; o[kRtT2] = cx.T.durs[3 * di + 1]; o[kRtT3] = cx.T.durs[3 * di + 2];
        o[kRtR2] = cx.T.rdurs[2 * di]; o[kRtR3] = cx.T.rdurs[2 * di + 1];
        o[kRtR1] = 0.0; o[kRtR4] = 0.0; o[kRtEnd] = 0.0;
      } else {
        const RtFoot ft = cx.T.rt_feet[pm >> 24];
        const double Td = pdur[ft.phase0 + (pm & 0xFF)] / static_cast<double>((pm >> 8) & 0xFF);
        const double T2 = Td * Td, T3 = Cube(Td);
        o[kRtT] = Td; o[kRtT2] = T2; o[kRtT3] = T3; o[kRtR2] = 1.0 / T2; o[kRtR3] = 1.0 / T3;
        o[kRtR1] = 1.0 / Td; o[kRtR4] = 1.0 / Pow4(Td);
      }
    }
    for (int k = 0; k < L.n_rt_splines; ++k) {
      const RtSpline sp = cx.T.rt_splines[k];
      double acc = 0.0;
      for (int j = 0; j < sp.n_polys; ++j) {
        double* o = rtd.data() + (sp.poly0 + j) * kRtd;
        acc += o[kRtT];
        o[kRtEnd] = acc;
      }
    }
  }
  for (int i = 0; i < 3 * L.n_polys; ++i) {
    int poly = i / 3, d = i % 3;
    uint32_t pr = cx.T.polys[poly];
    int n0 = pr & 0xFFFF;
    double T, T2, T3, r2, r3;
    if (RT) {
      const double* o = rtd.data() + poly * kRtd;
      T = o[kRtT]; T2 = o[kRtT2]; T3 = o[kRtT3]; r2 = o[kRtR2]; r3 = o[kRtR3];
    } else {
      const double* du = cx.T.durs + 3 * (pr >> 16);
      const double* rd = cx.T.rdurs + 2 * (pr >> 16);
      T = du[0]; T2 = du[1]; T3 = du[2]; r2 = rd[0]; r3 = rd[1];
    }
    double p0 = nodes[n0 * 6 + d], v0 = nodes[n0 * 6 + 3 + d], p1 = nodes[n0 * 6 + 6 + d], v1 = nodes[n0 * 6 + 9 + d];
    coef[poly * 6 + d] = ExactDiv(-(3.0 * (p0 - p1) + T * (2.0 * v0 + v1)), T2, r2);
    coef[poly * 6 + 3 + d] = ExactDiv(2.0 * (p0 - p1) + T * (v0 + v1), T3, r3);
  }
  cx.nodes = nodes.data(); cx.coef = coef.data(); cx.G = G.data(); cx.g = g; cx.pc = &pc;
  cx.rtd = rtd.data(); cx.pdur = pdur.data();
  for (const SegmentRec& sg : part.segments) {
    std::fill(G.begin(), G.end(), std::nan(""));
    G[kGOne] = 1.0; G[kGMass] = pc.mass; G[kGMinusOne] = -1.0;
    for (int i = 0; i < sg.n_items; ++i) RunItem<Lockstep<1, RT, CURV>>(cx, cx.T.items[sg.item0 + i]);   // level-0 items come first
    // the device descriptors, as the kernel's stage B reads them
    for (int e = 0; e < sg.n_entries; ++e) {
      const uint32_t w = hp.entries_dev[sg.entry0 + e];
      const double second = (w & 0x10000u) ? G[((w >> 16) & 0xFFF8u) / 8] : cx.T.B[(w >> 16) / 8];
      vals[sg.entry0 + e] = G[(w & 0xFFFF) / 8] * second;
    }
  }
}

extern "C" int emu_eval(const towr_b200_spec* s, const double* x, double* g, double* vals) {
  try {
    HostProgram hp;
    BuildProgram(*s, hp);
    const bool curv = hp.curvature_dims != 0;
    if (hp.optimize_durations) { if (curv) EmuEval<true, true>(hp, s, x, g, vals); else EmuEval<true, false>(hp, s, x, g, vals); }
    else { if (curv) EmuEval<false, true>(hp, s, x, g, vals); else EmuEval<false, false>(hp, s, x, g, vals); }
    return 0;
  } catch (const std::exception& e) {
    g_err = e.what();
    return -1;
  }
}

// ---- eval_f / eval_grad_f: the NodeCost column table (cost_kernel.cu CostKernel) and the time-discretised terms
// (sampled_cost.cuh) replayed serially for one problem
static inline int atomicMin(int* p, int v) { const int o = *p; if (v < o) *p = v; return o; }
static inline int atomicMax(int* p, int v) { const int o = *p; if (v > o) *p = v; return o; }
#include "../../towr2025_b200/csrc/sampled_cost.cuh"

extern "C" int emu_eval_cost(const towr_b200_spec* s, const double* x, double* f_out, double* grad) {
  try {
    HostProgram hp;
    BuildProgram(*s, hp);
    CostTables ct;
    BuildCostTables(hp, *s, ct);
    double fsum = 0.0;
    for (int c = 0; c < ct.n_vars; ++c) {
      double g = 0.0;
      for (int k = ct.col_ptr[c]; k < ct.col_ptr[c + 1]; ++k) {
        const double w = ct.weights[k], val = x[c];
        g = g + (w * 2.0) * val;
        fsum = fsum + w * (val * val);
      }
      grad[c] = g;
    }
    *f_out = fsum;
    std::vector<SampledCost> parts;                        // one kernel launch each on the device, in term order
    BuildSampledCostParts(hp, *s, parts);
    for (const SampledCost& sc : parts) {
      const SampledView V = MakeSampledView(sc.blob.data());
      const SampledHeader& H = *V.h;
      std::vector<double> work((sc.shared_bytes - H.bytes) / sizeof(double) + 8);
      const SampledWork W = SampledCarve(H, reinterpret_cast<unsigned char*>(work.data()));
      ProblemConst pc;
      FillProblemConst(*s, pc);
      for (int i = 0; i < 6 * H.n_nodes; ++i) SampledNode(V, W, x, i);
      if (H.rt) for (int e = 0; e < H.n_ee; ++e) SampledPhaseDurations(V, W, x, e);
      for (int poly = 0; poly < H.n_polys + H.n_phase_slots; ++poly) SampledPolyRecord(V, W, poly);
      for (int i = 0; i < 3 * H.n_polys + H.n_splines; ++i) SampledCoeff(V, W, i);
      double pref[12];
      SwingReferenceB(*s, pref);
      for (int it = 0; it < H.n_items; ++it) SampledItem(V, W, pc.I, pref, it);
      for (int k = 0; k < H.n_times; ++k) SampledSummand(V, W, k);
      for (int ti = 0; ti < H.n_terms; ++ti) SampledTermSum(V, W, ti);
      for (int j = 0; j < H.n_jobs; ++j) SampledJobGradient(V, W, grad, j);
      for (int i = 0; i < H.n_comb; ++i) SampledCombine(V, W, grad, i);
      for (int ti = 0; ti < H.n_terms; ++ti) *f_out = *f_out + W.fterm[ti];
    }
    return 0;
  } catch (const std::exception& e) {
    g_err = e.what();
    return -1;
  }
}

// shared memory one CTA of the sampled-cost kernel needs for this class (cost.h: SampledSharedBytes), blob bytes in *blob
extern "C" long emu_cost_shared_bytes(const towr_b200_spec* s, int* blob) {
  try {
    HostProgram hp;
    BuildProgram(*s, hp);
    std::vector<SampledCost> parts;
    BuildSampledCostParts(hp, *s, parts);
    long worst = 0;
    if (blob) *blob = static_cast<int>(parts.size());   // number of launches
    for (const SampledCost& sc : parts) worst = std::max(worst, static_cast<long>(sc.shared_bytes));
    return worst;
  } catch (const std::exception& e) {
    g_err = e.what();
    return -1;
  }
}
