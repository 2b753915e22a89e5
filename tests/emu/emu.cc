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

// x: n_vars; g: n_rows; vals: nnz
extern "C" int emu_eval(const towr_b200_spec* s, const double* x, double* g, double* vals) {
  try {
    HostProgram hp;
    BuildProgram(*s, hp);
    DeviceProgram P{};
    P.n_vars = hp.n_vars; P.n_rows = hp.n_rows; P.nnz = hp.nnz; P.n_ee = hp.n_ee;
    P.n_node_values = hp.n_node_values; P.n_polys = (int)hp.polys.size(); P.n_segments = (int)hp.segments.size();
    P.g_capacity = hp.g_capacity;
    P.node_src = hp.node_src.data(); P.polys = hp.polys.data(); P.samples = hp.samples.data();
    P.items = hp.items.data(); P.entries = hp.entries.data(); P.B = hp.B.data(); P.segments = hp.segments.data();
    ProblemConst pc;
    FillProblemConst(*s, pc);
    std::vector<double> nodes(P.n_node_values), coef(6 * P.n_polys), G(P.g_capacity);
    for (int i = 0; i < P.n_node_values; ++i) nodes[i] = P.node_src[i] >= 0 ? x[P.node_src[i]] : 0.0;
    for (int i = 0; i < 3 * P.n_polys; ++i) {
      int poly = i / 3, d = i % 3;
      const PolyRec& pr = P.polys[poly];
      int n0 = pr.node0;
      double p0 = nodes[n0 * 6 + d], v0 = nodes[n0 * 6 + 3 + d], p1 = nodes[n0 * 6 + 6 + d], v1 = nodes[n0 * 6 + 9 + d];
      coef[poly * 6 + d] = -(3.0 * (p0 - p1) + pr.T * (2.0 * v0 + v1)) / pr.T2;
      coef[poly * 6 + 3 + d] = (2.0 * (p0 - p1) + pr.T * (v0 + v1)) / pr.T3;
    }
    ItemCtx cx;
    cx.P = &P; cx.nodes = nodes.data(); cx.coef = coef.data(); cx.G = G.data(); cx.g = g; cx.pc = &pc;
    for (int sgi = 0; sgi < P.n_segments; ++sgi) {
      const SegmentRec sg = P.segments[sgi];
      std::fill(G.begin(), G.end(), std::nan(""));
      G[kGOne] = 1.0; G[kGMass] = pc.mass;
      for (int i = 0; i < sg.n_items; ++i) RunItem(cx, P.items[sg.item0 + i]);
      for (int e = 0; e < sg.n_entries; ++e) {
        const EntryRec w = P.entries[sg.entry0 + e];
        int gi = w.w0 & 0xFFFF, gs = (w.w0 >> 16) & 0xFFF, nt = ((w.w0 >> 28) & 3) + 1;
        int bi = w.w1 & 0x3FFFFFFF, bs = w.w1 >> 31;
        double v = G[gi] * P.B[bi];
        for (int t = 1; t < nt; ++t) v += G[gi + t * gs] * P.B[bi + t * bs];
        vals[sg.entry0 + e] = v;
      }
    }
    return 0;
  } catch (const std::exception& e) {
    g_err = e.what();
    return -1;
  }
}
