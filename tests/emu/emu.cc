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
    const BlobLayout& L = hp.blob;
    ProblemConst pc;
    FillProblemConst(*s, pc);
    ItemCtx<Lockstep<1>> cx;
    cx.T = MakeTables(hp.blob_data.data(), L);
    std::vector<double> nodes(L.n_node_values), coef(6 * L.n_polys), G(hp.g_capacity);
    for (int i = 0; i < L.n_node_values; ++i) nodes[i] = cx.T.node_src[i] != 0xFFFF ? x[cx.T.node_src[i]] : 0.0;
    for (int i = 0; i < 3 * L.n_polys; ++i) {
      int poly = i / 3, d = i % 3;
      uint32_t pr = cx.T.polys[poly];
      int n0 = pr & 0xFFFF;
      const double* du = cx.T.durs + 3 * (pr >> 16);
      double p0 = nodes[n0 * 6 + d], v0 = nodes[n0 * 6 + 3 + d], p1 = nodes[n0 * 6 + 6 + d], v1 = nodes[n0 * 6 + 9 + d];
      const double* rd = cx.T.rdurs + 2 * (pr >> 16);
      coef[poly * 6 + d] = ExactDiv(-(3.0 * (p0 - p1) + du[0] * (2.0 * v0 + v1)), du[1], rd[0]);
      coef[poly * 6 + 3 + d] = ExactDiv(2.0 * (p0 - p1) + du[0] * (v0 + v1), du[2], rd[1]);
    }
    cx.nodes = nodes.data(); cx.coef = coef.data(); cx.G = G.data(); cx.g = g; cx.pc = &pc;
    for (const SegmentRec& sg : hp.segments) {
      std::fill(G.begin(), G.end(), std::nan(""));
      G[kGOne] = 1.0; G[kGMass] = pc.mass;
      for (int i = 0; i < sg.n_items; ++i) RunItem<Lockstep<1>>(cx, cx.T.items[sg.item0 + i]);   // level-0 items come first
      // the device descriptors, as the kernel's stage B reads them
      for (int e = 0; e < sg.n_entries; ++e) {
        const uint32_t w = hp.entries_dev[sg.entry0 + e];
        vals[sg.entry0 + e] = G[(w & 0xFFFF) / 8] * cx.T.B[(w >> 16) / 8];
      }
    }
    return 0;
  } catch (const std::exception& e) {
    g_err = e.what();
    return -1;
  }
}
