// cost.h — cost terms of a structure class (NodeCost, costs/node_cost.cc) as a column-compressed weight table.
#pragma once
#include <cstdint>
#include <vector>

#ifdef __CUDACC__
#include <cuda_runtime.h>
#endif

#include "program.h"
#include "../../include/towr_b200.h"

namespace tb {

struct CostTables {              // host side
  int n_vars = 0, n_terms = 0;
  std::vector<int> col_ptr;      // n_vars + 1
  std::vector<double> weights;   // one per (variable, node value) pair of a cost term, grouped by column, term order kept
};
struct CostDevice {
  int n_vars;
  const int* col_ptr;
  const double* weights;
};

// NlpFormulation::GetCosts over params_.costs_ (nlp_formulation.cc:605-681): the NodeCost families; EnergyCostID and
// AngMomCostID are skipped here (BuildSampledCost)
void BuildCostTables(const HostProgram& hp, const towr_b200_spec& s, CostTables& out);

// ---- time-discretised cost terms: EnergyCost (costs/energy_cost.cc), AngularMomentumCost (costs/angular_momentum_cost.cc)
// Both sample splines at t = 0, dt, ... <= T + 1e-9.  The class tables are ONE blob (header + arrays at byte offsets),
// copied to the device once and staged into shared memory by the kernel.
enum SampledKind { kSampledEnergy = 0, kSampledAngMom = 1 };
struct SampledTerm {     // one cost term
  int kind, spline0, n_splines, time0, n_samples, job0, n_jobs, pad;
  double weight, dtf;    // weight_, (dt_ > 0 ? dt_ : 1)
};
struct SampledSpline {   // a spline the term samples (force / torque of a foot, or base-ang)
  int node0, poly0, n_polys, foot, rt, is_torque;
  double scale;          // 1, or torque_weight_ for a torque spline (energy_cost.cc:84)
  double mscale;         // 2 wdt, or 2 wdt torque_weight_ (energy_cost.cc:112,121)
};
struct SampledFoot { int phase0, n_phases, col0, pad; double total; };
enum SampledJobKind { kJobNode = 0, kJobSchedule = 1 };
struct SampledJob {      // one gradient column
  int col;
  uint16_t kind, spline;         // spline: index within the term (kJobSchedule: the foot's force spline; its torque spline
  uint16_t node, deriv_dim;      //         follows it); kJobNode: node, deriv * 4 + dim; kJobSchedule: node = phase,
};                               //         deriv_dim = 1 when the torque spline contributes too
struct SampledHeader {
  int n_vars, n_ee, rt, n_terms, n_splines, n_nodes, n_polys, n_phases, n_times, n_jobs, max_items, max_samples;
  int off_terms, off_splines, off_feet, off_node_src, off_poly_T, off_poly_meta, off_poly_spline, off_times, off_jobs;
  int bytes;
  double I[9];           // the class spec's inertia (used when a call has no per-problem constants)
};
struct SampledCost {     // host side
  std::vector<unsigned char> blob;   // empty: the class has no time-discretised cost term
  size_t shared_bytes = 0;
};
constexpr int kSampledRec = 7;   // doubles per node / coefficient record in shared memory (6 used; odd stride, fewer bank conflicts)
constexpr size_t kSampledMaxShared = 200 * 1024;
inline size_t SampledSharedBytes(const SampledHeader& h) {
  const size_t blob = (static_cast<size_t>(h.bytes) + 15) & ~static_cast<size_t>(15);
  const size_t doubles = static_cast<size_t>(kSampledRec) * (h.n_nodes + h.n_polys) + h.n_polys + h.n_phases +
                         8 * static_cast<size_t>(h.max_items) + h.max_samples + 4;
  return blob + sizeof(double) * doubles + sizeof(int) * static_cast<size_t>(h.max_items);
}
// throws "unsupported" when a term's samples do not fit one SM's shared memory
void BuildSampledCost(const HostProgram& hp, const towr_b200_spec& s, SampledCost& out);
#ifdef __CUDACC__
cudaError_t LaunchCost(const CostDevice& c, int what, const double* x, double* f, double* grad, int n_problems, int max_ctas,
                       cudaStream_t stream);
// allocation-free variant: the tables go into a file-scope device array of cost_kernel.cu (no cudaMalloc / cudaFree);
// *why names the limit or call that failed
cudaError_t LaunchCostStatic(const CostTables& h, const SampledCost& sc, int device, int what, const double* x, double* f,
                             double* grad, int n_problems, cudaStream_t stream, const char** why);
// adds the time-discretised terms to f / grad written by LaunchCost (same stream, after it).  pcs: the batch's
// per-problem constants (inertia), or nullptr to use the class spec's
cudaError_t LaunchSampledCost(const unsigned char* blob_dev, size_t shared_bytes, const ProblemConst* pcs, int what,
                              const double* x, double* f, double* grad, int n_problems, int sms, cudaStream_t stream);
#endif

}  // namespace tb
