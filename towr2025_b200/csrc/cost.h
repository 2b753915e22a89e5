// cost.h — cost terms of a structure class: NodeCost (costs/node_cost.cc) as a column-compressed weight table, and the
// time-discretised terms (EnergyCost, AngularMomentumCost, EEBasePosCost) as one blob of class tables (SampledCost).
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
enum SampledKind { kSampledEnergy = 0, kSampledAngMom = 1, kSampledEEBasePos = 2 };
struct SampledTerm {     // one cost term
  int kind, spline0, n_splines, time0, n_samples, job0, n_jobs;
  int item0;             // first item of the term among all terms' items
  int item_splines;      // items per sample: one per spline (energy) or one for the whole sample (ang-mom, ee-base-pos)
  int scratch0;          // first double of the term's item arrays in the kernel's item scratch
  int rec;               // doubles kept per item: energy 4 (value, local time; 10 with optimised durations), ang-mom 8,
                         // ee-base-pos 9
  int foot, pad;         // ee-base-pos: the foot (EEBasePosCost::ee_)
  double weight, dtf;    // weight_, (dt_ > 0 ? dt_ : 1)
};
struct SampledSpline {   // a spline the term samples (force / torque of a foot, or base-ang)
  int node0, poly0, n_polys, foot, rt, is_torque;
  int ps0;               // optimised durations: first (spline, phase) slot of the spline (first sample of each phase)
  int fix0;              // class-constant durations: the spline's first entry in fix_t / fix_seg (one per sample); -1 else
  double scale;          // 1, or torque_weight_ for a torque spline (energy_cost.cc:84)
  double mscale;         // 2 wdt, or 2 wdt torque_weight_ (energy_cost.cc:112,121)
};
struct SampledFoot { int phase0, n_phases, col0, contact_at_start; double total; };
enum SampledJobKind { kJobNode = 0, kJobSchedule = 1, kJobColumn = 2 };
struct SampledJob {      // kJobNode: the columns of one node's values, col[deriv * 3 + dim] (-1: not optimised);
  int col[6];            // kJobSchedule: col[0] = the ee-schedule column
  uint16_t kind, spline; // kind: SampledJobKind | term << 8.  spline: index within the term (kJobSchedule: the foot's
  int slot;              // -1: the job adds to the gradient itself; else its sums go to W.scratch[slot + i] (i as in col[])
                         // and SampledCombine adds them in term order: columns that several terms write
  uint16_t node, aux;    // force spline; its torque spline follows it).  kJobNode: node; kJobSchedule: node = phase,
};                       // aux = 1 when the torque spline contributes too.  kJobColumn (variables that write several node
                         // values: ee-motion): col[0] = the column, aux = refs, col[1 + i] = node | deriv << 16 | dim << 20
struct SampledHeader {
  int n_vars, n_ee, rt, n_terms, n_splines, n_nodes, n_polys, n_phases, n_times, n_jobs;
  int n_phase_slots;           // (spline, phase) pairs of the splines with optimised durations
  int n_scratch, n_comb;       // doubles of deferred job sums; columns that SampledCombine finishes
  int off_comb_col, off_comb_ptr, off_comb_src;   // per such column: the column, its sources scratch[comb_src[ptr[i] .. ptr[i+1])]
  int n_items, item_doubles;   // items (spline, sample) of all terms; doubles of their arrays (rec each)
  int off_terms, off_splines, off_feet, off_node_src, off_poly_T, off_poly_meta, off_poly_spline, off_times, off_jobs;
  int off_fix_t, off_fix_seg;  // per sample of a spline with class-constant durations (SampledSpline::fix0 + sample): local
                               // time and polynomial (spline.cc:48-78); bit 15 of fix_seg: the foot is in contact
  int bytes;             // all of it is staged in shared memory
  double I[9];           // the class spec's inertia (used when a call has no per-problem constants)
  double p_ref_B[12];    // ... and its EEBasePosCost reference positions, 3 per foot
};
struct SampledCost {     // host side: the tables of one launch
  std::vector<unsigned char> blob;
  size_t shared_bytes = 0;
};
// one term of the class, in the order ifopt adds them: the EnergyCost / AngularMomentumCost entries of params_.costs_, then
// one EEBasePosCost per foot (nlp_formulation.cc:605-626)
struct SampledTermDef { int kind; double weight, dt; int foot; };
std::vector<SampledTermDef> SampledTermDefs(const HostProgram& hp, const towr_b200_spec& s);
constexpr int kSampledRec = 7;   // doubles per node / coefficient record in shared memory (6 used; odd stride, fewer bank conflicts)
constexpr int kSampledPolyRec = 9;   // doubles per polynomial record (T, T^2, T^3, T^4, their reciprocals, end time)
constexpr size_t kSampledMaxShared = 200 * 1024;
inline size_t SampledSharedBytes(const SampledHeader& h) {
  const size_t blob = (static_cast<size_t>(h.bytes) + 15) & ~static_cast<size_t>(15);
  const size_t doubles = static_cast<size_t>(kSampledRec) * (h.n_nodes + h.n_polys) + static_cast<size_t>(kSampledPolyRec) * h.n_polys +
                         h.n_phases + static_cast<size_t>(h.item_doubles) + h.n_times + 8 + h.n_scratch;
  return blob + sizeof(double) * doubles +
         sizeof(int) * (static_cast<size_t>(h.n_items) + 2 * static_cast<size_t>(h.n_polys) + static_cast<size_t>(h.n_phase_slots) + 4);
}
// shared memory per CTA up to which two CTAs of the kernel fit one SM (227 KB, 1 KB reserved per CTA)
constexpr size_t kSampledTwoCtaShared = 112 * 1024;
// tables of the terms defs[first, first + count); throws "unsupported" when they do not fit one SM's shared memory
// (size_only_ok: do not throw, the caller looks at out.shared_bytes and retries with fewer terms)
void BuildSampledCost(const HostProgram& hp, const towr_b200_spec& s, const std::vector<SampledTermDef>& defs, int first,
                      int count, SampledCost& out, bool size_only_ok = false);
// the class's terms as one launch, or - where that would leave one CTA per SM - as several launches of consecutive terms
// that each leave room for two (run in order on one stream: f and the gradient accumulate in term order either way);
// empty: the class has no time-discretised cost term
void BuildSampledCostParts(const HostProgram& hp, const towr_b200_spec& s, std::vector<SampledCost>& out);
// true when the class has EEBasePosCost terms (params_.enable_swing_ee_base_pos_tracking with a positive weight)
inline bool HasSwingTracking(const towr_b200_spec& s) { return s.enable_swing_ee_base_pos_tracking && s.swing_ee_base_pos_tracking_weight > 0.0; }
// r_B0 = b_R_w0 (initial_ee_W - initial_base.lin.p) per foot (nlp_formulation.cc:616-620); formulation_init.cc
void SwingReferenceB(const towr_b200_spec& s, double out[12]);
#ifdef __CUDACC__
cudaError_t LaunchCost(const CostDevice& c, int what, const double* x, double* f, double* grad, int n_problems, int max_ctas,
                       cudaStream_t stream);
// allocation-free variant: the tables go into a file-scope device array of cost_kernel.cu (no cudaMalloc / cudaFree);
// *why names the limit or call that failed
cudaError_t LaunchCostStatic(const CostTables& h, const std::vector<SampledCost>& sc, int device, int what, const double* x, double* f,
                             double* grad, int n_problems, cudaStream_t stream, const char** why);
// adds the time-discretised terms to f / grad written by LaunchCost (same stream, after it).  pcs: the batch's
// per-problem constants (inertia), or nullptr to use the class spec's
// pref: per problem 12 doubles, the EEBasePosCost reference positions (SwingReferenceB), or nullptr likewise
cudaError_t LaunchSampledCost(const unsigned char* blob_dev, size_t shared_bytes, const ProblemConst* pcs, const double* pref,
                              int what, const double* x, double* f, double* grad, int n_problems, int sms, cudaStream_t stream);
#endif

}  // namespace tb
