// capi.cu — the extern "C" boundary declared in include/towr_b200.h.
#include <cuda_runtime.h>
#include <sched.h>

#include <cctype>
#include <cerrno>

#include <algorithm>
#include <cstdio>
#include <fstream>
#include <iomanip>
#include <map>
#include <cstdlib>
#include <cstring>
#include <exception>
#include <memory>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <vector>

#include "kernels.h"
#include "structure.h"
#include "terrain.h"
#include "trajectory.h"
#include "cost.h"

namespace tb {
void SpecDefaults(towr_b200_spec* s);
void SpecRobot(towr_b200_spec* s, int robot);
void SpecGait(towr_b200_spec* s, int combo, double total_time);
void SpecNominalStates(towr_b200_spec* s, double gx, double gy, double gyaw);
}  // namespace tb

struct towr_b200_structure {
  tb::HostProgram hp;
  towr_b200_spec spec;     // the class's defining spec (a batch may rebuild the program with a smaller G capacity)
};

namespace {
thread_local std::string g_err;

struct CudaError : std::runtime_error {
  explicit CudaError(const std::string& m) : std::runtime_error(m) {}
};
struct StructureError : std::runtime_error {
  explicit StructureError(const std::string& m) : std::runtime_error(m) {}
};

void Check(cudaError_t e, const char* what) {
  if (e != cudaSuccess) throw CudaError(std::string(what) + ": " + cudaGetErrorString(e));
}

template <class F>
int Guard(F&& f) {
  try {
    f();
    return TOWR_B200_OK;
  } catch (const CudaError& e) {
    g_err = e.what();
    return TOWR_B200_ERR_CUDA;
  } catch (const StructureError& e) {
    g_err = e.what();
    return TOWR_B200_ERR_STRUCTURE;
  } catch (const std::exception& e) {
    g_err = e.what();
    return g_err.rfind("unsupported", 0) == 0 ? TOWR_B200_ERR_UNSUPPORTED : TOWR_B200_ERR_ARG;
  } catch (...) {
    g_err = "unknown error";
    return TOWR_B200_ERR_ARG;
  }
}

template <class T>
T* Upload(const std::vector<T>& v) {
  T* d = nullptr;
  size_t bytes = sizeof(T) * (v.empty() ? 1 : v.size());
  Check(cudaMalloc(&d, bytes), "cudaMalloc(table)");
  if (!v.empty()) Check(cudaMemcpy(d, v.data(), sizeof(T) * v.size(), cudaMemcpyHostToDevice), "cudaMemcpy(table)");
  return d;
}
}  // namespace

struct towr_b200_batch {
  const towr_b200_structure* st = nullptr;
  int n = 0, device = 0;
  cudaStream_t stream = nullptr, stream2 = nullptr;   // host-buffer evals alternate their chunks between the two
  cudaEvent_t consts_ready = nullptr;
  // eval_device launches on caller streams read pc_dev / write status_dev: the last one's completion, so that a later
  // H2D of the constants or a status read-out is ordered after it
  cudaEvent_t device_eval_done = nullptr;
  bool device_eval_pending = false;
  tb::DeviceProgram dp{};                 // part 0 (sizes, flags); classes with large tables have further parts
  std::vector<tb::DeviceProgram> dps;     // one per ProgramPart (pack.cc): one kernel launch each per evaluation
  // evaluates problems 0..n-1 of the given device arrays: every part's kernel, in order, on `s`
  cudaError_t LaunchAll(const tb::ProblemConst* pcs, int what, const double* x, double* g, double* vals, int* status, int n,
                        cudaStream_t s) {
    for (const tb::DeviceProgram& part : dps) {
      cudaError_t e = tb::LaunchEval(part, pcs, what, x, g, vals, status, n, var, max_ctas, s);
      if (e != cudaSuccess) return e;
      launches += 1;
    }
    return cudaSuccess;
  }
  std::vector<void*> owned;            // device allocations
  std::vector<tb::ProblemConst> pc_host;
  tb::ProblemConst* pc_dev = nullptr;
  bool pc_dirty = true;
  int* status_dev = nullptr;
  // staging for the host-buffer path
  double *x_dev = nullptr, *g_dev = nullptr, *v_dev = nullptr;
  tb::ProblemConst* pc_gather_dev = nullptr;
  int max_ctas = 0;
  tb::EvalVariant var{0, 1, false, 0};   // kernel variant (kernels.h); lanes = 0 until chosen
  std::unique_ptr<tb::HostProgram> hp_local;   // program rebuilt with a smaller G capacity (large class tables)
  long long launches = 0;
  // trajectory sampler tables per sampling period (trajectory.h)
  struct Traj {                        // owns its device tables (freed when the entry is evicted or the batch dies)
    tb::TrajTables host;
    tb::TrajDevice dev;
    std::vector<void*> owned;
    ~Traj() { for (void* p : owned) cudaFree(p); }
  };
  static constexpr size_t kMaxTrajTables = 8;   // sampling periods cached per batch
  std::map<double, std::unique_ptr<Traj>> traj;
  double* traj_out_dev = nullptr;
  size_t traj_out_cap = 0;
  // cost tables (cost.h), built at the first eval_cost call
  bool cost_ready = false;
  tb::CostDevice cost_dev{};
  // EnergyCost / AngularMomentumCost / EEBasePosCost tables (cost.h): one entry per launch, in term order
  struct SampledPart { const unsigned char* dev; size_t shared; };
  std::vector<SampledPart> sampled;
  // EEBasePosCost reference positions per problem (cost.h: SwingReferenceB), kept only by classes with that cost
  std::vector<double> pref_host;
  double* pref_dev = nullptr;
  double *cost_f_dev = nullptr, *cost_grad_dev = nullptr;

  void SyncConsts() {
    if (!pc_dirty) return;
    if (device_eval_pending) Check(cudaStreamWaitEvent(stream, device_eval_done, 0), "order consts after eval_device");
    Check(cudaMemcpyAsync(pc_dev, pc_host.data(), sizeof(tb::ProblemConst) * n, cudaMemcpyHostToDevice, stream), "H2D consts");
    if (pref_dev) Check(cudaMemcpyAsync(pref_dev, pref_host.data(), sizeof(double) * pref_host.size(), cudaMemcpyHostToDevice, stream), "H2D consts");
    pc_dirty = false;
  }
  ~towr_b200_batch() {
    cudaSetDevice(device);
    for (void* p : owned) cudaFree(p);
    if (traj_out_dev) cudaFree(traj_out_dev);
    traj.clear();
    if (consts_ready) cudaEventDestroy(consts_ready);
    if (device_eval_done) cudaEventDestroy(device_eval_done);
    if (stream2) cudaStreamDestroy(stream2);
    if (stream) cudaStreamDestroy(stream);
  }
};

extern "C" {

const char* towr_b200_last_error(void) { return g_err.c_str(); }

int towr_b200_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

int towr_b200_spec_defaults(towr_b200_spec* s) { return Guard([&] { if (!s) throw std::runtime_error("null spec"); tb::SpecDefaults(s); }); }
int towr_b200_spec_robot(towr_b200_spec* s, int robot) { return Guard([&] { if (!s) throw std::runtime_error("null spec"); tb::SpecRobot(s, robot); }); }
int towr_b200_spec_gait(towr_b200_spec* s, int combo, double T) { return Guard([&] { if (!s) throw std::runtime_error("null spec"); tb::SpecGait(s, combo, T); }); }
int towr_b200_spec_nominal_states(towr_b200_spec* s, double gx, double gy, double gyaw) {
  return Guard([&] { if (!s) throw std::runtime_error("null spec"); tb::SpecNominalStates(s, gx, gy, gyaw); });
}
int towr_b200_spec_optimize_durations(towr_b200_spec* s) {
  return Guard([&] {
    if (!s) throw std::runtime_error("null spec");
    if (s->n_constraints >= TOWR_B200_MAX_CONSTRAINTS) throw std::runtime_error("constraint list full");
    s->constraints[s->n_constraints++] = TOWR_B200_C_TOTAL_TIME;
  });
}

int towr_b200_structure_create(const towr_b200_spec* s, towr_b200_structure** out) {
  return Guard([&] {
    if (!s || !out) throw std::runtime_error("null argument");
    auto st = std::make_unique<towr_b200_structure>();
    tb::BuildProgram(*s, st->hp);
    st->spec = *s;
    *out = st.release();
  });
}
void towr_b200_structure_destroy(towr_b200_structure* st) { delete st; }
int towr_b200_structure_n_vars(const towr_b200_structure* st) { return st->hp.n_vars; }
int towr_b200_structure_n_rows(const towr_b200_structure* st) { return st->hp.n_rows; }
int towr_b200_structure_nnz(const towr_b200_structure* st) { return st->hp.nnz; }
int towr_b200_structure_n_var_sets(const towr_b200_structure* st) { return static_cast<int>(st->hp.var_sets.size()); }
int towr_b200_structure_n_con_sets(const towr_b200_structure* st) { return static_cast<int>(st->hp.con_sets.size()); }
const char* towr_b200_structure_var_set_name(const towr_b200_structure* st, int i) { return st->hp.var_sets.at(i).name.c_str(); }
int towr_b200_structure_var_set_rows(const towr_b200_structure* st, int i) { return st->hp.var_sets.at(i).n_vars; }
const char* towr_b200_structure_con_set_name(const towr_b200_structure* st, int i) { return st->hp.con_sets.at(i).name.c_str(); }
int towr_b200_structure_con_set_rows(const towr_b200_structure* st, int i) { return st->hp.con_sets.at(i).n_rows; }
int towr_b200_structure_pattern(const towr_b200_structure* st, int* rows, int* cols) {
  return Guard([&] {
    if (!st) throw std::runtime_error("null structure");
    if (rows) std::memcpy(rows, st->hp.pat_rows.data(), sizeof(int) * st->hp.nnz);
    if (cols) std::memcpy(cols, st->hp.pat_cols.data(), sizeof(int) * st->hp.nnz);
  });
}
int towr_b200_structure_accepts(const towr_b200_structure* st, const towr_b200_spec* s) {
  return Guard([&] {
    if (!st || !s) throw std::runtime_error("null argument");
    tb::ValidateSpec(*s);
    if (tb::TerrainCurvatureDims(s->terrain_id) & ~st->hp.curvature_dims)
      throw StructureError("the problem's terrain has curvature (Gap) and this structure class does not declare the curvature block: "
                           "create the class from a spec on that terrain");
    if (s->terrain_id == TOWR_B200_TERRAIN_TABLE && st->hp.curvature_dims == 0)
      throw StructureError("the problem's terrain is given as a table and this structure class was created on an analytic terrain: "
                           "create the class from a spec with a table terrain (or on Gap)");
    if (tb::StructureSignature(*s, st->hp.curvature_dims) != st->hp.signature)
      throw StructureError("spec belongs to a different structure class");
  });
}

int towr_b200_problem_x0(const towr_b200_structure* st, const towr_b200_spec* s, double* x) {
  return Guard([&] {
    if (towr_b200_structure_accepts(st, s) != 0) throw StructureError(g_err);
    tb::ProblemX0(st->hp, *s, x);
  });
}
int towr_b200_problem_var_bounds(const towr_b200_structure* st, const towr_b200_spec* s, double* lo, double* up) {
  return Guard([&] {
    if (towr_b200_structure_accepts(st, s) != 0) throw StructureError(g_err);
    tb::ProblemVarBounds(st->hp, *s, lo, up);
  });
}
int towr_b200_problem_con_bounds(const towr_b200_structure* st, const towr_b200_spec* s, double* lo, double* up) {
  return Guard([&] {
    if (towr_b200_structure_accepts(st, s) != 0) throw StructureError(g_err);
    tb::ProblemConBounds(st->hp, *s, lo, up);
  });
}

int towr_b200_batch_create(const towr_b200_structure* st, int n_problems, int device, towr_b200_batch** out) {
  return Guard([&] {
    if (!st || !out || n_problems < 1) throw std::runtime_error("bad argument to batch_create");
    if (towr_b200_device_count() < 1) throw CudaError("no CUDA device visible: this backend has no CPU fallback");
    Check(cudaSetDevice(device), "cudaSetDevice");
    auto b = std::make_unique<towr_b200_batch>();
    b->st = st;
    b->n = n_problems;
    b->device = device;
    Check(cudaStreamCreateWithFlags(&b->stream, cudaStreamNonBlocking), "cudaStreamCreate");
    Check(cudaStreamCreateWithFlags(&b->stream2, cudaStreamNonBlocking), "cudaStreamCreate");
    Check(cudaEventCreateWithFlags(&b->consts_ready, cudaEventDisableTiming), "cudaEventCreate");
    Check(cudaEventCreateWithFlags(&b->device_eval_done, cudaEventDisableTiming), "cudaEventCreate");
    cudaDeviceProp prop;
    Check(cudaGetDeviceProperties(&prop, device), "cudaGetDeviceProperties");
    // The class tables, one problem's spline arrays and one G buffer must fit the SM's shared memory.  Classes with many
    // samples have large tables: rebuild the program with the G capacity that is left (more, smaller segments).
    const tb::HostProgram* hpp = &st->hp;
    const long long optin = static_cast<long long>(prop.sharedMemPerBlockOptin);
    auto need = [](const tb::HostProgram& h, int g_cap, const tb::EvalVariant& v) {
      long long worst = 0;
      for (const tb::ProgramPart& part : h.parts) {
        tb::DeviceProgram t{};
        t.blob = part.blob;
        t.g_capacity = g_cap;
        worst = std::max(worst, static_cast<long long>(tb::EvalSharedBytes(t, v.lanes, v.streams, v.mode())) + 1024);   // + the kernel's static shared memory
      }
      return worst;
    };
    // rebuilds the program with at most `cap` G slots per segment (more, smaller segments); true when `v` then fits
    auto rebuild = [&](long long cap, const tb::EvalVariant& v) {
      if (cap < 256) return false;
      auto h = std::make_unique<tb::HostProgram>();
      tb::BuildProgram(st->spec, *h, static_cast<int>(cap));
      if (need(*h, h->g_capacity, v) > optin) return false;
      b->hp_local = std::move(h);
      hpp = b->hp_local.get();
      return true;
    };
    const tb::EvalVariant one_lane{1, 1, false, 0};
    if (need(*hpp, hpp->g_capacity, one_lane) > optin) {
      const long long left = optin - need(*hpp, 0, one_lane) - 2048;
      if (left < 8 * 256) throw std::runtime_error("unsupported: the class tables alone exceed the SM's shared memory");
      if (!rebuild(left / 8, one_lane))
        throw std::runtime_error("unsupported: structure class needs more shared memory per CTA than the device has");
    }
    int forced = 0;
    if (const char* env = std::getenv("TOWR_B200_LANES")) forced = std::atoi(env);
    // kernel variant (eval_kernel.cuh).  TOWR_B200_KERNEL = lockstep | ring | streams | streamsring selects one for
    // experiments; the default is chosen below.
    const char* kenv = std::getenv("TOWR_B200_KERNEL");
    tb::EvalVariant want{2, 1, false, 0};
    bool forced_variant = false;
    if (kenv && *kenv && std::strcmp(kenv, "lockstep") != 0) {
      forced_variant = true;
      if (std::strcmp(kenv, "ring") == 0) want = tb::EvalVariant{1, 1, true, 0};
      else if (std::strcmp(kenv, "streams") == 0) want = tb::EvalVariant{1, 2, false, 0};
      else if (std::strcmp(kenv, "streamsring") == 0) want = tb::EvalVariant{1, 2, true, 0};
      else if (std::strcmp(kenv, "split") == 0) want = tb::EvalVariant{2, 1, false, 0, true};
      else throw std::runtime_error(std::string("unknown TOWR_B200_KERNEL: ") + kenv);
    }
    if (const char* env = std::getenv("TOWR_B200_SKEW")) want.skew = std::atoi(env);
    // Two problems per CTA share the class tables and the descriptor stream.  Where the default program only fits one
    // lane (classes with optimised durations have large G footprints), a lower G capacity - more, smaller segments -
    // that lets two lanes fit is faster as long as it keeps about half of the default capacity (HyQ gallop with
    // optimised durations: 4.40 ms with 2 lanes / 17 segments against 6.69 ms with 1 lane / 14 segments).
    if (!std::getenv("TOWR_B200_GCAP") && !b->hp_local && (forced_variant || !forced) && need(*hpp, hpp->g_capacity, want) > optin) {
      const int default_cap = hpp->g_capacity;
      long long cap = (optin - need(*hpp, 0, want) - 6144) / (8LL * want.lanes * want.streams * (want.split ? 2 : 1));   // tables grow a little with the segment count
      for (int attempt = 0; attempt < 4 && (forced_variant || cap * 100 >= 45LL * default_cap); ++attempt, cap = cap * 31 / 32)
        if (rebuild(cap, want)) break;
    }
    const tb::HostProgram& hp = *hpp;
    tb::DeviceProgram& dp = b->dp;
    dp.n_vars = hp.n_vars; dp.n_rows = hp.n_rows; dp.nnz = hp.nnz; dp.n_ee = hp.n_ee;
    dp.n_segments = static_cast<int>(hp.segments.size());
    dp.total_time = hp.total_time;
    dp.g_capacity = hp.g_capacity;
    dp.curvature_dims = hp.curvature_dims;
    auto keep = [&](auto* p) { b->owned.push_back(const_cast<void*>(static_cast<const void*>(p))); return p; };
    dp.entries = keep(Upload(hp.entries_dev));
    for (const tb::ProgramPart& part : hp.parts) {
      tb::DeviceProgram d = dp;                            // the sizes and flags are the class's
      d.n_segments = static_cast<int>(part.segments.size());
      d.blob = part.blob;
      d.blob_data = keep(Upload(part.blob_data));
      d.segments = keep(Upload(part.segments));
      b->dps.push_back(d);
    }
    dp = b->dps.front();
    b->pc_host.assign(n_problems, tb::ProblemConst{});
    Check(cudaMalloc(&b->pc_dev, sizeof(tb::ProblemConst) * n_problems), "cudaMalloc(consts)");
    keep(b->pc_dev);
    if (tb::HasSwingTracking(st->spec)) {
      b->pref_host.assign(12 * static_cast<size_t>(n_problems), 0.0);
      Check(cudaMalloc(&b->pref_dev, sizeof(double) * b->pref_host.size()), "cudaMalloc(consts)");
      keep(b->pref_dev);
    }
    Check(cudaMalloc(&b->status_dev, sizeof(int) * n_problems), "cudaMalloc(status)");
    keep(b->status_dev);
    Check(cudaMemset(b->status_dev, 0, sizeof(int) * n_problems), "cudaMemset(status)");
    auto try_variant = [&](const tb::EvalVariant& v) {
      if (b->var.lanes || !tb::EvalVariantBuilt(dp, v)) return;
      if (need(hp, hp.g_capacity, v) > optin) return;
      int occ = 1 << 30;
      for (const tb::DeviceProgram& part : b->dps) occ = std::min(occ, tb::EvalOccupancy(part, v));
      if (occ < 1) return;
      b->var = v;
      b->max_ctas = prop.multiProcessorCount * occ;
    };
    if (forced_variant) {
      try_variant(want);
      if (!b->var.lanes) throw std::runtime_error(std::string("unsupported: kernel variant ") + kenv + " does not fit this structure class");
    }
    // lanes per CTA: as many problems side by side as shared memory allows (the class tables are held once per CTA)
    for (int lanes : {4, 2, 1}) {
      if (forced && lanes != forced) continue;
      try_variant(tb::EvalVariant{lanes, 1, false, want.skew});
    }
    if (b->var.lanes == 0)
      throw std::runtime_error("unsupported: structure class needs more shared memory per CTA than the device has");
    if (std::getenv("TOWR_B200_DEBUG"))
      std::fprintf(stderr, "batch: %d lane(s) x %d stream(s)%s, skew %d, %zu B shared (blob %d B, G capacity %d), %d CTAs, %d segments\n",
                   b->var.lanes, b->var.streams, b->var.ring ? ", bulk-store ring" : b->var.split ? ", producer/consumer split" : "", b->var.skew,
                   tb::EvalSharedBytes(dp, b->var.lanes, b->var.streams, b->var.mode()), dp.blob.bytes, dp.g_capacity, b->max_ctas,
                   dp.n_segments);
    *out = b.release();
  });
}
void towr_b200_batch_destroy(towr_b200_batch* b) { delete b; }
int towr_b200_batch_size(const towr_b200_batch* b) { return b ? b->n : 0; }

int towr_b200_batch_set_problem(towr_b200_batch* b, int i, const towr_b200_spec* s) {
  return Guard([&] {
    if (!b || i < 0 || i >= b->n) throw std::runtime_error("problem index out of range");
    if (towr_b200_structure_accepts(b->st, s) != 0) throw StructureError(g_err);
    tb::FillProblemConst(*s, b->pc_host[i]);
    if (b->pref_dev) tb::SwingReferenceB(*s, b->pref_host.data() + 12 * static_cast<size_t>(i));
    b->pc_dirty = true;
  });
}
int towr_b200_batch_set_all_problems(towr_b200_batch* b, const towr_b200_spec* s) {
  return Guard([&] {
    if (!b) throw std::runtime_error("null batch");
    if (towr_b200_structure_accepts(b->st, s) != 0) throw StructureError(g_err);
    tb::ProblemConst pc;
    tb::FillProblemConst(*s, pc);
    for (auto& p : b->pc_host) p = pc;
    if (b->pref_dev)
      for (int i = 0; i < b->n; ++i) tb::SwingReferenceB(*s, b->pref_host.data() + 12 * static_cast<size_t>(i));
    b->pc_dirty = true;
  });
}

}  // extern "C"
namespace {
// every problem of the batch on a terrain without slope: the terrain basis is a constant too
bool FlatBasis(const towr_b200_batch* b) {
  for (const tb::ProblemConst& pc : b->pc_host)
    if (tb::TerrainHasSlope(pc.terrain_id, pc.table)) return false;
  return true;
}
const std::vector<int>& DynIndex(const towr_b200_batch* b) {
  const tb::HostProgram& hp = b->st->hp;
  return FlatBasis(b) ? hp.dyn_index_flat : hp.dyn_index;
}
// Rows of the EVAL_JAC_DYNAMIC transfer: runs of x-dependent entries.  Runs separated by fewer than kMergeGap constant
// entries travel as one row: short rows make poor DMA (measured: 61 rows of 0.3-1 KB per problem reach 38 GB/s, the
// merged rows the link's rate); the constants in between are rewritten with the value the kernel stores for them, which
// is the constant itself.
struct DynRun { int pos, len; };
std::vector<DynRun> DynRuns(const towr_b200_batch* b) {
  constexpr int kMergeGap = 64;
  std::vector<DynRun> runs;
  const std::vector<int>& idx = DynIndex(b);
  for (size_t k = 0; k < idx.size();) {
    size_t e = k + 1;
    while (e < idx.size() && idx[e] - idx[e - 1] <= kMergeGap) ++e;
    runs.push_back(DynRun{idx[k], idx[e - 1] - idx[k] + 1});
    k = e;
  }
  return runs;
}
}  // namespace
extern "C" {

int towr_b200_batch_n_dynamic(const towr_b200_batch* b) { return b ? static_cast<int>(DynIndex(b).size()) : -1; }
int towr_b200_batch_dynamic_transfer_entries(const towr_b200_batch* b) {
  if (!b) return -1;
  int n = 0;
  for (const DynRun& r : DynRuns(b)) n += r.len;
  return n;
}
int towr_b200_batch_dynamic_index(const towr_b200_batch* b, int* index) {
  return Guard([&] {
    if (!b || !index) throw std::runtime_error("null argument");
    const std::vector<int>& idx = DynIndex(b);
    std::memcpy(index, idx.data(), sizeof(int) * idx.size());
  });
}
int towr_b200_batch_constant_values(const towr_b200_batch* b, int i, double* values) {
  return Guard([&] {
    if (!b || i < 0 || i >= b->n || !values) throw std::runtime_error("bad argument to batch_constant_values");
    tb::ProblemConstantValues(b->st->hp, b->pc_host[i].mass, b->st->spec.torque_k_friction, FlatBasis(b), values);
  });
}
int towr_b200_batch_eval_device(towr_b200_batch* b, int n, int what, const double* d_x, double* d_g,
                                double* d_jac_values, void* stream) {
  return Guard([&] {
    if (!b || n < 0 || n > b->n) throw std::runtime_error("bad problem count");
    if (what < 1 || what > 3) throw std::runtime_error("what must be EVAL_G, EVAL_JAC or EVAL_BOTH");
    if (!d_x || ((what & TOWR_B200_EVAL_G) && !d_g) || ((what & TOWR_B200_EVAL_JAC) && !d_jac_values))
      throw std::runtime_error("null device buffer");
    Check(cudaSetDevice(b->device), "cudaSetDevice");
    cudaStream_t s = static_cast<cudaStream_t>(stream);   // NULL = CUDA's default stream, as everywhere in the runtime API
    if (b->pc_dirty) {
      b->SyncConsts();
      Check(cudaStreamSynchronize(b->stream), "sync consts");
    }
    Check(b->LaunchAll(b->pc_dev, what, d_x, d_g, d_jac_values, b->status_dev, n, s), "launch");
    Check(cudaEventRecord(b->device_eval_done, s), "record eval_device");
    b->device_eval_pending = true;
  });
}

int towr_b200_batch_eval(towr_b200_batch* b, const int* ids, int n, int what, const double* x, double* g_out,
                         double* jac_out) {
  return Guard([&] {
    if (!b || n < 0 || n > b->n) throw std::runtime_error("bad problem count");
    const bool want_g = (what & TOWR_B200_EVAL_G) != 0, want_jac = (what & TOWR_B200_EVAL_JAC) != 0;
    const bool want_dyn = (what & TOWR_B200_EVAL_JAC_DYNAMIC) != 0;
    if (what < 1 || what > 5 || (want_jac && want_dyn))
      throw std::runtime_error("what must be EVAL_G, EVAL_JAC, EVAL_BOTH or EVAL_JAC_DYNAMIC [| EVAL_G]");
    if (!x || (want_g && !g_out) || ((want_jac || want_dyn) && !jac_out)) throw std::runtime_error("null host buffer");
    if (n == 0) return;
    Check(cudaSetDevice(b->device), "cudaSetDevice");
    const tb::DeviceProgram& dp = b->dp;
    const int kernel_what = (want_g ? TOWR_B200_EVAL_G : 0) | ((want_jac || want_dyn) ? TOWR_B200_EVAL_JAC : 0);
    auto ensure = [&](auto*& p, size_t per) {
      if (!p) {
        Check(cudaMalloc(&p, sizeof(*p) * per * b->n), "cudaMalloc(staging)");
        b->owned.push_back(p);
      }
    };
    ensure(b->x_dev, dp.n_vars);
    if (want_g) ensure(b->g_dev, dp.n_rows);
    if (want_jac || want_dyn) ensure(b->v_dev, dp.nnz);
    // EVAL_JAC_DYNAMIC: runs of consecutive x-dependent entries; each goes back with one strided copy for all problems of
    // a chunk, straight to its place in the caller's n x nnz array (whose constant entries the caller filled once)
    std::vector<DynRun> runs;
    if (want_dyn) runs = DynRuns(b);
    const tb::ProblemConst* pcs = b->pc_dev;
    if (ids) {
      // gather the constants of the requested problems (the callers' x rows are already in request order)
      ensure(b->pc_gather_dev, 1);
      std::vector<tb::ProblemConst> sel(n);
      for (int i = 0; i < n; ++i) {
        if (ids[i] < 0 || ids[i] >= b->n) throw std::runtime_error("problem id out of range");
        sel[i] = b->pc_host[ids[i]];
      }
      Check(cudaMemcpyAsync(b->pc_gather_dev, sel.data(), sizeof(tb::ProblemConst) * n, cudaMemcpyHostToDevice, b->stream), "H2D consts");
      Check(cudaStreamSynchronize(b->stream), "sync consts");
      pcs = b->pc_gather_dev;
    } else {
      b->SyncConsts();
    }
    Check(cudaEventRecord(b->consts_ready, b->stream), "record consts");
    Check(cudaStreamWaitEvent(b->stream2, b->consts_ready, 0), "wait consts");
    // chunks alternate between two streams: D2H of chunk k (the long pole: PCIe) overlaps H2D + kernel of chunk k + 1
    int chunk = n;
    if (n > 1536) chunk = 1024;
    if (const char* env = std::getenv("TOWR_B200_CHUNK")) chunk = std::max(1, std::atoi(env));
    int k = 0;
    for (int c0 = 0; c0 < n; c0 += chunk, ++k) {
      const int m = std::min(chunk, n - c0);
      cudaStream_t st = (k & 1) ? b->stream2 : b->stream;
      const size_t o = static_cast<size_t>(c0);
      Check(cudaMemcpyAsync(b->x_dev + o * dp.n_vars, x + o * dp.n_vars, sizeof(double) * dp.n_vars * m, cudaMemcpyHostToDevice, st), "H2D x");
      Check(b->LaunchAll(pcs + c0, kernel_what, b->x_dev + o * dp.n_vars, want_g ? b->g_dev + o * dp.n_rows : nullptr,
                         (want_jac || want_dyn) ? b->v_dev + o * dp.nnz : nullptr, b->status_dev + c0, m, st), "launch");
      if (want_g)
        Check(cudaMemcpyAsync(g_out + o * dp.n_rows, b->g_dev + o * dp.n_rows, sizeof(double) * dp.n_rows * m, cudaMemcpyDeviceToHost, st), "D2H g");
      if (want_jac)
        Check(cudaMemcpyAsync(jac_out + o * dp.nnz, b->v_dev + o * dp.nnz, sizeof(double) * dp.nnz * m, cudaMemcpyDeviceToHost, st), "D2H values");
      for (const DynRun& r : runs)
        Check(cudaMemcpy2DAsync(jac_out + o * dp.nnz + r.pos, sizeof(double) * dp.nnz, b->v_dev + o * dp.nnz + r.pos, sizeof(double) * dp.nnz,
                                sizeof(double) * r.len, m, cudaMemcpyDeviceToHost, st), "D2H dynamic values");
    }
    Check(cudaStreamSynchronize(b->stream), "eval");
    Check(cudaStreamSynchronize(b->stream2), "eval");
  });
}

int towr_b200_batch_status(towr_b200_batch* b, int n, int* status_out) {
  return Guard([&] {
    if (!b || n < 0 || n > b->n || !status_out) throw std::runtime_error("bad argument");
    Check(cudaSetDevice(b->device), "cudaSetDevice");
    // ordered after every eval that may still be writing the flags: host-buffer evals have synchronised already, the
    // last eval_device launch (on a caller stream) is waited for through its event
    if (b->device_eval_pending) Check(cudaStreamWaitEvent(b->stream, b->device_eval_done, 0), "order status after eval_device");
    Check(cudaMemcpyAsync(status_out, b->status_dev, sizeof(int) * n, cudaMemcpyDeviceToHost, b->stream), "D2H status");
    Check(cudaMemsetAsync(b->status_dev, 0, sizeof(int) * n, b->stream), "reset status");
    Check(cudaStreamSynchronize(b->stream), "status");
    b->device_eval_pending = false;
  });
}

// ---- eval_f / eval_grad_f --------------------------------------------------------------------------------------------
}  // extern "C"
namespace {
void EnsureCost(towr_b200_batch* b) {
  if (b->cost_ready) return;
  tb::CostTables h;
  tb::BuildCostTables(b->st->hp, b->st->spec, h);
  auto up = [&](const auto& v) {
    using T = typename std::decay<decltype(v)>::type::value_type;
    T* d = nullptr;
    Check(cudaMalloc(&d, sizeof(T) * std::max<size_t>(1, v.size())), "cudaMalloc(cost tables)");
    b->owned.push_back(d);
    if (!v.empty()) Check(cudaMemcpy(d, v.data(), sizeof(T) * v.size(), cudaMemcpyHostToDevice), "H2D cost tables");
    return d;
  };
  b->cost_dev.n_vars = h.n_vars;
  b->cost_dev.col_ptr = up(h.col_ptr);
  b->cost_dev.weights = up(h.weights);
  std::vector<tb::SampledCost> parts;
  tb::BuildSampledCostParts(b->st->hp, b->st->spec, parts);
  for (const tb::SampledCost& sc : parts) b->sampled.push_back({up(sc.blob), sc.shared_bytes});
  b->cost_ready = true;
}
}  // namespace
extern "C" {

int towr_b200_batch_eval_cost_device(towr_b200_batch* b, int n, int what, const double* d_x, double* d_f, double* d_grad,
                                     void* stream) {
  return Guard([&] {
    if (!b || n < 0 || n > b->n) throw std::runtime_error("bad problem count");
    if (what < 1 || what > 3) throw std::runtime_error("what must be 1 (f), 2 (gradient) or 3 (both)");
    if (!d_x || ((what & 1) && !d_f) || ((what & 2) && !d_grad)) throw std::runtime_error("null device buffer");
    Check(cudaSetDevice(b->device), "cudaSetDevice");
    EnsureCost(b);
    int sms = 0;
    Check(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, b->device), "device attribute");
    Check(tb::LaunchCost(b->cost_dev, what, d_x, d_f, d_grad, n, 16 * sms, static_cast<cudaStream_t>(stream)), "launch");
    b->launches += 1;
    if (!b->sampled.empty()) {
      if (b->pc_dirty) {                        // per-problem inertia: the batch's constants
        b->SyncConsts();
        Check(cudaStreamSynchronize(b->stream), "sync consts");
      }
      for (const auto& part : b->sampled) {
        Check(tb::LaunchSampledCost(part.dev, part.shared, b->pc_dev, b->pref_dev, what, d_x, d_f, d_grad, n, sms,
                                    static_cast<cudaStream_t>(stream)), "launch");
        b->launches += 1;
      }
      // the kernel reads the batch's constants on the caller's stream: a later set_problem must not overwrite them first
      Check(cudaEventRecord(b->device_eval_done, static_cast<cudaStream_t>(stream)), "record eval_cost_device");
      b->device_eval_pending = true;
    }
  });
}

int towr_b200_batch_eval_cost(towr_b200_batch* b, int n, int what, const double* x, double* f_out, double* grad_out) {
  return Guard([&] {
    if (!b || n < 0 || n > b->n) throw std::runtime_error("bad problem count");
    if (what < 1 || what > 3) throw std::runtime_error("what must be 1 (f), 2 (gradient) or 3 (both)");
    if (!x || ((what & 1) && !f_out) || ((what & 2) && !grad_out)) throw std::runtime_error("null host buffer");
    if (n == 0) return;
    Check(cudaSetDevice(b->device), "cudaSetDevice");
    EnsureCost(b);
    auto ensure = [&](double*& p, size_t per) {
      if (!p) {
        Check(cudaMalloc(&p, sizeof(double) * per * b->n), "cudaMalloc(staging)");
        b->owned.push_back(p);
      }
    };
    ensure(b->x_dev, b->dp.n_vars);
    ensure(b->cost_f_dev, 1);
    if (what & 2) ensure(b->cost_grad_dev, b->dp.n_vars);
    Check(cudaMemcpyAsync(b->x_dev, x, sizeof(double) * b->dp.n_vars * n, cudaMemcpyHostToDevice, b->stream), "H2D x");
    int sms = 0;
    Check(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, b->device), "device attribute");
    Check(tb::LaunchCost(b->cost_dev, what, b->x_dev, b->cost_f_dev, b->cost_grad_dev, n, 16 * sms, b->stream), "launch");
    b->launches += 1;
    if (!b->sampled.empty()) b->SyncConsts();
    for (const auto& part : b->sampled) {
      Check(tb::LaunchSampledCost(part.dev, part.shared, b->pc_dev, b->pref_dev, what, b->x_dev, b->cost_f_dev,
                                  b->cost_grad_dev, n, sms, b->stream), "launch");
      b->launches += 1;
    }
    if (what & 1) Check(cudaMemcpyAsync(f_out, b->cost_f_dev, sizeof(double) * n, cudaMemcpyDeviceToHost, b->stream), "D2H f");
    if (what & 2)
      Check(cudaMemcpyAsync(grad_out, b->cost_grad_dev, sizeof(double) * b->dp.n_vars * n, cudaMemcpyDeviceToHost, b->stream), "D2H gradient");
    Check(cudaStreamSynchronize(b->stream), "eval_cost");
  });
}

// Allocation-free variant (no counterpart in the reference; see include/towr_b200.h): no batch, no cudaMalloc / cudaFree.
int towr_b200_eval_cost_static(const towr_b200_structure* st, int device, int n, int what, const double* d_x, double* d_f,
                               double* d_grad, void* stream) {
  return Guard([&] {
    if (!st || n < 0) throw std::runtime_error("bad argument to eval_cost_static");
    if (what < 1 || what > 3) throw std::runtime_error("what must be 1 (f), 2 (gradient) or 3 (both)");
    if (!d_x || ((what & 1) && !d_f) || ((what & 2) && !d_grad)) throw std::runtime_error("null device buffer");
    if (towr_b200_device_count() < 1) throw CudaError("no CUDA device visible: this backend has no CPU fallback");
    tb::CostTables h;                                   // host tables only
    tb::BuildCostTables(st->hp, st->spec, h);
    std::vector<tb::SampledCost> sc;
    tb::BuildSampledCostParts(st->hp, st->spec, sc);
    const char* why = "";
    const cudaError_t err = tb::LaunchCostStatic(h, sc, device, what, d_x, d_f, d_grad, n, static_cast<cudaStream_t>(stream), &why);
    if (err == cudaErrorInvalidValue && *why) throw std::runtime_error(std::string("eval_cost_static: ") + why);
    if (err != cudaSuccess) throw CudaError(std::string("eval_cost_static") + (*why ? std::string(" (") + why + ")" : std::string()) + ": " + cudaGetErrorString(err));
  });
}

// ---- trajectory sampling -------------------------------------------------------------------------------------------
int towr_b200_trajectory_width(const towr_b200_structure* st) { return st ? tb::TrajectoryWidth(st->hp.n_ee) : -1; }

int towr_b200_trajectory_samples(const towr_b200_structure* st, double dt) {
  int n = -1;
  const int rc = Guard([&] {
    if (!st) throw std::runtime_error("null structure");
    n = static_cast<int>(tb::TrajectoryTimes(st->hp, dt).size());
  });
  return rc == TOWR_B200_OK ? n : -1;
}

}  // extern "C"

namespace {
template <class T>
const T* UploadVec(std::vector<void*>& owned, const std::vector<T>& v) {
  if (v.empty()) return nullptr;
  T* d = nullptr;
  Check(cudaMalloc(&d, sizeof(T) * v.size()), "cudaMalloc(trajectory tables)");
  owned.push_back(d);
  Check(cudaMemcpy(d, v.data(), sizeof(T) * v.size(), cudaMemcpyHostToDevice), "H2D trajectory tables");
  return d;
}

// uploads t->host (filled by the caller) and checks that the sampler's tables fit an SM
void UploadTraj(towr_b200_batch* b, towr_b200_batch::Traj* t) {
  const tb::TrajTables& h = t->host;
  tb::TrajDevice& d = t->dev;
  d.n_vars = h.n_vars; d.n_ee = h.n_ee; d.width = h.width; d.n_samples = h.n_samples;
  d.n_splines = static_cast<int>(h.splines.size());
  d.n_nodes = h.n_nodes; d.n_polys = h.n_polys; d.n_phases = h.n_phases; d.rt = h.rt ? 1 : 0;
  d.total_time = h.total_time;
  d.times = UploadVec(t->owned, h.times);
  d.splines = UploadVec(t->owned, h.splines);
  d.feet = UploadVec(t->owned, h.feet);
  d.node_src = UploadVec(t->owned, h.node_src);
  d.poly_T = UploadVec(t->owned, h.poly_T);
  d.poly_meta = UploadVec(t->owned, h.poly_meta);
  d.poly_spline = UploadVec(t->owned, h.poly_spline);
  d.samp_poly = UploadVec(t->owned, h.samp_poly);
  d.samp_t = UploadVec(t->owned, h.samp_t);
  d.contact = UploadVec(t->owned, h.contact);
  cudaDeviceProp prop;
  Check(cudaGetDeviceProperties(&prop, b->device), "cudaGetDeviceProperties");
  if (tb::TrajSharedBytes(d) > prop.sharedMemPerBlockOptin)
    throw std::runtime_error("unsupported: the problem's spline tables exceed the SM's shared memory");
}

void SampleHost(towr_b200_batch* b, const towr_b200_batch::Traj& t, int n, const double* x, double* rows_out);

const towr_b200_batch::Traj& TrajFor(towr_b200_batch* b, double dt) {
  auto it = b->traj.find(dt);
  if (it != b->traj.end()) return *it->second;
  auto t = std::make_unique<towr_b200_batch::Traj>();
  tb::BuildTrajTables(b->st->hp, b->st->spec, dt, t->host);
  UploadTraj(b, t.get());
  if (b->traj.size() >= towr_b200_batch::kMaxTrajTables) {
    // evict one cached sampling period (its kernels have completed: the sampler entry points synchronise or run on a
    // caller stream the caller is responsible for; be safe and drain the device first)
    Check(cudaDeviceSynchronize(), "sync before evicting trajectory tables");
    b->traj.erase(b->traj.begin());
  }
  return *(b->traj[dt] = std::move(t));
}
}  // namespace

extern "C" {

int towr_b200_batch_sample_trajectory_device(towr_b200_batch* b, int n, double dt, const double* d_x, double* d_rows,
                                             void* stream) {
  return Guard([&] {
    if (!b || n < 0 || n > b->n) throw std::runtime_error("bad problem count");
    if (!d_x || !d_rows) throw std::runtime_error("null device buffer");
    Check(cudaSetDevice(b->device), "cudaSetDevice");
    const towr_b200_batch::Traj& t = TrajFor(b, dt);
    int sms = 0;
    Check(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, b->device), "device attribute");
    Check(tb::LaunchSampleTrajectory(t.dev, d_x, d_rows, n, sms, static_cast<cudaStream_t>(stream)), "launch");
    b->launches += 1;
  });
}

int towr_b200_batch_sample_trajectory(towr_b200_batch* b, int n, double dt, const double* x, double* rows_out) {
  return Guard([&] {
    if (!b || n < 0 || n > b->n) throw std::runtime_error("bad problem count");
    if (!x || !rows_out) throw std::runtime_error("null host buffer");
    if (n == 0) return;
    Check(cudaSetDevice(b->device), "cudaSetDevice");
    SampleHost(b, TrajFor(b, dt), n, x, rows_out);
  });
}

int towr_b200_batch_sample_points(towr_b200_batch* b, int n, const double* times, int n_times, const double* x,
                                  double* rows_out) {
  return Guard([&] {
    if (!b || n < 0 || n > b->n) throw std::runtime_error("bad problem count");
    if (!times || n_times < 0 || !x || !rows_out) throw std::runtime_error("null host buffer");
    if (n == 0 || n_times == 0) return;
    Check(cudaSetDevice(b->device), "cudaSetDevice");
    towr_b200_batch::Traj t;                             // not cached: the time list is the caller's
    tb::BuildTrajTablesAt(b->st->hp, b->st->spec, std::vector<double>(times, times + n_times), t.host);
    UploadTraj(b, &t);
    SampleHost(b, t, n, x, rows_out);                    // synchronises before t's tables are freed
  });
}

}  // extern "C"
namespace {
void SampleHost(towr_b200_batch* b, const towr_b200_batch::Traj& t, int n, const double* x, double* rows_out) {
  {
    const size_t per = static_cast<size_t>(t.dev.n_samples) * t.dev.width;
    if (!b->x_dev) {
      Check(cudaMalloc(&b->x_dev, sizeof(double) * b->dp.n_vars * b->n), "cudaMalloc(staging)");
      b->owned.push_back(b->x_dev);
    }
    if (b->traj_out_cap < per * n) {
      if (b->traj_out_dev) {                               // superseded: the previous call synchronised, nothing reads it
        Check(cudaFree(b->traj_out_dev), "cudaFree(trajectory rows)");
        b->traj_out_dev = nullptr;
        b->traj_out_cap = 0;
      }
      double* p = nullptr;
      Check(cudaMalloc(&p, sizeof(double) * per * b->n), "cudaMalloc(trajectory rows)");
      b->traj_out_dev = p;
      b->traj_out_cap = per * b->n;
    }
    Check(cudaMemcpyAsync(b->x_dev, x, sizeof(double) * b->dp.n_vars * n, cudaMemcpyHostToDevice, b->stream), "H2D x");
    int sms = 0;
    Check(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, b->device), "device attribute");
    Check(tb::LaunchSampleTrajectory(t.dev, b->x_dev, b->traj_out_dev, n, sms, b->stream), "launch");
    b->launches += 1;
    Check(cudaMemcpyAsync(rows_out, b->traj_out_dev, sizeof(double) * per * n, cudaMemcpyDeviceToHost, b->stream), "D2H rows");
    Check(cudaStreamSynchronize(b->stream), "sample_trajectory");
  }
}
}  // namespace
extern "C" {

int towr_b200_trajectory_write_csv(const towr_b200_structure* st, const double* rows, int n_rows, const char* filename) {
  return Guard([&] {
    if (!st || !rows || n_rows < 0 || !filename) throw std::runtime_error("bad argument");
    const int E = st->hp.n_ee, W = tb::TrajectoryWidth(E);
    std::ofstream f(filename);
    if (!f.is_open()) throw std::runtime_error(std::string("cannot create ") + filename);
    f << std::fixed << std::setprecision(6);
    f << "time,base_pos_x,base_pos_y,base_pos_z,base_vel_x,base_vel_y,base_vel_z,base_acc_x,base_acc_y,base_acc_z,"
      << "base_euler_roll,base_euler_pitch,base_euler_yaw,base_omega_x,base_omega_y,base_omega_z,"
      << "base_omegadot_x,base_omegadot_y,base_omegadot_z";
    static const char* kCols[8] = {"ee_pos", "ee_vel", "ee_acc", "ee_euler", "ee_omega", "ee_omegadot", "contact_force", "contact_torque"};
    static const char* kAxes[2][3] = {{"x", "y", "z"}, {"roll", "pitch", "yaw"}};
    for (int i = 0; i < E; ++i) {
      for (int c = 0; c < 8; ++c)
        for (int d = 0; d < 3; ++d) f << "," << kCols[c] << "_" << kAxes[c == 3 ? 1 : 0][d] << "_" << i;
      f << ",is_contact_phase_" << i;
    }
    f << std::endl;
    for (int k = 0; k < n_rows; ++k) {
      const double* r = rows + static_cast<size_t>(k) * W;
      for (int c = 0; c < W; ++c) {
        if (c) f << ",";
        if (c >= 19 && (c - 19) % 25 == 24) f << (r[c] != 0.0 ? 1 : 0);   // is_contact is written as an integer
        else f << r[c];
      }
      f << std::endl;
    }
  });
}

int towr_b200_host_alloc(size_t bytes, void** out) {
  return Guard([&] {
    if (!out) throw std::runtime_error("null argument");
    *out = nullptr;
    if (towr_b200_device_count() < 1) throw CudaError("no CUDA device visible");
    Check(cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocPortable), "cudaHostAlloc");
  });
}
void towr_b200_host_free(void* p) {
  if (p) cudaFreeHost(p);
}
int towr_b200_device_numa_node(int device) {
  char bus[32] = {0};
  if (cudaDeviceGetPCIBusId(bus, sizeof(bus), device) != cudaSuccess) { cudaGetLastError(); return -1; }
  for (char* c = bus; *c; ++c) *c = static_cast<char>(std::tolower(static_cast<unsigned char>(*c)));
  const std::string path = std::string("/sys/bus/pci/devices/") + bus + "/numa_node";
  int node = -1;
  if (FILE* f = std::fopen(path.c_str(), "r")) {
    if (std::fscanf(f, "%d", &node) != 1) node = -1;
    std::fclose(f);
  }
  return node;
}
int towr_b200_bind_thread_to_device_node(int device) {
  int result = 0;
  const int rc = Guard([&] {
    result = [&]() -> int {
    const int node = towr_b200_device_numa_node(device);
    if (node < 0) return 0;
    const std::string path = "/sys/devices/system/node/node" + std::to_string(node) + "/cpulist";
    FILE* f = std::fopen(path.c_str(), "r");
    if (!f) return 0;
    char buf[4096] = {0};
    const bool ok = std::fgets(buf, sizeof(buf), f) != nullptr;
    std::fclose(f);
    if (!ok) return 0;
    cpu_set_t set;
    CPU_ZERO(&set);
    int n = 0;
    for (char* tok = std::strtok(buf, ",\n"); tok; tok = std::strtok(nullptr, ",\n")) {   // "0-15,32-47"
      int a = 0, b = 0;
      const int got = std::sscanf(tok, "%d-%d", &a, &b);
      if (got < 1) continue;
      if (got == 1) b = a;
      for (int c = a; c <= b && c < CPU_SETSIZE; ++c) { CPU_SET(c, &set); ++n; }
    }
    if (n == 0) return 0;
    if (sched_setaffinity(0, sizeof(set), &set) != 0) throw std::runtime_error(std::string("sched_setaffinity: ") + std::strerror(errno));
    return n;
    }();
  });
  return rc != TOWR_B200_OK ? rc : result;
}
long long towr_b200_batch_launch_count(const towr_b200_batch* b) { return b ? b->launches : 0; }
int towr_b200_batch_kernel_info(const towr_b200_batch* b, char* buf, int n) {
  return Guard([&] {
    if (!b || !buf || n < 1) throw std::runtime_error("bad argument");
    std::snprintf(buf, static_cast<size_t>(n), "tb::EvalKernel%s<lanes=%d, optimised_durations=%d, streams=%d, ring=%d, curvature=%d>: "
                  "%d CTAs x %d threads, %zu B shared, %d segments", b->var.split ? "Split" : "", b->var.lanes, b->dp.blob.rt,
                  b->var.streams, b->var.ring ? 1 : 0, b->dp.curvature_dims != 0, b->max_ctas,
                  (b->var.split ? 384 : 256) * b->var.lanes * b->var.streams,
                  tb::EvalSharedBytes(b->dp, b->var.lanes, b->var.streams, b->var.mode()), b->dp.n_segments);
    if (b->dps.size() > 1) {                               // classes with large tables: one launch per part
      const size_t used = std::strlen(buf);
      std::snprintf(buf + used, static_cast<size_t>(n) - used, " in the first of %zu parts (one launch each)", b->dps.size());
    }
  });
}
void* towr_b200_batch_stream(const towr_b200_batch* b) { return b ? static_cast<void*>(b->stream) : nullptr; }

}  // extern "C"
