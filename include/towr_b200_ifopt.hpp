// towr_b200_ifopt.hpp — header-only C++ mirror of the ifopt component surface over the C ABI of towr_b200.h.
//
// The reference evaluates one NLP through ifopt's component interface:
//   ifopt::Component::{GetName, GetRows}                    ifopt/composite.h (ifopt 2.0.1, not vendored)
//   ifopt::VariableSet::{GetValues, SetVariables, GetBounds}   e.g. towr/src/variables/nodes_variables.cc:56-122
//   ifopt::ConstraintSet::{GetValues, GetBounds, FillJacobianBlock}
//                                                           e.g. towr/src/constraints/time_discretization_constraint.cc:65-96
//   ifopt::Problem::{GetNumberOfOptimizationVariables, GetNumberOfConstraints, GetBoundsOnOptimizationVariables,
//                    GetBoundsOnConstraints, GetVariableValues, SetVariables, EvaluateConstraints, GetJacobianOfConstraints}
// The classes below keep those names, argument meaning and error behaviour (std::runtime_error), but every problem of
// a batch is evaluated by ONE launch: a view asks its BatchedProblem, which evaluates the whole batch lazily when the
// variables changed (the same caching role ifopt's Composite plays for one problem).  No Eigen dependency: vectors are
// std::vector<double>, the Jacobian is (row, col, value) triplets in ifopt's row-major storage order.
//
// There is no CPU fallback: Evaluate() throws when the CUDA backend cannot run.
#pragma once
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "towr_b200.h"
#include "towr_b200_formulation.hpp"
#include "towr_b200_spline_holder.hpp"

namespace towr_b200 {

struct Bounds {            // ifopt::Bounds (ifopt/bounds.h)
  double lower_, upper_;
};
using VecBound = std::vector<Bounds>;
using VectorXd = std::vector<double>;
struct Triplet { int row, col; double value; };

inline void Check(int rc) {
  if (rc != TOWR_B200_OK) throw std::runtime_error(towr_b200_last_error());
}

// towr::NlpFormulation + towr::Parameters, as the plain spec the backend reads (include/towr/nlp_formulation.h:100-105)
class NlpFormulation {
 public:
  NlpFormulation() { Check(towr_b200_spec_defaults(&spec_)); }
  // from the reference's own member layout (towr_b200_formulation.hpp: initial_base_, final_base_, initial_ee_W_, model_,
  // terrain_, params_), or from a finished spec
  explicit NlpFormulation(const ref::NlpFormulation& f) : spec_(ToSpec(f)) {}
  explicit NlpFormulation(const towr_b200_spec& s) : spec_(s) {}
  // towr::RobotModel(Robot) ; towr::GaitGenerator::SetCombo + GetPhaseDurations / GetContactSchedule
  void SetRobot(int robot) { Check(towr_b200_spec_robot(&spec_, robot)); }
  void SetGait(int combo, double total_time) { Check(towr_b200_spec_gait(&spec_, combo, total_time)); }
  // initial_base_, final_base_, initial_ee_W_ in the style of towr_ros_app.cc:47-117
  void SetNominalStates(double goal_x, double goal_y, double goal_yaw) {
    Check(towr_b200_spec_nominal_states(&spec_, goal_x, goal_y, goal_yaw));
  }
  void SetTerrain(int terrain_id) { spec_.terrain_id = terrain_id; }
  void OptimizePhaseDurations() { Check(towr_b200_spec_optimize_durations(&spec_)); }   // parameters.cc:107-111
  void AddConstraint(int name) {
    if (spec_.n_constraints >= TOWR_B200_MAX_CONSTRAINTS) throw std::runtime_error("too many constraints");
    spec_.constraints[spec_.n_constraints++] = name;
  }
  // params_.costs_.push_back({name, weight})   (parameters.h:163,192; hopper_example.cc:147-148)
  void AddCost(int name, double weight) {
    if (spec_.n_costs >= TOWR_B200_MAX_COSTS) throw std::runtime_error("too many costs");
    spec_.cost_ids[spec_.n_costs] = name;
    spec_.cost_weights[spec_.n_costs++] = weight;
  }
  // nlp.AddCostSet(std::make_shared<NodeCost>(nodes_id, deriv, dim, weight))   (node_cost.cc:36-45; the reference's
  // hand-built examples, test/squat_jump_example.cc:205-211); nodes: TOWR_B200_NODES_*, ee: the foot of an ee set
  void AddNodeCost(int nodes, int ee, int deriv, int dim, double weight) {
    if (spec_.n_node_costs >= TOWR_B200_MAX_NODE_COSTS) throw std::runtime_error("too many NodeCost terms");
    const int i = spec_.n_node_costs++;
    spec_.node_cost_set[i] = nodes;
    spec_.node_cost_ee[i] = ee;
    spec_.node_cost_deriv[i] = deriv;
    spec_.node_cost_dim[i] = dim;
    spec_.node_cost_weight[i] = weight;
  }
  towr_b200_spec& params() { return spec_; }
  const towr_b200_spec& params() const { return spec_; }

 private:
  towr_b200_spec spec_;
};

class BatchedProblem;

// ifopt::VariableSet of problem `p` (name, rows, current values, bounds)
class VariableSetView {
 public:
  std::string GetName() const { return name_; }
  int GetRows() const { return rows_; }
  VectorXd GetValues() const;
  void SetVariables(const VectorXd& x);
  VecBound GetBounds() const;

 private:
  friend class BatchedProblem;
  VariableSetView(BatchedProblem* b, int p, std::string name, int col0, int rows)
      : b_(b), p_(p), name_(std::move(name)), col0_(col0), rows_(rows) {}
  BatchedProblem* b_;
  int p_;
  std::string name_;
  int col0_, rows_;
};

// ifopt::ConstraintSet of problem `p`
class ConstraintSetView {
 public:
  std::string GetName() const { return name_; }
  int GetRows() const { return rows_; }
  VectorXd GetValues() const;
  VecBound GetBounds() const;
  // entries of this set's rows that fall into the columns of variable set `var_set`, rows/cols local to the two sets
  void FillJacobianBlock(const std::string& var_set, std::vector<Triplet>& jac_block) const;

 private:
  friend class BatchedProblem;
  ConstraintSetView(BatchedProblem* b, int p, std::string name, int row0, int rows)
      : b_(b), p_(p), name_(std::move(name)), row0_(row0), rows_(rows) {}
  BatchedProblem* b_;
  int p_;
  std::string name_;
  int row0_, rows_;
};

// n ifopt::Problem instances of one structure class, resident on one GPU
class BatchedProblem {
 public:
  // all problems share the formulation's structure class; per-problem specs (goals, masses, terrain ids) may differ
  BatchedProblem(const std::vector<NlpFormulation>& problems, int device = 0) : n_(static_cast<int>(problems.size())) {
    if (problems.empty()) throw std::runtime_error("empty batch");
    Check(towr_b200_structure_create(&problems.front().params(), &st_));
    n_vars_ = towr_b200_structure_n_vars(st_);
    n_rows_ = towr_b200_structure_n_rows(st_);
    nnz_ = towr_b200_structure_nnz(st_);
    rows_.resize(nnz_);
    cols_.resize(nnz_);
    Check(towr_b200_structure_pattern(st_, rows_.data(), cols_.data()));
    x_.assign(static_cast<size_t>(n_) * n_vars_, 0.0);
    specs_.reserve(n_);
    for (int p = 0; p < n_; ++p) {
      specs_.push_back(problems[p].params());
      Check(towr_b200_structure_accepts(st_, &specs_[p]));
      Check(towr_b200_problem_x0(st_, &specs_[p], x_.data() + static_cast<size_t>(p) * n_vars_));   // initial guess
    }
    int col = 0;
    for (int i = 0; i < towr_b200_structure_n_var_sets(st_); ++i) {
      var_sets_.push_back({towr_b200_structure_var_set_name(st_, i), col, towr_b200_structure_var_set_rows(st_, i)});
      col += var_sets_.back().rows;
    }
    int row = 0;
    for (int i = 0; i < towr_b200_structure_n_con_sets(st_); ++i) {
      con_sets_.push_back({towr_b200_structure_con_set_name(st_, i), row, towr_b200_structure_con_set_rows(st_, i)});
      row += con_sets_.back().rows;
    }
    device_ = device;
  }
  ~BatchedProblem() {
    if (batch_) towr_b200_batch_destroy(batch_);
    if (st_) towr_b200_structure_destroy(st_);
  }
  BatchedProblem(const BatchedProblem&) = delete;
  BatchedProblem& operator=(const BatchedProblem&) = delete;

  int GetBatchSize() const { return n_; }
  // ---- ifopt::Problem, per problem p
  int GetNumberOfOptimizationVariables() const { return n_vars_; }
  int GetNumberOfConstraints() const { return n_rows_; }
  int GetNumberOfJacobianNonzeros() const { return nnz_; }
  VectorXd GetVariableValues(int p) const { return VectorXd(xp(p), xp(p) + n_vars_); }
  void SetVariables(int p, const double* x) {
    std::copy(x, x + n_vars_, x_.begin() + static_cast<size_t>(p) * n_vars_);
    dirty_ = true;
    cost_dirty_ = true;
  }
  VecBound GetBoundsOnOptimizationVariables(int p) const {
    VectorXd lo(n_vars_), up(n_vars_);
    Check(towr_b200_problem_var_bounds(st_, &specs_[p], lo.data(), up.data()));
    return Zip(lo, up);
  }
  VecBound GetBoundsOnConstraints(int p) const {
    VectorXd lo(n_rows_), up(n_rows_);
    Check(towr_b200_problem_con_bounds(st_, &specs_[p], lo.data(), up.data()));
    return Zip(lo, up);
  }
  // eval_g: sets x of problem p, returns its g (evaluates the whole batch if anything changed)
  VectorXd EvaluateConstraints(int p, const double* x) {
    SetVariables(p, x);
    Evaluate();
    return VectorXd(g_.begin() + static_cast<size_t>(p) * n_rows_, g_.begin() + static_cast<size_t>(p + 1) * n_rows_);
  }
  // eval_jac_g: pattern (same for every problem, ifopt's row-major order) and the values of problem p
  const std::vector<int>& JacobianRows() const { return rows_; }
  const std::vector<int>& JacobianCols() const { return cols_; }
  const double* GetJacobianOfConstraints(int p) {
    Evaluate();
    return v_.data() + static_cast<size_t>(p) * nnz_;
  }
  // eval_f / eval_grad_f (ifopt::Problem::EvaluateCostFunction / EvaluateCostFunctionGradient): one launch for the
  // whole batch, like Evaluate()
  double EvaluateCostFunction(int p, const double* x) {
    SetVariables(p, x);
    EvaluateCosts();
    return f_[p];
  }
  VectorXd EvaluateCostFunctionGradient(int p, const double* x) {
    SetVariables(p, x);
    EvaluateCosts();
    return VectorXd(grad_.begin() + static_cast<size_t>(p) * n_vars_, grad_.begin() + static_cast<size_t>(p + 1) * n_vars_);
  }
  void EvaluateCosts() {
    if (!cost_dirty_ && !f_.empty()) return;
    EnsureBatch();
    f_.resize(n_);
    grad_.resize(static_cast<size_t>(n_) * n_vars_);
    Check(towr_b200_batch_eval_cost(batch_, n_, 3, x_.data(), f_.data(), grad_.data()));
    cost_dirty_ = false;
  }
  // towr::SaveTrajectoryToCSV's sampling loop (save_data.cpp:54-119) for every problem's current variables:
  // n x samples x width doubles, row layout in include/towr_b200.h.  SplineHolder::GetPoint(t) of problem p is the
  // row of sample k = t / T_sample.
  int GetTrajectoryWidth() const { return towr_b200_trajectory_width(st_); }
  int GetTrajectorySamples(double T_sample) const { return towr_b200_trajectory_samples(st_, T_sample); }
  std::vector<double> SampleTrajectories(double T_sample) {
    EnsureBatch();
    const int samples = GetTrajectorySamples(T_sample);
    if (samples < 0) throw std::runtime_error(towr_b200_last_error());
    std::vector<double> rows(static_cast<size_t>(n_) * samples * GetTrajectoryWidth());
    Check(towr_b200_batch_sample_trajectory(batch_, n_, T_sample, x_.data(), rows.data()));
    return rows;
  }
  // towr::SplineHolder of problem p's current variables (what GetVariableSets hands back in the reference,
  // nlp_formulation.cc:76-119): base_linear_, ee_motion_.at(ee), ... ->GetPoint(t), phase_durations_.at(ee)->IsContactPhase(t)
  SplineHolder GetSplineHolder(int p) {
    EnsureBatch();
    return SplineHolder(batch_, st_, specs_[p], xp(p));
  }
  bool SaveTrajectoryToCSV(int p, const std::string& filename, double T_sample) {
    const std::vector<double> rows = SampleTrajectories(T_sample);
    const size_t per = static_cast<size_t>(GetTrajectorySamples(T_sample)) * GetTrajectoryWidth();
    return towr_b200_trajectory_write_csv(st_, rows.data() + p * per, GetTrajectorySamples(T_sample), filename.c_str()) == 0;
  }
  // one launch for the whole batch (no-op when nothing changed since the last one)
  void Evaluate() {
    if (!dirty_ && !g_.empty()) return;
    EnsureBatch();
    g_.resize(static_cast<size_t>(n_) * n_rows_);
    v_.resize(static_cast<size_t>(n_) * nnz_);
    Check(towr_b200_batch_eval(batch_, nullptr, n_, TOWR_B200_EVAL_BOTH, x_.data(), g_.data(), v_.data()));
    dirty_ = false;
  }

  // ---- ifopt::Composite::GetComponent
  VariableSetView GetVariableSet(int p, const std::string& name) {
    for (const auto& s : var_sets_)
      if (s.name == name) return VariableSetView(this, p, s.name, s.first, s.rows);
    throw std::runtime_error("variable set " + name + " not found");
  }
  ConstraintSetView GetConstraintSet(int p, const std::string& name) {
    for (const auto& s : con_sets_)
      if (s.name == name) return ConstraintSetView(this, p, s.name, s.first, s.rows);
    throw std::runtime_error("constraint set " + name + " not found");
  }
  std::vector<std::string> GetVariableSetNames() const { return Names(var_sets_); }
  std::vector<std::string> GetConstraintSetNames() const { return Names(con_sets_); }

 private:
  friend class VariableSetView;
  friend class ConstraintSetView;
  struct SetInfo { std::string name; int first, rows; };
  static VecBound Zip(const VectorXd& lo, const VectorXd& up) {
    VecBound b(lo.size());
    for (size_t i = 0; i < lo.size(); ++i) b[i] = Bounds{lo[i], up[i]};
    return b;
  }
  static std::vector<std::string> Names(const std::vector<SetInfo>& v) {
    std::vector<std::string> n;
    for (const auto& s : v) n.push_back(s.name);
    return n;
  }
  const double* xp(int p) const { return x_.data() + static_cast<size_t>(p) * n_vars_; }
  const SetInfo& VarSet(const std::string& name) const {
    for (const auto& s : var_sets_)
      if (s.name == name) return s;
    throw std::runtime_error("variable set " + name + " not found");
  }

  towr_b200_structure* st_ = nullptr;
  towr_b200_batch* batch_ = nullptr;
  int n_, n_vars_ = 0, n_rows_ = 0, nnz_ = 0, device_ = 0;
  std::vector<towr_b200_spec> specs_;
  std::vector<SetInfo> var_sets_, con_sets_;
  std::vector<int> rows_, cols_;
  std::vector<double> x_, g_, v_;
  bool dirty_ = true, cost_dirty_ = true;
  std::vector<double> f_, grad_;
  void EnsureBatch() {
    if (batch_) return;
    Check(towr_b200_batch_create(st_, n_, device_, &batch_));
    for (int p = 0; p < n_; ++p) Check(towr_b200_batch_set_problem(batch_, p, &specs_[p]));
  }
};

inline VectorXd VariableSetView::GetValues() const {
  const double* x = b_->xp(p_) + col0_;
  return VectorXd(x, x + rows_);
}
inline void VariableSetView::SetVariables(const VectorXd& x) {
  if (static_cast<int>(x.size()) != rows_) throw std::runtime_error("SetVariables: wrong size for " + name_);
  std::copy(x.begin(), x.end(), b_->x_.begin() + static_cast<size_t>(p_) * b_->n_vars_ + col0_);
  b_->dirty_ = true;
  b_->cost_dirty_ = true;
}
inline VecBound VariableSetView::GetBounds() const {
  VecBound all = b_->GetBoundsOnOptimizationVariables(p_);
  return VecBound(all.begin() + col0_, all.begin() + col0_ + rows_);
}
inline VectorXd ConstraintSetView::GetValues() const {
  b_->Evaluate();
  const double* g = b_->g_.data() + static_cast<size_t>(p_) * b_->n_rows_ + row0_;
  return VectorXd(g, g + rows_);
}
inline VecBound ConstraintSetView::GetBounds() const {
  VecBound all = b_->GetBoundsOnConstraints(p_);
  return VecBound(all.begin() + row0_, all.begin() + row0_ + rows_);
}
inline void ConstraintSetView::FillJacobianBlock(const std::string& var_set, std::vector<Triplet>& jac_block) const {
  const auto& vs = b_->VarSet(var_set);
  const double* v = b_->GetJacobianOfConstraints(p_);
  jac_block.clear();
  for (int e = 0; e < b_->nnz_; ++e) {
    const int r = b_->rows_[e], c = b_->cols_[e];
    if (r >= row0_ && r < row0_ + rows_ && c >= vs.first && c < vs.first + vs.rows)
      jac_block.push_back(Triplet{r - row0_, c - vs.first, v[e]});
  }
}

}  // namespace towr_b200
