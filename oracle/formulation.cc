// ORACLE — TEST INFRASTRUCTURE ONLY (see towr_oracle.h header).
// Parameters, NlpFormulation (variable sets, constraint list, bounds, initial guess)
// and the ifopt::Problem assembly restated from its call sites.
#include <algorithm>
#include <cmath>
#include <limits>
#include <numeric>
#include <stdexcept>

#include "constraints.h"

namespace orc {

// ================================================================ parameters.cc
// :40-105
Parameters::Parameters() {
  duration_base_polynomial_ = 0.1;
  force_polynomials_per_stance_phase_ = 3;
  torque_polynomials_per_stance_phase_ = 3;
  ee_polynomials_per_swing_phase_ = 2;
  force_limit_in_normal_direction_ = 1000;
  torque_tx_min_ = -100.0; torque_tx_max_ = 100.0;
  torque_ty_min_ = -100.0; torque_ty_max_ = 100.0;
  torque_k_friction_ = 2.0 / 3.0;
  dt_constraint_range_of_motion_ = 0.08;
  dt_constraint_dynamic_ = 0.1;
  dt_constraint_base_motion_ = duration_base_polynomial_ / 4.;
  dt_constraint_force_ = 0.02;
  dt_constraint_torque_ = 0.02;
  dt_cost_energy_ = 0.02;               // parameters.cc:69-71
  energy_cost_torque_weight_ = 1.0;
  dt_cost_ang_mom_ = 0.02;
  bound_phase_duration_ = std::make_pair(0.2, 1.0);
  constraints_ = {Terrain, Dynamic, BaseAcc, EndeffectorRom, Force, Swing, BaseHeight};
  bounds_final_lin_pos_ = {X, Y, Z};
  bounds_final_lin_vel_ = {X, Y, Z};
  bounds_final_ang_pos_ = {X, Y, Z};
  bounds_final_ang_vel_ = {X, Y, Z};
  base_rom_ax = NoBound();
  base_rom_ay = NoBound();
  base_rom_lz = NoBound();
}
// :114-130
Vec Parameters::GetBasePolyDurations() const {
  Vec base_spline_timings;
  double dt = duration_base_polynomial_;
  double t_left = GetTotalTime();
  double eps = 1e-10;
  while (t_left > eps) {
    double duration = t_left > dt ? dt : t_left;
    base_spline_timings.push_back(duration);
    t_left -= dt;
  }
  return base_spline_timings;
}
// :144-158
double Parameters::GetTotalTime() const {
  std::vector<double> T_feet;
  for (const auto& v : ee_phase_durations_) T_feet.push_back(std::accumulate(v.begin(), v.end(), 0.0));
  return T_feet.empty() ? 0.0 : T_feet.front();
}
// :160-167
bool Parameters::IsOptimizeTimings() const {
  return std::find(constraints_.begin(), constraints_.end(), TotalTime) != constraints_.end();
}

// ========================================================== nlp_formulation.cc
// :76-119 GetVariableSets with :121-363 Make*Variables inlined; SplineHolder per helpers/spline_holder.cc:35-69
std::vector<std::shared_ptr<VariableSet>> NlpFormulation::GetVariableSets(SplineHolder& spline_holder) {
  std::vector<std::shared_ptr<VariableSet>> vars;
  const double T = params_.GetTotalTime();
  const int n_ee = params_.GetEECount();
  using PB = NodesVariablesPhaseBased;

  // ---- MakeBaseVariables :121-186
  int n_nodes = static_cast<int>(params_.GetBasePolyDurations().size()) + 1;
  auto spline_lin = std::make_shared<NodesVariablesAll>(n_nodes, "base-lin");
  {
    double x = final_base_.lin.p.x();
    double y = final_base_.lin.p.y();
    double z = terrain_->GetHeight(x, y) - model_.kinematic_model_->nominal_stance_.front().z();
    V3 final_pos(x, y, z);
    spline_lin->SetByLinearInterpolation(initial_base_.lin.p, final_pos, T);
    spline_lin->AddStartBound(kPos, {X, Y, Z}, initial_base_.lin.p);
    spline_lin->AddStartBound(kVel, {X, Y, Z}, initial_base_.lin.v);
    // AddStartBound/AddFinalBound(kAcc, ...) in the reference match no node value: no-ops
    spline_lin->AddFinalBound(kPos, params_.bounds_final_lin_pos_, final_base_.lin.p);
    spline_lin->AddFinalBound(kVel, params_.bounds_final_lin_vel_, final_base_.lin.v);
  }
  auto spline_ang = std::make_shared<NodesVariablesAll>(n_nodes, "base-ang");
  spline_ang->SetByLinearInterpolation(initial_base_.ang.p, final_base_.ang.p, T);
  spline_ang->AddStartBound(kPos, {X, Y, Z}, initial_base_.ang.p);
  spline_ang->AddStartBound(kVel, {X, Y, Z}, initial_base_.ang.v);
  spline_ang->AddFinalBound(kPos, params_.bounds_final_ang_pos_, final_base_.ang.p);
  spline_ang->AddFinalBound(kVel, params_.bounds_final_ang_vel_, final_base_.ang.v);
  vars.push_back(spline_lin);
  vars.push_back(spline_ang);

  // ---- MakeEndeffectorVariables :188-242
  std::vector<std::shared_ptr<PB>> ee_motion, ee_ang, ee_force, ee_torque;
  for (int ee = 0; ee < n_ee; ee++) {
    auto nodes = std::make_shared<PB>(PB::Motion, params_.GetPhaseCount(ee), params_.ee_in_contact_at_start_.at(ee),
                                      "ee-motion_" + std::to_string(ee), params_.ee_polynomials_per_swing_phase_);
    double yaw = final_base_.ang.p.z();
    V3 euler(0.0, 0.0, yaw);
    M3 w_R_b = EulerConverter::GetRotationMatrixBaseToWorldDense(euler);
    V3 final_ee_pos_W = final_base_.lin.p + w_R_b * model_.kinematic_model_->nominal_stance_.at(ee);
    double x = final_ee_pos_W.x();
    double y = final_ee_pos_W.y();
    double z = terrain_->GetHeight(x, y);
    nodes->SetByLinearInterpolationRelativeToBase(initial_ee_W_.at(ee), V3(x, y, z),
                                                  initial_base_.lin.p, final_base_.lin.p,
                                                  initial_base_.ang.p, final_base_.ang.p, T);
    if (params_.enable_stance_tracking) {
      int phase_count = params_.GetPhaseCount(ee);
      int stance_id = 0;
      for (int phase = 0; phase < phase_count; phase++) {
        int node_id = nodes->GetNodeIDAtStartOfPhase(phase);
        if (!nodes->IsConstantNode(node_id)) continue;
        const auto& sp = params_.ee_stance_position_.at(ee).at(stance_id);   // throws std::out_of_range like the reference
        nodes->AddBounds(node_id, kPos, {X, Y}, V3(sp.at(0), sp.at(1), 0.0));
        stance_id++;
        phase++;
      }
    } else {
      nodes->AddStartBound(kPos, {X, Y, Z}, initial_ee_W_.at(ee));
      nodes->AddFinalBound(kPos, {X, Y, Z}, V3(x, y, z));
    }
    ee_motion.push_back(nodes);
  }
  // ---- MakeEndeffectorAngVariables :244-287 (no ee_stance_rpy_ given: start/final bounds branch)
  for (int ee = 0; ee < n_ee; ee++) {
    auto nodes = std::make_shared<PB>(PB::Ang, params_.GetPhaseCount(ee), params_.ee_in_contact_at_start_.at(ee),
                                      "ee-ang_" + std::to_string(ee), params_.ee_polynomials_per_swing_phase_);
    nodes->SetByLinearInterpolation(initial_base_.ang.p, final_base_.ang.p, T);
    nodes->AddStartBound(kPos, {0, 1, 2}, initial_base_.ang.p);
    nodes->AddFinalBound(kPos, {0, 1, 2}, final_base_.ang.p);
    ee_ang.push_back(nodes);
  }
  // ---- MakeForceVariables :289-318
  for (int ee = 0; ee < n_ee; ee++) {
    auto nodes = std::make_shared<PB>(PB::Force, params_.GetPhaseCount(ee), params_.ee_in_contact_at_start_.at(ee),
                                      "ee-force_" + std::to_string(ee), params_.force_polynomials_per_stance_phase_);
    double m = model_.dynamic_model_->m();
    double g = model_.dynamic_model_->g();
    V3 f_stance(0.0, 0.0, m * g / n_ee);
    nodes->SetByLinearInterpolation(f_stance, f_stance, T);
    nodes->AddStartBound(kPos, {X, Y, Z}, V3(0.0, 0.0, m * g / n_ee));
    nodes->AddFinalBound(kPos, {X, Y, Z}, V3(0.0, 0.0, m * g / n_ee));
    nodes->AddStartBound(kVel, {X, Y, Z}, V3(0.0, 0.0, 0.0));
    nodes->AddFinalBound(kVel, {X, Y, Z}, V3(0.0, 0.0, 0.0));
    ee_force.push_back(nodes);
  }
  // ---- MakeTorqueVariables :320-346
  for (int ee = 0; ee < n_ee; ee++) {
    auto nodes = std::make_shared<PB>(PB::Torque, params_.GetPhaseCount(ee), params_.ee_in_contact_at_start_.at(ee),
                                      "ee-torque_" + std::to_string(ee), params_.torque_polynomials_per_stance_phase_);
    V3 tau_stance(0.0, 0.0, 0.0);
    nodes->SetByLinearInterpolation(tau_stance, tau_stance, T);
    nodes->AddStartBound(kPos, {X, Y, Z}, V3());
    nodes->AddFinalBound(kPos, {X, Y, Z}, V3());
    nodes->AddStartBound(kVel, {X, Y, Z}, V3());
    nodes->AddFinalBound(kVel, {X, Y, Z}, V3());
    ee_torque.push_back(nodes);
  }
  for (auto& v : ee_motion) vars.push_back(v);
  for (auto& v : ee_ang) vars.push_back(v);
  for (auto& v : ee_force) vars.push_back(v);
  for (auto& v : ee_torque) vars.push_back(v);

  // ---- MakeContactScheduleVariables :348-363 ; only part of the NLP when timings are optimised (:99-101)
  std::vector<std::shared_ptr<PhaseDurations>> contact_schedule;
  for (int ee = 0; ee < n_ee; ee++)
    contact_schedule.push_back(std::make_shared<PhaseDurations>(ee, params_.ee_phase_durations_.at(ee),
                                                                params_.ee_in_contact_at_start_.at(ee),
                                                                params_.bound_phase_duration_.first,
                                                                params_.bound_phase_duration_.second));
  bool durations_change = params_.IsOptimizeTimings();
  if (durations_change)
    for (auto& v : contact_schedule) vars.push_back(v);

  // ---- SplineHolder ctor, helpers/spline_holder.cc:35-69
  SplineHolder sh;
  sh.base_linear_ = std::make_shared<NodeSpline>(spline_lin.get(), params_.GetBasePolyDurations());
  sh.base_angular_ = std::make_shared<NodeSpline>(spline_ang.get(), params_.GetBasePolyDurations());
  sh.phase_durations_ = contact_schedule;
  for (int ee = 0; ee < n_ee; ++ee) {
    auto make = [&](const std::shared_ptr<PB>& nodes) -> std::shared_ptr<NodeSpline> {
      if (durations_change) return std::make_shared<PhaseSpline>(nodes, contact_schedule.at(ee).get());
      return std::make_shared<NodeSpline>(nodes.get(), nodes->ConvertPhaseToPolyDurations(contact_schedule.at(ee)->GetPhaseDurations()));
    };
    sh.ee_motion_.push_back(make(ee_motion.at(ee)));
    sh.ee_ang_.push_back(make(ee_ang.at(ee)));
    sh.ee_force_.push_back(make(ee_force.at(ee)));
    sh.ee_torque_.push_back(make(ee_torque.at(ee)));
  }
  sh.angular_converter_ = std::make_shared<EulerConverter>(sh.base_angular_);   // :113-116 (EulerZYX)
  spline_holder = sh;
  spline_holder_for_costs_ = sh;
  keep_alive_ = vars;
  keep_alive_.insert(keep_alive_.end(), contact_schedule.begin(), contact_schedule.end());
  return vars;
}

// :365-378
std::vector<std::shared_ptr<ConstraintSet>> NlpFormulation::GetConstraints(const SplineHolder& s) const {
  std::vector<std::shared_ptr<ConstraintSet>> constraints;
  for (auto name : params_.constraints_)
    for (auto& c : GetConstraint(name, s)) constraints.push_back(c);
  for (const auto& def : params_.ee_linear_constraints_) {            // :373-375, MakeEELinearConstraint :682-704
    std::vector<EELinearConstraint::Term> terms;
    for (const auto& t : def.terms) terms.push_back({t.ee, t.dim, t.coeff});
    const bool motion = def.target == Parameters::EELinearConstraintDef::Motion;
    const auto& splines = motion ? s.ee_motion_ : s.ee_ang_;
    std::vector<std::string> var_names;
    for (int ee = 0; ee < params_.GetEECount(); ++ee) var_names.push_back((motion ? "ee-motion_" : "ee-ang_") + std::to_string(ee));
    constraints.push_back(std::make_shared<EELinearConstraint>(terms, splines, var_names, def.deriv, def.tolerance,
                                                               params_.GetTotalTime(), def.dt));
  }
  return constraints;
}

// :380-602
std::vector<std::shared_ptr<ConstraintSet>> NlpFormulation::GetConstraint(Parameters::ConstraintName name, const SplineHolder& s) const {
  std::vector<std::shared_ptr<ConstraintSet>> c;
  const double T = params_.GetTotalTime();
  const int n_ee = params_.GetEECount();
  switch (name) {
    case Parameters::Dynamic:
      c.push_back(std::make_shared<DynamicConstraint>(model_.dynamic_model_, T, params_.dt_constraint_dynamic_, s));
      break;
    case Parameters::EndeffectorRom:
      for (int ee = 0; ee < n_ee; ee++)
        c.push_back(std::make_shared<RangeOfMotionConstraint>(model_.kinematic_model_, T, params_.dt_constraint_range_of_motion_, ee, s));
      break;
    case Parameters::BaseRom:
      c.push_back(std::make_shared<BaseMotionConstraint>(T, params_.dt_constraint_base_motion_, params_.base_rom_ax, params_.base_rom_ay, params_.base_rom_lz, s));
      break;
    case Parameters::TotalTime:
      for (int ee = 0; ee < n_ee; ee++) c.push_back(std::make_shared<TotalDurationConstraint>(T, ee));
      break;
    case Parameters::Terrain:
      for (int ee = 0; ee < n_ee; ee++) {
        double min_height = 0.02;
        double max_height = std::numeric_limits<double>::infinity();
        if (ee < static_cast<int>(params_.ee_swing_height_min_.size())) min_height = params_.ee_swing_height_min_.at(ee);
        if (ee < static_cast<int>(params_.ee_swing_height_max_.size())) max_height = params_.ee_swing_height_max_.at(ee);
        if (min_height < 0.0) throw std::runtime_error("Swing height minimum must be >= 0.0");
        if (max_height <= min_height) throw std::runtime_error("Swing height maximum must be > minimum");
        c.push_back(std::make_shared<TerrainConstraint>(terrain_, "ee-motion_" + std::to_string(ee), min_height, max_height));
      }
      break;
    case Parameters::TerrainHard:
      for (int ee = 0; ee < n_ee; ee++)
        c.push_back(std::make_shared<TerrainConstraintHard>(terrain_, T, params_.dt_constraint_range_of_motion_, ee, s));
      break;
    case Parameters::Force:
      for (int ee = 0; ee < n_ee; ee++) {
        if (params_.dt_constraint_force_ > 0.0)
          c.push_back(std::make_shared<ForceConstraintDiscretized>(terrain_, T, params_.dt_constraint_force_,
                                                                   params_.force_limit_in_normal_direction_, ee, spline_holder_for_costs_));
        else
          c.push_back(std::make_shared<ForceConstraint>(terrain_, params_.force_limit_in_normal_direction_, ee));
      }
      break;
    case Parameters::Torque:
      for (int ee = 0; ee < n_ee; ee++) {
        if (params_.dt_constraint_torque_ > 0.0)
          c.push_back(std::make_shared<TorqueConstraintDiscretized>(terrain_, T, params_.dt_constraint_torque_,
                                                                    params_.torque_tx_min_, params_.torque_tx_max_,
                                                                    params_.torque_ty_min_, params_.torque_ty_max_,
                                                                    params_.torque_k_friction_, ee, spline_holder_for_costs_));
        else
          c.push_back(std::make_shared<TorqueConstraint>(terrain_, params_.torque_tx_min_, params_.torque_tx_max_,
                                                         params_.torque_ty_min_, params_.torque_ty_max_,
                                                         params_.torque_k_friction_, ee));
      }
      break;
    case Parameters::Swing:
      for (int ee = 0; ee < n_ee; ee++) c.push_back(std::make_shared<SwingConstraint>("ee-motion_" + std::to_string(ee)));
      break;
    case Parameters::BaseAcc:
      c.push_back(std::make_shared<SplineAccConstraint>(s.base_linear_, "base-lin"));
      c.push_back(std::make_shared<SplineAccConstraint>(s.base_angular_, "base-ang"));
      break;
    case Parameters::BaseHeight:
      c.push_back(std::make_shared<BaseHeightConstraint>(terrain_, 0.4, "base-lin"));   // :597
      break;
    default: throw std::runtime_error("constraint not defined!");
  }
  return c;
}

// ============================================================== ifopt::Problem
std::shared_ptr<VariableSet> Variables::GetComponent(const std::string& name) const {
  for (auto& s : sets_)
    if (s->GetName() == name) return s;
  throw std::runtime_error("variable set not found: " + name);
}

Vec Problem::GetVariableValues() const {
  Vec x;
  for (auto& s : variables_->GetComponents()) { Vec v = s->GetValues(); x.insert(x.end(), v.begin(), v.end()); }
  return x;
}
void Problem::SetVariables(const double* x) {
  int off = 0;
  for (auto& s : variables_->GetComponents()) {
    int n = s->GetRows();
    s->SetVariables(Vec(x + off, x + off + n));
    off += n;
  }
}
Vec Problem::EvaluateConstraints(const double* x) {
  SetVariables(x);
  Vec g;
  for (auto& c : constraints_) { Vec v = c->GetValues(); g.insert(g.end(), v.begin(), v.end()); }
  return g;
}
// ConstraintSet::GetJacobian: one FillJacobianBlock per variable set on a fresh (rows x n) block, triplets shifted
// by the column offset; Composite::GetJacobian shifts by the row offset; setFromTriplets gives row-major sorted storage.
void Problem::EvalJacobian(const double* x, std::vector<int>* rows, std::vector<int>* cols, std::vector<double>* vals) {
  SetVariables(x);
  if (rows) rows->clear();
  if (cols) cols->clear();
  if (vals) vals->clear();
  int row0 = 0;
  for (auto& c : constraints_) {
    int nr = c->GetRows();
    std::vector<std::vector<Entry>> merged(nr);
    int col0 = 0;
    for (auto& vs : variables_->GetComponents()) {
      int n = vs->GetRows();
      SpMat jac(nr, n);
      c->FillJacobianBlock(vs->GetName(), jac);
      for (int r = 0; r < nr; ++r)
        for (auto& en : jac.r[r]) merged[r].emplace_back(col0 + en.first, en.second);   // blocks are disjoint & ascending in col0
      col0 += n;
    }
    for (int r = 0; r < nr; ++r)
      for (auto& en : merged[r]) {
        if (rows) rows->push_back(row0 + r);
        if (cols) cols->push_back(en.first);
        if (vals) vals->push_back(en.second);
      }
    row0 += nr;
  }
}
VecBound Problem::GetBoundsOnOptimizationVariables() const {
  VecBound b;
  for (auto& s : variables_->GetComponents()) { VecBound v = s->GetBounds(); b.insert(b.end(), v.begin(), v.end()); }
  return b;
}
VecBound Problem::GetBoundsOnConstraints() const {
  VecBound b;
  for (auto& c : constraints_) { VecBound v = c->GetBounds(); b.insert(b.end(), v.begin(), v.end()); }
  return b;
}

}  // namespace orc
