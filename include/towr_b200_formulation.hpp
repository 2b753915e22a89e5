// towr_b200_formulation.hpp — the reference's public formulation types as plain data, and ToSpec().
//
// The reference describes one problem through the six public members of towr::NlpFormulation
// (include/towr/nlp_formulation.h:100-105: initial_base_, final_base_, initial_ee_W_, model_, terrain_, params_) and
// towr::Parameters (include/towr/parameters.h:135-336).  Its headers need Eigen and ifopt, which are not installed
// here, so the structs below carry the SAME member names on std::array / std::vector; an application
// that fills a towr::NlpFormulation fills one of these with the same statements (test/hopper_example.cc:93-150 is
// restated verbatim in tests/cpp/ifopt_surface.cc).  With the real headers present the maintainer's ToSpec copies
// the same members out of towr::NlpFormulation instead.
//
// ToSpec() flattens the formulation into the towr_b200_spec the C ABI reads and refuses, with the reference's own
// exception type where it has one, what the backend does not evaluate (std::runtime_error("unsupported: ...")).
#pragma once
#include <array>
#include <cmath>
#include <limits>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "towr_b200.h"

namespace towr_b200 {
namespace ref {

enum Dx { kPos = 0, kVel, kAcc };                       // variables/state.h:41
enum Dim3D { X = 0, Y, Z };                             // variables/cartesian_dimensions.h
using Vector3d = std::array<double, 3>;

// towr::State / towr::Node (variables/state.h:49-123): at(kPos), at(kVel), p(), v()
struct Node {
  Vector3d pos{{0.0, 0.0, 0.0}}, vel{{0.0, 0.0, 0.0}};
  Vector3d& at(Dx d) {
    if (d == kPos) return pos;
    if (d == kVel) return vel;
    throw std::out_of_range("Node::at: only kPos and kVel are node values");
  }
  const Vector3d& at(Dx d) const { return const_cast<Node*>(this)->at(d); }
  const Vector3d& p() const { return pos; }
  const Vector3d& v() const { return vel; }
};
struct BaseState {                                      // variables/state.h:130-139
  Node lin, ang;
};

// towr::KinematicModel (models/kinematic_model.h) and the constants of towr::SingleRigidBodyDynamics
struct KinematicModel {
  using Ptr = std::shared_ptr<KinematicModel>;
  explicit KinematicModel(int n_ee) : nominal_stance_(n_ee), max_dev_from_nominal_(n_ee), min_dev_from_nominal_(n_ee) {}
  std::vector<Vector3d> GetNominalStanceInBase() const { return nominal_stance_; }
  std::vector<Vector3d> GetMaximumDeviationFromNominal() const { return max_dev_from_nominal_; }
  std::vector<Vector3d> GetMinimumDeviationFromNominal() const { return min_dev_from_nominal_; }
  int GetNumberOfEndeffectors() const { return static_cast<int>(nominal_stance_.size()); }
  std::vector<Vector3d> nominal_stance_, max_dev_from_nominal_, min_dev_from_nominal_;
};
struct DynamicModel {                                   // SingleRigidBodyDynamics(mass, Ixx, Iyy, Izz, Ixy, Ixz, Iyz, ee_count)
  using Ptr = std::shared_ptr<DynamicModel>;
  DynamicModel(double mass, double Ixx, double Iyy, double Izz, double Ixy, double Ixz, double Iyz, int ee_count)
      : m_(mass), inertia_{{Ixx, Iyy, Izz, Ixy, Ixz, Iyz}}, ee_count_(ee_count) {}
  double m() const { return m_; }
  double g() const { return 9.80665; }                  // models/dynamic_model.cc:37
  int GetEECount() const { return ee_count_; }
  double m_;
  std::array<double, 6> inertia_;
  int ee_count_;
};
// towr::RobotModel (models/robot_model.h:63-82); the four robots' constants come from towr_b200_spec_robot
struct RobotModel {
  enum Robot { Monoped, Biped, Hyq, Anymal, ROBOT_COUNT };
  RobotModel() = default;
  explicit RobotModel(Robot robot) {
    towr_b200_spec s;
    if (towr_b200_spec_defaults(&s) != TOWR_B200_OK || towr_b200_spec_robot(&s, static_cast<int>(robot)) != TOWR_B200_OK)
      throw std::runtime_error(towr_b200_last_error());
    kinematic_model_ = std::make_shared<KinematicModel>(s.n_ee);
    for (int e = 0; e < s.n_ee; ++e)
      for (int d = 0; d < 3; ++d) {
        kinematic_model_->nominal_stance_[e][d] = s.nominal_stance[e][d];
        kinematic_model_->max_dev_from_nominal_[e][d] = s.max_dev[e][d];
        kinematic_model_->min_dev_from_nominal_[e][d] = s.min_dev[e][d];
      }
    dynamic_model_ = std::make_shared<DynamicModel>(s.mass, s.inertia[0], s.inertia[1], s.inertia[2], s.inertia[3],
                                                    s.inertia[4], s.inertia[5], s.n_ee);
  }
  KinematicModel::Ptr kinematic_model_;
  DynamicModel::Ptr dynamic_model_;
};

// towr::HeightMap::TerrainID (terrain/height_map.h) + the FiveStepStairs of test/hopper_example.cc:53-91.  The
// reference holds a HeightMap::Ptr; the analytic terrains are selected by id here (height_map.cc:39-52 MakeTerrain).
struct HeightMap {
  using Ptr = std::shared_ptr<HeightMap>;
  enum TerrainID { FlatID, BlockID, StairsID, GapID, SlopeID, ChimneyID, ChimneyLRID, FiveStepStairsID, TERRAIN_COUNT };
  explicit HeightMap(TerrainID id) : id_(id) {}
  static Ptr MakeTerrain(TerrainID type) { return std::make_shared<HeightMap>(type); }
  TerrainID id_;
};

// towr::Parameters (include/towr/parameters.h:135-336, defaults of src/parameters.cc:40-105)
class Parameters {
 public:
  enum ConstraintName { Dynamic, EndeffectorRom, TotalTime, Terrain, TerrainHard, Force, Torque, Swing, BaseRom, BaseAcc,
                        BaseHeight };
  enum CostName { ForcesCostID, EEMotionCostID, EnergyCostID, AngMomCostID };
  using CostWeights = std::vector<std::pair<CostName, double>>;
  using UsedConstraints = std::vector<ConstraintName>;
  using VecTimes = std::vector<double>;
  using VecPos = std::vector<std::vector<double>>;
  using EEID = unsigned int;

  Parameters() {
    towr_b200_spec d;
    if (towr_b200_spec_defaults(&d) != TOWR_B200_OK) throw std::runtime_error(towr_b200_last_error());
    duration_base_polynomial_ = d.duration_base_polynomial;
    force_polynomials_per_stance_phase_ = d.force_polynomials_per_stance_phase;
    torque_polynomials_per_stance_phase_ = d.torque_polynomials_per_stance_phase;
    ee_polynomials_per_swing_phase_ = d.ee_polynomials_per_swing_phase;
    force_limit_in_normal_direction_ = d.force_limit_in_normal_direction;
    torque_tx_min_ = d.torque_tx_min; torque_tx_max_ = d.torque_tx_max;
    torque_ty_min_ = d.torque_ty_min; torque_ty_max_ = d.torque_ty_max;
    torque_k_friction_ = d.torque_k_friction;
    dt_constraint_range_of_motion_ = d.dt_constraint_range_of_motion;
    dt_constraint_dynamic_ = d.dt_constraint_dynamic;
    dt_constraint_base_motion_ = d.dt_constraint_base_motion;
    dt_constraint_force_ = d.dt_constraint_force;
    dt_constraint_torque_ = d.dt_constraint_torque;
    dt_cost_energy_ = d.dt_cost_energy;                    // parameters.cc:69-71
    energy_cost_torque_weight_ = d.energy_cost_torque_weight;
    dt_cost_ang_mom_ = d.dt_cost_ang_mom;
    bound_phase_duration_ = std::make_pair(d.bound_phase_duration[0], d.bound_phase_duration[1]);
    for (int i = 0; i < d.n_constraints; ++i) constraints_.push_back(static_cast<ConstraintName>(d.constraints[i]));
    bounds_final_lin_pos_ = bounds_final_lin_vel_ = bounds_final_ang_pos_ = bounds_final_ang_vel_ = {X, Y, Z};
  }
  virtual ~Parameters() = default;

  std::vector<VecTimes> ee_phase_durations_;
  std::vector<VecPos> ee_stance_position_;
  std::vector<VecPos> ee_stance_rpy_;
  std::vector<bool> ee_in_contact_at_start_;
  UsedConstraints constraints_;
  CostWeights costs_;
  double dt_constraint_dynamic_, dt_constraint_range_of_motion_, dt_constraint_base_motion_;
  double dt_constraint_force_, dt_constraint_torque_;
  double dt_cost_energy_, energy_cost_torque_weight_, dt_cost_ang_mom_;   // parameters.h:245-247
  double duration_base_polynomial_;
  int ee_polynomials_per_swing_phase_, force_polynomials_per_stance_phase_, torque_polynomials_per_stance_phase_;
  double force_limit_in_normal_direction_;
  bool enable_stance_tracking = true;                   // the fork's default (parameters.h:225)
  bool enable_stance_rpy_tracking = true;
  bool enable_swing_ee_base_pos_tracking = false;
  double swing_ee_base_pos_tracking_weight_ = 1e-2;     // parameters.h:234-235
  double dt_cost_swing_ee_base_pos_tracking_ = 0.05;
  bool constrain_base_pitch_ = false;
  double torque_tx_min_, torque_tx_max_, torque_ty_min_, torque_ty_max_, torque_k_friction_;
  std::vector<double> ee_swing_height_min_, ee_swing_height_max_;
  std::vector<int> bounds_final_lin_pos_, bounds_final_lin_vel_, bounds_final_ang_pos_, bounds_final_ang_vel_;
  std::pair<double, double> bound_phase_duration_;
  std::vector<int> rom_swing_relax_dims_;
  enum AngularRepresentation { EulerZYX, RotationVector };
  AngularRepresentation angular_rep_ = EulerZYX;
  // parameters.h:315-327
  struct EELinearConstraintDef {
    struct Term { int ee; int dim; double coeff; };
    enum Target { Motion, Angle };
    Target target = Motion;
    Dx deriv = kPos;
    std::vector<Term> terms;
    double tolerance = 0.0;
    double dt = 0.1;
  };
  std::vector<EELinearConstraintDef> ee_linear_constraints_;
  // count only: a non-empty list is refused by ToSpec (base waypoints are not evaluated)
  int n_base_waypoints_ = 0;

  void OptimizePhaseDurations() { constraints_.push_back(TotalTime); }          // parameters.cc:107-111
  int GetPhaseCount(EEID ee) const { return static_cast<int>(ee_phase_durations_.at(ee).size()); }
  int GetEECount() const { return static_cast<int>(ee_in_contact_at_start_.size()); }
  bool IsOptimizeTimings() const {
    for (ConstraintName c : constraints_)
      if (c == TotalTime) return true;
    return false;
  }
  double GetTotalTime() const {                                                 // parameters.cc:146-159
    if (ee_phase_durations_.empty()) return 0.0;
    double T = 0.0;
    for (double d : ee_phase_durations_.front()) T += d;
    return T;
  }
};

// towr::NlpFormulation's public data members (include/towr/nlp_formulation.h:100-105)
struct NlpFormulation {
  using EEPos = std::vector<Vector3d>;
  BaseState initial_base_;
  BaseState final_base_;
  EEPos initial_ee_W_;
  RobotModel model_;
  HeightMap::Ptr terrain_;
  Parameters params_;
};

}  // namespace ref

// The spec of one problem.  Throws what the reference throws for the same input where it throws (std::out_of_range
// for missing stance positions, nlp_formulation.cc:224), std::runtime_error("unsupported: ...") for members the
// backend does not evaluate, std::length_error beyond the fixed capacities of towr_b200_spec.
inline towr_b200_spec ToSpec(const ref::NlpFormulation& f) {
  const ref::Parameters& P = f.params_;
  towr_b200_spec s;
  if (towr_b200_spec_defaults(&s) != TOWR_B200_OK) throw std::runtime_error(towr_b200_last_error());
  if (!f.model_.kinematic_model_ || !f.model_.dynamic_model_) throw std::runtime_error("NlpFormulation::model_ is not set");
  if (!f.terrain_) throw std::runtime_error("NlpFormulation::terrain_ is not set");
  const int n_ee = P.GetEECount();
  if (n_ee < 1 || n_ee > TOWR_B200_MAX_EE) throw std::length_error("endeffector count outside 1..TOWR_B200_MAX_EE");
  if (f.model_.kinematic_model_->GetNumberOfEndeffectors() != n_ee || static_cast<int>(f.initial_ee_W_.size()) != n_ee ||
      static_cast<int>(P.ee_phase_durations_.size()) != n_ee)
    throw std::runtime_error("model_, initial_ee_W_, ee_phase_durations_ and ee_in_contact_at_start_ disagree on the endeffector count");
  // members without a counterpart in the backend
  if (P.angular_rep_ != ref::Parameters::EulerZYX) throw std::runtime_error("unsupported: RotationVector base orientation");
  if (P.constrain_base_pitch_) throw std::runtime_error("unsupported: constrain_base_pitch_");
  if (!P.ee_swing_height_min_.empty() || !P.ee_swing_height_max_.empty()) throw std::runtime_error("unsupported: per-foot swing height bounds");
  if (!P.rom_swing_relax_dims_.empty()) throw std::runtime_error("unsupported: rom_swing_relax_dims_");
  if (P.n_base_waypoints_ != 0) throw std::runtime_error("unsupported: base waypoints");
  if (static_cast<int>(P.ee_linear_constraints_.size()) > TOWR_B200_MAX_EE_LINEAR) throw std::length_error("too many EELinearConstraint definitions");
  if (P.enable_stance_rpy_tracking)
    for (const auto& v : P.ee_stance_rpy_)
      if (!v.empty()) throw std::runtime_error("unsupported: ee_stance_rpy_ tracking");
  s.n_ee = n_ee;
  s.mass = f.model_.dynamic_model_->m();
  for (int i = 0; i < 6; ++i) s.inertia[i] = f.model_.dynamic_model_->inertia_[i];
  s.terrain_id = static_cast<int>(f.terrain_->id_);
  for (int e = 0; e < n_ee; ++e) {
    for (int d = 0; d < 3; ++d) {
      s.nominal_stance[e][d] = f.model_.kinematic_model_->nominal_stance_[e][d];
      s.max_dev[e][d] = f.model_.kinematic_model_->max_dev_from_nominal_[e][d];
      s.min_dev[e][d] = f.model_.kinematic_model_->min_dev_from_nominal_[e][d];
      s.initial_ee_W[e][d] = f.initial_ee_W_[e][d];
    }
    const auto& ph = P.ee_phase_durations_[e];
    if (ph.empty() || static_cast<int>(ph.size()) > TOWR_B200_MAX_PHASES) throw std::length_error("phase count outside 1..TOWR_B200_MAX_PHASES");
    s.n_phases[e] = static_cast<int>(ph.size());
    for (size_t i = 0; i < ph.size(); ++i) s.phase_durations[e][i] = ph[i];
    s.in_contact_at_start[e] = P.ee_in_contact_at_start_[e] ? 1 : 0;
  }
  if (static_cast<int>(P.constraints_.size()) > TOWR_B200_MAX_CONSTRAINTS) throw std::length_error("too many constraints");
  s.n_constraints = static_cast<int>(P.constraints_.size());
  for (int i = 0; i < s.n_constraints; ++i) s.constraints[i] = static_cast<int>(P.constraints_[i]);
  if (static_cast<int>(P.costs_.size()) > TOWR_B200_MAX_COSTS) throw std::length_error("too many costs");
  s.n_costs = static_cast<int>(P.costs_.size());
  for (int i = 0; i < s.n_costs; ++i) {
    s.cost_ids[i] = static_cast<int>(P.costs_[i].first);
    s.cost_weights[i] = P.costs_[i].second;
  }
  s.dt_constraint_dynamic = P.dt_constraint_dynamic_;
  s.dt_constraint_range_of_motion = P.dt_constraint_range_of_motion_;
  s.dt_constraint_base_motion = P.dt_constraint_base_motion_;
  s.dt_constraint_force = P.dt_constraint_force_;
  s.dt_constraint_torque = P.dt_constraint_torque_;
  s.n_ee_linear = static_cast<int>(P.ee_linear_constraints_.size());      // the backend refuses Angle targets / optimised durations
  for (int i = 0; i < s.n_ee_linear; ++i) {
    const ref::Parameters::EELinearConstraintDef& d = P.ee_linear_constraints_[i];
    if (static_cast<int>(d.terms.size()) > TOWR_B200_MAX_EE_LINEAR_TERMS) throw std::length_error("too many EELinearConstraint terms");
    s.ee_linear_target[i] = d.target == ref::Parameters::EELinearConstraintDef::Angle ? 1 : 0;
    s.ee_linear_deriv[i] = static_cast<int>(d.deriv);
    s.ee_linear_n_terms[i] = static_cast<int>(d.terms.size());
    for (size_t j = 0; j < d.terms.size(); ++j) {
      s.ee_linear_term_ee[i][j] = d.terms[j].ee;
      s.ee_linear_term_dim[i][j] = d.terms[j].dim;
      s.ee_linear_term_coeff[i][j] = d.terms[j].coeff;
    }
    s.ee_linear_tolerance[i] = d.tolerance;
    s.ee_linear_dt[i] = d.dt;
  }
  s.dt_cost_energy = P.dt_cost_energy_;
  s.energy_cost_torque_weight = P.energy_cost_torque_weight_;
  s.dt_cost_ang_mom = P.dt_cost_ang_mom_;
  s.enable_swing_ee_base_pos_tracking = P.enable_swing_ee_base_pos_tracking ? 1 : 0;
  s.swing_ee_base_pos_tracking_weight = P.swing_ee_base_pos_tracking_weight_;
  s.dt_cost_swing_ee_base_pos_tracking = P.dt_cost_swing_ee_base_pos_tracking_;
  s.duration_base_polynomial = P.duration_base_polynomial_;
  s.ee_polynomials_per_swing_phase = P.ee_polynomials_per_swing_phase_;
  s.force_polynomials_per_stance_phase = P.force_polynomials_per_stance_phase_;
  s.torque_polynomials_per_stance_phase = P.torque_polynomials_per_stance_phase_;
  s.force_limit_in_normal_direction = P.force_limit_in_normal_direction_;
  s.torque_tx_min = P.torque_tx_min_; s.torque_tx_max = P.torque_tx_max_;
  s.torque_ty_min = P.torque_ty_min_; s.torque_ty_max = P.torque_ty_max_;
  s.torque_k_friction = P.torque_k_friction_;
  s.bound_phase_duration[0] = P.bound_phase_duration_.first;
  s.bound_phase_duration[1] = P.bound_phase_duration_.second;
  auto flags = [](const std::vector<int>& dims, int (&out)[3]) {
    for (int d = 0; d < 3; ++d) out[d] = 0;
    for (int d : dims) {
      if (d < 0 || d > 2) throw std::out_of_range("bounds_final_*: dimension outside X, Y, Z");
      out[d] = 1;
    }
  };
  flags(P.bounds_final_lin_pos_, s.bounds_final_lin_pos);
  flags(P.bounds_final_lin_vel_, s.bounds_final_lin_vel);
  flags(P.bounds_final_ang_pos_, s.bounds_final_ang_pos);
  flags(P.bounds_final_ang_vel_, s.bounds_final_ang_vel);
  auto state = [](const ref::Node& n, double (&out)[6]) {
    for (int d = 0; d < 3; ++d) { out[d] = n.pos[d]; out[3 + d] = n.vel[d]; }
  };
  state(f.initial_base_.lin, s.initial_base_lin);
  state(f.initial_base_.ang, s.initial_base_ang);
  state(f.final_base_.lin, s.final_base_lin);
  state(f.final_base_.ang, s.final_base_ang);
  s.enable_stance_tracking = P.enable_stance_tracking ? 1 : 0;
  if (P.enable_stance_tracking) {
    if (static_cast<int>(P.ee_stance_position_.size()) < n_ee)
      throw std::out_of_range("ee_stance_position_.at(ee)");                     // vector::at in nlp_formulation.cc:224
    for (int e = 0; e < n_ee; ++e) {
      const auto& v = P.ee_stance_position_[e];
      if (static_cast<int>(v.size()) > TOWR_B200_MAX_STANCES) throw std::length_error("more than TOWR_B200_MAX_STANCES stance positions");
      s.n_stance_positions[e] = static_cast<int>(v.size());
      for (size_t k = 0; k < v.size(); ++k) {
        if (v[k].size() < 2) throw std::out_of_range("ee_stance_position_ entries need x and y");
        s.ee_stance_position[e][k][0] = v[k][0];
        s.ee_stance_position[e][k][1] = v[k][1];
      }
    }
  }
  return s;
}

}  // namespace towr_b200
