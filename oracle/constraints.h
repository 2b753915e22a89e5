// ORACLE — TEST INFRASTRUCTURE ONLY (see towr_oracle.h header).
// Constraint sets of the hot path plus the opt-in ones (§8f rank 1).
#pragma once
#include "towr_oracle.h"

namespace orc {

// constraints/dynamic_constraint.cc:38-148
class DynamicConstraint : public TimeDiscretizationConstraint {
 public:
  DynamicConstraint(const std::shared_ptr<SingleRigidBodyDynamics>& m, double T, double dt, const SplineHolder& s);
 private:
  int GetRow(int k, int dim) const { return k6D * k + dim; }
  void UpdateConstraintAtInstance(double t, int k, Vec& g) const override;
  void UpdateBoundsAtInstance(double t, int k, VecBound& b) const override;
  void UpdateJacobianAtInstance(double t, int k, const std::string& var_set, SpMat& jac) const override;
  void UpdateModel(double t) const;
  std::shared_ptr<NodeSpline> base_linear_;
  std::shared_ptr<EulerConverter> base_angular_;
  std::vector<std::shared_ptr<NodeSpline>> ee_forces_, ee_torques_, ee_motion_;
  mutable std::shared_ptr<SingleRigidBodyDynamics> model_;
};

// constraints/range_of_motion_constraint.cc:37-131
class RangeOfMotionConstraint : public TimeDiscretizationConstraint {
 public:
  RangeOfMotionConstraint(const std::shared_ptr<KinematicModel>& model, double T, double dt, int ee, const SplineHolder& s);
 private:
  int GetRow(int node, int dim) const { return node * k3D + dim; }
  void UpdateConstraintAtInstance(double t, int k, Vec& g) const override;
  void UpdateBoundsAtInstance(double t, int k, VecBound& b) const override;
  void UpdateJacobianAtInstance(double t, int k, const std::string& var_set, SpMat& jac) const override;
  std::shared_ptr<NodeSpline> base_linear_, ee_motion_;
  std::shared_ptr<EulerConverter> base_angular_;
  V3 max_deviation_from_nominal_, min_deviation_from_nominal_, nominal_ee_pos_B_;
  int ee_;
};

// constraints/force_constraint_discretized.cc:38-221
class ForceConstraintDiscretized : public TimeDiscretizationConstraint {
 public:
  ForceConstraintDiscretized(const std::shared_ptr<HeightMap>& terrain, double T, double dt, double force_limit, int ee, const SplineHolder& s);
 private:
  int GetRow(int k, int local_row) const { return k * n_constraints_per_instance_ + local_row; }
  void UpdateConstraintAtInstance(double t, int k, Vec& g) const override;
  void UpdateBoundsAtInstance(double t, int k, VecBound& b) const override;
  void UpdateJacobianAtInstance(double t, int k, const std::string& var_set, SpMat& jac) const override;
  std::shared_ptr<HeightMap> terrain_;
  std::shared_ptr<NodeSpline> ee_force_, ee_motion_;
  double fn_max_, mu_;
  int ee_, n_constraints_per_instance_;
};

// constraints/torque_constraint_discretized.cc:68-235
class TorqueConstraintDiscretized : public TimeDiscretizationConstraint {
 public:
  TorqueConstraintDiscretized(const std::shared_ptr<HeightMap>& terrain, double T, double dt,
                              double tx_min, double tx_max, double ty_min, double ty_max, double k_friction,
                              int ee, const SplineHolder& s);
 private:
  int GetRow(int k, int local_row) const { return k * 4 + local_row; }
  void UpdateConstraintAtInstance(double t, int k, Vec& g) const override;
  void UpdateBoundsAtInstance(double t, int k, VecBound& b) const override;
  void UpdateJacobianAtInstance(double t, int k, const std::string& var_set, SpMat& jac) const override;
  std::shared_ptr<HeightMap> terrain_;
  std::shared_ptr<NodeSpline> ee_torque_, ee_force_, ee_motion_;
  double tx_min_, tx_max_, ty_min_, ty_max_, k_friction_, mu_;
  int ee_;
};

// constraints/terrain_constraint_hard.cc:37-132
class TerrainConstraintHard : public TimeDiscretizationConstraint {
 public:
  TerrainConstraintHard(const std::shared_ptr<HeightMap>& terrain, double T, double dt, int ee, const SplineHolder& s);
 private:
  void UpdateConstraintAtInstance(double t, int k, Vec& g) const override;
  void UpdateBoundsAtInstance(double t, int k, VecBound& b) const override;
  void UpdateJacobianAtInstance(double t, int k, const std::string& var_set, SpMat& jac) const override;
  std::shared_ptr<HeightMap> terrain_;
  std::shared_ptr<NodeSpline> ee_motion_;
  int ee_;
  double k_coeff_;
};

// constraints/base_motion_constraint.cc:38-91
class BaseMotionConstraint : public TimeDiscretizationConstraint {
 public:
  BaseMotionConstraint(double T, double dt, Bounds ax, Bounds ay, Bounds lz, const SplineHolder& s);
 private:
  int GetRow(int node, int dim) const { return node * static_cast<int>(node_bounds_.size()) + dim; }
  void UpdateConstraintAtInstance(double t, int k, Vec& g) const override;
  void UpdateBoundsAtInstance(double t, int k, VecBound& b) const override;
  void UpdateJacobianAtInstance(double t, int k, const std::string& var_set, SpMat& jac) const override;
  std::shared_ptr<NodeSpline> base_linear_, base_angular_;
  VecBound node_bounds_;
};

// constraints/ee_linear_constraint.cc:5-48
class EELinearConstraint : public TimeDiscretizationConstraint {
 public:
  struct Term { int ee; int dim; double coeff; };
  EELinearConstraint(const std::vector<Term>& terms, const std::vector<std::shared_ptr<NodeSpline>>& splines,
                     const std::vector<std::string>& var_names, Dx deriv, double tolerance, double T, double dt);
 private:
  void UpdateConstraintAtInstance(double t, int k, Vec& g) const override;
  void UpdateBoundsAtInstance(double t, int k, VecBound& b) const override;
  void UpdateJacobianAtInstance(double t, int k, const std::string& var_set, SpMat& jac) const override;
  std::vector<Term> terms_;
  std::vector<std::shared_ptr<NodeSpline>> splines_;
  std::vector<std::string> var_names_;
  Dx deriv_;
  double tolerance_;
};

// constraints/terrain_constraint.cc:36-111
class TerrainConstraint : public ConstraintSet {
 public:
  TerrainConstraint(const std::shared_ptr<HeightMap>& terrain, const std::string& ee_motion, double min_h, double max_h);
  void InitVariableDependedQuantities(const std::shared_ptr<Variables>& x) override;
  Vec GetValues() const override;
  VecBound GetBounds() const override;
  void FillJacobianBlock(const std::string& var_set, SpMat& jac) const override;
 private:
  std::shared_ptr<NodesVariablesPhaseBased> ee_motion_;
  std::shared_ptr<HeightMap> terrain_;
  std::string ee_motion_id_;
  std::vector<int> node_ids_;
  double min_swing_height_, max_swing_height_;
};

// constraints/swing_constraint.cc:35-108 ; t_swing_avg_ = 0.3 (swing_constraint.h:68)
class SwingConstraint : public ConstraintSet {
 public:
  explicit SwingConstraint(const std::string& ee_motion);
  void InitVariableDependedQuantities(const std::shared_ptr<Variables>& x) override;
  Vec GetValues() const override;
  VecBound GetBounds() const override;
  void FillJacobianBlock(const std::string& var_set, SpMat& jac) const override;
 private:
  std::shared_ptr<NodesVariablesPhaseBased> ee_motion_;
  double t_swing_avg_ = .3;
  std::string ee_motion_id_;
  std::vector<int> pure_swing_node_ids_;
};

// constraints/spline_acc_constraint.cc:34-86
class SplineAccConstraint : public ConstraintSet {
 public:
  SplineAccConstraint(const std::shared_ptr<NodeSpline>& spline, const std::string& node_variable_name);
  Vec GetValues() const override;
  VecBound GetBounds() const override;
  void FillJacobianBlock(const std::string& var_set, SpMat& jac) const override;
 private:
  std::shared_ptr<NodeSpline> spline_;
  std::string node_variables_id_;
  int n_junctions_, n_dim_;
  Vec T_;
};

// constraints/base_height_constraint.cc:35-110
class BaseHeightConstraint : public ConstraintSet {
 public:
  BaseHeightConstraint(const std::shared_ptr<HeightMap>& terrain, double safety_distance, const std::string& base_lin_id);
  void InitVariableDependedQuantities(const std::shared_ptr<Variables>& x) override;
  Vec GetValues() const override;
  VecBound GetBounds() const override;
  void FillJacobianBlock(const std::string& var_set, SpMat& jac) const override;
 private:
  std::shared_ptr<NodesVariablesAll> base_linear_;
  std::shared_ptr<HeightMap> terrain_;
  std::string base_lin_id_;
  std::vector<int> node_ids_;
  double safety_distance_;
};

// constraints/total_duration_constraint.cc:36-72
class TotalDurationConstraint : public ConstraintSet {
 public:
  TotalDurationConstraint(double T_total, int ee);
  void InitVariableDependedQuantities(const std::shared_ptr<Variables>& x) override;
  Vec GetValues() const override;
  VecBound GetBounds() const override;
  void FillJacobianBlock(const std::string& var_set, SpMat& jac) const override;
 private:
  std::shared_ptr<PhaseDurations> phase_durations_;
  double T_total_;
  int ee_;
};

// constraints/force_constraint.cc:37-171 (node-based variant, used when dt_constraint_force_ <= 0)
class ForceConstraint : public ConstraintSet {
 public:
  ForceConstraint(const std::shared_ptr<HeightMap>& terrain, double force_limit, int ee);
  void InitVariableDependedQuantities(const std::shared_ptr<Variables>& x) override;
  Vec GetValues() const override;
  VecBound GetBounds() const override;
  void FillJacobianBlock(const std::string& var_set, SpMat& jac) const override;
 private:
  std::shared_ptr<NodesVariablesPhaseBased> ee_force_, ee_motion_;
  std::shared_ptr<HeightMap> terrain_;
  double fn_max_, mu_;
  int n_constraints_per_node_, ee_;
  std::vector<int> pure_stance_force_node_ids_;
};

// constraints/torque_constraint.cc:34-159 (node-based; used when dt_constraint_torque_ <= 0, nlp_formulation.cc:543-567)
class TorqueConstraint : public ConstraintSet {
 public:
  TorqueConstraint(const std::shared_ptr<HeightMap>& terrain, double tx_min, double tx_max, double ty_min, double ty_max,
                   double k_friction, int ee);
  void InitVariableDependedQuantities(const std::shared_ptr<Variables>& x) override;
  Vec GetValues() const override;
  VecBound GetBounds() const override;
  void FillJacobianBlock(const std::string& var_set, SpMat& jac) const override;
 private:
  std::shared_ptr<NodesVariablesPhaseBased> ee_torque_, ee_motion_, ee_force_;
  std::shared_ptr<HeightMap> terrain_;
  double tx_min_, tx_max_, ty_min_, ty_max_, k_friction_, mu_;
  int n_constraints_per_node_, ee_;
  std::vector<int> pure_stance_torque_node_ids_;
};


// costs/energy_cost.cc, costs/angular_momentum_cost.cc (costs.cc)
std::vector<double> CostSampleTimes(const SplineHolder& s, double dt);
double EnergyCostValue(const SplineHolder& s, double weight, double torque_weight, double dt);
void EnergyCostFillBlock(const SplineHolder& s, double weight, double torque_weight, double dt, const std::string& var_set,
                         double* block);
double EEBasePosCostValue(const SplineHolder& s, int ee, const V3& p_ref_B, double weight, double dt);
void EEBasePosCostFillBlock(const SplineHolder& s, int ee, const V3& p_ref_B, double weight, double dt,
                            const std::string& var_set, double* block);
double AngMomCostValue(const SplineHolder& s, const SingleRigidBodyDynamics& model, double weight, double dt);
void AngMomCostFillBlock(const SplineHolder& s, const SingleRigidBodyDynamics& model, double weight, double dt,
                         const std::string& var_set, double* block);

}  // namespace orc
