// ORACLE — TEST INFRASTRUCTURE ONLY (see towr_oracle.h header).
// Node-based constraint sets (plain ifopt::ConstraintSet in the reference).
#include <cmath>

#include "constraints.h"

namespace orc {

using NVI = NodesVariables::NodeValueInfo;

// ====================================================== terrain_constraint.cc
// :36-46
TerrainConstraint::TerrainConstraint(const std::shared_ptr<HeightMap>& terrain, const std::string& ee_motion, double min_h, double max_h)
    : ConstraintSet(-1, "terrain-" + ee_motion) {
  ee_motion_id_ = ee_motion;
  terrain_ = terrain;
  min_swing_height_ = min_h;
  max_swing_height_ = max_h;
}
// :48-59  (first node skipped: constrained by the initial stance)
void TerrainConstraint::InitVariableDependedQuantities(const std::shared_ptr<Variables>& x) {
  ee_motion_ = x->GetComponent<NodesVariablesPhaseBased>(ee_motion_id_);
  for (int id = 1; id < static_cast<int>(ee_motion_->GetNodes().size()); ++id) node_ids_.push_back(id);
  SetRows(static_cast<int>(node_ids_.size()));
}
// :61-74
Vec TerrainConstraint::GetValues() const {
  Vec g(GetRows());
  const auto& nodes = ee_motion_->GetNodes();
  int row = 0;
  for (int id : node_ids_) {
    V3 p = nodes.at(id).p;
    g[row++] = p.z() - terrain_->GetHeight(p.x(), p.y());
  }
  return g;
}
// :76-91
VecBound TerrainConstraint::GetBounds() const {
  VecBound bounds(GetRows());
  int row = 0;
  for (int id : node_ids_) {
    if (ee_motion_->IsConstantNode(id)) bounds.at(row) = BoundZero();
    else bounds.at(row) = Bounds(min_swing_height_, max_swing_height_);
    row++;
  }
  return bounds;
}
// :93-111
void TerrainConstraint::FillJacobianBlock(const std::string& var_set, SpMat& jac) const {
  if (var_set != ee_motion_->GetName()) return;
  const auto& nodes = ee_motion_->GetNodes();
  int row = 0;
  for (int id : node_ids_) {
    int idx = ee_motion_->GetOptIndex(NVI{id, kPos, Z});
    jac.coeffRef(row, idx) = 1.0;
    V3 p = nodes.at(id).p;
    for (int dim : {X, Y}) {
      int idx2 = ee_motion_->GetOptIndex(NVI{id, kPos, dim});
      jac.coeffRef(row, idx2) = -terrain_->GetDerivativeOfHeightWrt(dim, p.x(), p.y());
    }
    row++;
  }
}

// ======================================================== swing_constraint.cc
SwingConstraint::SwingConstraint(const std::string& ee_motion) : ConstraintSet(-1, "swing-" + ee_motion) { ee_motion_id_ = ee_motion; }
// :41-52
void SwingConstraint::InitVariableDependedQuantities(const std::shared_ptr<Variables>& x) {
  ee_motion_ = x->GetComponent<NodesVariablesPhaseBased>(ee_motion_id_);
  pure_swing_node_ids_ = ee_motion_->GetIndicesOfNonConstantNodes();
  SetRows(static_cast<int>(pure_swing_node_ids_.size()) * 2 * k2D);
}
// :54-78
Vec SwingConstraint::GetValues() const {
  Vec g(GetRows());
  int row = 0;
  const auto& nodes = ee_motion_->GetNodes();
  for (int node_id : pure_swing_node_ids_) {
    const Node& curr = nodes.at(node_id);
    const V3& prev = nodes.at(node_id - 1).p;
    const V3& next = nodes.at(node_id + 1).p;
    for (int dim : {X, Y}) {
      double distance = next(dim) - prev(dim);
      double center = prev(dim) + 0.5 * distance;
      double des_vel_center = distance / t_swing_avg_;
      g[row++] = curr.p(dim) - center;
      g[row++] = curr.v(dim) - des_vel_center;
    }
  }
  return g;
}
// :80-84
VecBound SwingConstraint::GetBounds() const { return VecBound(GetRows(), BoundZero()); }
// :86-108
void SwingConstraint::FillJacobianBlock(const std::string& var_set, SpMat& jac) const {
  if (var_set != ee_motion_->GetName()) return;
  int row = 0;
  for (int node_id : pure_swing_node_ids_)
    for (int dim : {X, Y}) {
      jac.coeffRef(row, ee_motion_->GetOptIndex(NVI{node_id, kPos, dim})) = 1.0;
      jac.coeffRef(row, ee_motion_->GetOptIndex(NVI{node_id + 1, kPos, dim})) = -0.5;
      jac.coeffRef(row, ee_motion_->GetOptIndex(NVI{node_id - 1, kPos, dim})) = -0.5;
      row++;
      jac.coeffRef(row, ee_motion_->GetOptIndex(NVI{node_id, kVel, dim})) = 1.0;
      jac.coeffRef(row, ee_motion_->GetOptIndex(NVI{node_id + 1, kPos, dim})) = -1.0 / t_swing_avg_;
      jac.coeffRef(row, ee_motion_->GetOptIndex(NVI{node_id - 1, kPos, dim})) = +1.0 / t_swing_avg_;
      row++;
    }
}

// =================================================== spline_acc_constraint.cc
// :34-46
SplineAccConstraint::SplineAccConstraint(const std::shared_ptr<NodeSpline>& spline, const std::string& node_variable_name)
    : ConstraintSet(-1, "splineacc-" + node_variable_name) {
  spline_ = spline;
  node_variables_id_ = node_variable_name;
  n_dim_ = k3D;
  n_junctions_ = spline->GetPolynomialCount() - 1;
  T_ = spline->GetPolyDurations();
  SetRows(n_dim_ * n_junctions_);
}
// :48-64
Vec SplineAccConstraint::GetValues() const {
  Vec g(GetRows());
  for (int j = 0; j < n_junctions_; ++j) {
    V3 acc_prev = spline_->GetPoint(j, T_.at(j)).a;
    V3 acc_next = spline_->GetPoint(j + 1, 0.0).a;
    V3 d = acc_prev - acc_next;
    for (int k = 0; k < n_dim_; ++k) g[j * n_dim_ + k] = d(k);
  }
  return g;
}
// :66-80
void SplineAccConstraint::FillJacobianBlock(const std::string& var_set, SpMat& jac) const {
  if (var_set != node_variables_id_) return;
  for (int j = 0; j < n_junctions_; ++j) {
    SpMat acc_prev = spline_->GetJacobianWrtNodes(j, T_.at(j), kAcc);
    SpMat acc_next = spline_->GetJacobianWrtNodes(j + 1, 0.0, kAcc);
    jac.setMiddleRows(j * n_dim_, acc_prev - acc_next);
  }
}
// :82-86
VecBound SplineAccConstraint::GetBounds() const { return VecBound(GetRows(), BoundZero()); }

// ================================================== base_height_constraint.cc
BaseHeightConstraint::BaseHeightConstraint(const std::shared_ptr<HeightMap>& terrain, double safety_distance, const std::string& base_lin_id)
    : ConstraintSet(-1, "base-height-" + base_lin_id) {
  base_lin_id_ = base_lin_id;
  terrain_ = terrain;
  safety_distance_ = safety_distance;
}
// :45-56
void BaseHeightConstraint::InitVariableDependedQuantities(const std::shared_ptr<Variables>& x) {
  base_linear_ = x->GetComponent<NodesVariablesAll>(base_lin_id_);
  for (int id = 1; id < static_cast<int>(base_linear_->GetNodes().size()); ++id) node_ids_.push_back(id);
  SetRows(static_cast<int>(node_ids_.size()));
}
// :58-71
Vec BaseHeightConstraint::GetValues() const {
  Vec g(GetRows());
  const auto& nodes = base_linear_->GetNodes();
  int row = 0;
  for (int id : node_ids_) {
    V3 p = nodes.at(id).p;
    g[row++] = p.z() - terrain_->GetHeight(p.x(), p.y()) - safety_distance_;
  }
  return g;
}
// :73-88
VecBound BaseHeightConstraint::GetBounds() const { return VecBound(GetRows(), Bounds(0.0, 1e20)); }
// :90-110
void BaseHeightConstraint::FillJacobianBlock(const std::string& var_set, SpMat& jac) const {
  if (var_set != base_lin_id_) return;
  const auto& nodes = base_linear_->GetNodes();
  int row = 0;
  for (int id : node_ids_) {
    int idx_z = base_linear_->GetOptIndex(NVI{id, kPos, Z});
    jac.coeffRef(row, idx_z) = 1.0;
    V3 p = nodes.at(id).p;
    for (int dim : {X, Y}) {
      int idx = base_linear_->GetOptIndex(NVI{id, kPos, dim});
      jac.coeffRef(row, idx) = -terrain_->GetDerivativeOfHeightWrt(dim, p.x(), p.y());
    }
    row++;
  }
}

// =============================================== total_duration_constraint.cc
TotalDurationConstraint::TotalDurationConstraint(double T_total, int ee)
    : ConstraintSet(1, "totalduration-" + std::to_string(ee)), T_total_(T_total), ee_(ee) {}
void TotalDurationConstraint::InitVariableDependedQuantities(const std::shared_ptr<Variables>& x) {
  phase_durations_ = x->GetComponent<PhaseDurations>("ee-schedule" + std::to_string(ee_));
}
// :48-54  (sum excludes the last, implied duration)
Vec TotalDurationConstraint::GetValues() const {
  Vec g(GetRows(), 0.0);
  double s = 0.0;
  for (double v : phase_durations_->GetValues()) s += v;
  g[0] = s;
  return g;
}
// :56-64
VecBound TotalDurationConstraint::GetBounds() const { return VecBound(GetRows(), Bounds(0.1, T_total_ - 0.2)); }
// :66-72
void TotalDurationConstraint::FillJacobianBlock(const std::string& var_set, SpMat& jac) const {
  if (var_set == phase_durations_->GetName())
    for (int col = 0; col < phase_durations_->GetRows(); ++col) jac.coeffRef(0, col) = 1.0;
}

// ======================================================== force_constraint.cc
// :37-48
ForceConstraint::ForceConstraint(const std::shared_ptr<HeightMap>& terrain, double force_limit, int ee)
    : ConstraintSet(-1, "force-ee-force_" + std::to_string(ee)) {
  terrain_ = terrain;
  fn_max_ = force_limit;
  mu_ = terrain->GetFrictionCoeff();
  ee_ = ee;
  n_constraints_per_node_ = 1 + 2 * k2D;
}
// :50-60
void ForceConstraint::InitVariableDependedQuantities(const std::shared_ptr<Variables>& x) {
  ee_force_ = x->GetComponent<NodesVariablesPhaseBased>("ee-force_" + std::to_string(ee_));
  ee_motion_ = x->GetComponent<NodesVariablesPhaseBased>("ee-motion_" + std::to_string(ee_));
  pure_stance_force_node_ids_ = ee_force_->GetIndicesOfNonConstantNodes();
  SetRows(static_cast<int>(pure_stance_force_node_ids_.size()) * n_constraints_per_node_);
}
// :62-89
Vec ForceConstraint::GetValues() const {
  Vec g(GetRows());
  int row = 0;
  const auto& force_nodes = ee_force_->GetNodes();
  for (int f_node_id : pure_stance_force_node_ids_) {
    int phase = ee_force_->GetPhase(f_node_id);
    V3 p = ee_motion_->GetValueAtStartOfPhase(phase);
    V3 n = terrain_->GetNormalizedBasis(HeightMap::Normal, p.x(), p.y());
    V3 f = force_nodes.at(f_node_id).p;
    g[row++] = dot(f, n);
    V3 t1 = terrain_->GetNormalizedBasis(HeightMap::Tangent1, p.x(), p.y());
    g[row++] = dot(f, t1 - mu_ * n);
    g[row++] = dot(f, t1 + mu_ * n);
    V3 t2 = terrain_->GetNormalizedBasis(HeightMap::Tangent2, p.x(), p.y());
    g[row++] = dot(f, t2 - mu_ * n);
    g[row++] = dot(f, t2 + mu_ * n);
  }
  return g;
}
// :91-105
VecBound ForceConstraint::GetBounds() const {
  VecBound bounds;
  for (size_t i = 0; i < pure_stance_force_node_ids_.size(); ++i) {
    bounds.push_back(Bounds(0.0, fn_max_));
    bounds.push_back(BoundSmallerZero());
    bounds.push_back(BoundGreaterZero());
    bounds.push_back(BoundSmallerZero());
    bounds.push_back(BoundGreaterZero());
  }
  return bounds;
}
// :107-171
void ForceConstraint::FillJacobianBlock(const std::string& var_set, SpMat& jac) const {
  if (var_set == ee_force_->GetName()) {
    int row = 0;
    for (int f_node_id : pure_stance_force_node_ids_) {
      int phase = ee_force_->GetPhase(f_node_id);
      V3 p = ee_motion_->GetValueAtStartOfPhase(phase);
      V3 n = terrain_->GetNormalizedBasis(HeightMap::Normal, p.x(), p.y());
      V3 t1 = terrain_->GetNormalizedBasis(HeightMap::Tangent1, p.x(), p.y());
      V3 t2 = terrain_->GetNormalizedBasis(HeightMap::Tangent2, p.x(), p.y());
      for (int dim : {X, Y, Z}) {
        int idx = ee_force_->GetOptIndex(NVI{f_node_id, kPos, dim});
        int rr = row;
        jac.coeffRef(rr++, idx) = n(dim);
        jac.coeffRef(rr++, idx) = t1(dim) - mu_ * n(dim);
        jac.coeffRef(rr++, idx) = t1(dim) + mu_ * n(dim);
        jac.coeffRef(rr++, idx) = t2(dim) - mu_ * n(dim);
        jac.coeffRef(rr++, idx) = t2(dim) + mu_ * n(dim);
      }
      row += n_constraints_per_node_;
    }
  }
  if (var_set == ee_motion_->GetName()) {
    int row = 0;
    const auto& force_nodes = ee_force_->GetNodes();
    for (int f_node_id : pure_stance_force_node_ids_) {
      int phase = ee_force_->GetPhase(f_node_id);
      int ee_node_id = ee_motion_->GetNodeIDAtStartOfPhase(phase);
      V3 p = ee_motion_->GetValueAtStartOfPhase(phase);
      V3 f = force_nodes.at(f_node_id).p;
      for (int dim : {0, 1}) {
        V3 dn = terrain_->GetDerivativeOfNormalizedBasisWrt(HeightMap::Normal, dim, p.x(), p.y());
        V3 dt1 = terrain_->GetDerivativeOfNormalizedBasisWrt(HeightMap::Tangent1, dim, p.x(), p.y());
        V3 dt2 = terrain_->GetDerivativeOfNormalizedBasisWrt(HeightMap::Tangent2, dim, p.x(), p.y());
        int idx = ee_motion_->GetOptIndex(NVI{ee_node_id, kPos, dim});
        int rr = row;
        jac.coeffRef(rr++, idx) = dot(f, dn);
        jac.coeffRef(rr++, idx) = dot(f, dt1 - mu_ * dn);
        jac.coeffRef(rr++, idx) = dot(f, dt1 + mu_ * dn);
        jac.coeffRef(rr++, idx) = dot(f, dt2 - mu_ * dn);
        jac.coeffRef(rr++, idx) = dot(f, dt2 + mu_ * dn);
      }
      row += n_constraints_per_node_;
    }
  }
}

// ------------------------------------------------------------------ TorqueConstraint (node-based)
// torque_constraint.cc:34-51
TorqueConstraint::TorqueConstraint(const std::shared_ptr<HeightMap>& terrain, double tx_min, double tx_max, double ty_min,
                                   double ty_max, double k_friction, int ee)
    : ConstraintSet(-1, "torque-ee-torque_" + std::to_string(ee)) {
  terrain_ = terrain;
  tx_min_ = tx_min; tx_max_ = tx_max; ty_min_ = ty_min; ty_max_ = ty_max;
  k_friction_ = k_friction;
  mu_ = terrain->GetFrictionCoeff();
  ee_ = ee;
  n_constraints_per_node_ = 3;
}
// :53-62
void TorqueConstraint::InitVariableDependedQuantities(const std::shared_ptr<Variables>& x) {
  ee_torque_ = x->GetComponent<NodesVariablesPhaseBased>("ee-torque_" + std::to_string(ee_));
  ee_motion_ = x->GetComponent<NodesVariablesPhaseBased>("ee-motion_" + std::to_string(ee_));
  ee_force_ = x->GetComponent<NodesVariablesPhaseBased>("ee-force_" + std::to_string(ee_));
  pure_stance_torque_node_ids_ = ee_torque_->GetIndicesOfNonConstantNodes();
  SetRows(static_cast<int>(pure_stance_torque_node_ids_.size()) * n_constraints_per_node_);
}
// :64-92
Vec TorqueConstraint::GetValues() const {
  Vec g(GetRows());
  int row = 0;
  const auto& torque_nodes = ee_torque_->GetNodes();
  for (int t_node_id : pure_stance_torque_node_ids_) {
    int phase = ee_torque_->GetPhase(t_node_id);
    V3 p = ee_motion_->GetValueAtStartOfPhase(phase);
    V3 tau = torque_nodes.at(t_node_id).p;
    V3 n = terrain_->GetNormalizedBasis(HeightMap::Normal, p.x(), p.y());
    V3 t1 = terrain_->GetNormalizedBasis(HeightMap::Tangent1, p.x(), p.y());
    V3 t2 = terrain_->GetNormalizedBasis(HeightMap::Tangent2, p.x(), p.y());
    g[row++] = dot(tau, t1);
    g[row++] = dot(tau, t2);
    g[row++] = dot(tau, n);
  }
  return g;
}
// :94-112 - the friction-moment bound uses the force at the start of the phase AS IT IS when the bounds are asked for
VecBound TorqueConstraint::GetBounds() const {
  VecBound bounds;
  for (int t_node_id : pure_stance_torque_node_ids_) {
    int phase = ee_torque_->GetPhase(t_node_id);
    V3 f = ee_force_->GetValueAtStartOfPhase(phase);
    V3 p = ee_motion_->GetValueAtStartOfPhase(phase);
    V3 n = terrain_->GetNormalizedBasis(HeightMap::Normal, p.x(), p.y());
    double f_n = dot(f, n);
    bounds.push_back(Bounds(tx_min_, tx_max_));
    bounds.push_back(Bounds(ty_min_, ty_max_));
    double tz_limit = k_friction_ * mu_ * f_n;
    bounds.push_back(Bounds(-tz_limit, tz_limit));
  }
  return bounds;
}
// :114-157 - the ee-motion block uses the torque at the START of the phase (:141), not the row's node; no ee-force block
void TorqueConstraint::FillJacobianBlock(const std::string& var_set, SpMat& jac) const {
  if (var_set == ee_torque_->GetName()) {
    int row = 0;
    for (int t_node_id : pure_stance_torque_node_ids_) {
      int phase = ee_torque_->GetPhase(t_node_id);
      V3 p = ee_motion_->GetValueAtStartOfPhase(phase);
      V3 n = terrain_->GetNormalizedBasis(HeightMap::Normal, p.x(), p.y());
      V3 t1 = terrain_->GetNormalizedBasis(HeightMap::Tangent1, p.x(), p.y());
      V3 t2 = terrain_->GetNormalizedBasis(HeightMap::Tangent2, p.x(), p.y());
      for (int dim : {X, Y, Z}) {
        int idx = ee_torque_->GetOptIndex(NVI{t_node_id, kPos, dim});
        int rr = row;
        jac.coeffRef(rr++, idx) = t1(dim);
        jac.coeffRef(rr++, idx) = t2(dim);
        jac.coeffRef(rr++, idx) = n(dim);
      }
      row += n_constraints_per_node_;
    }
  }
  if (var_set == ee_motion_->GetName()) {
    int row = 0;
    for (int t_node_id : pure_stance_torque_node_ids_) {
      int phase = ee_torque_->GetPhase(t_node_id);
      int ee_node_id = ee_motion_->GetNodeIDAtStartOfPhase(phase);
      V3 p = ee_motion_->GetValueAtStartOfPhase(phase);
      V3 tau = ee_torque_->GetValueAtStartOfPhase(phase);
      for (int dim : {0, 1}) {
        V3 dt1 = terrain_->GetDerivativeOfNormalizedBasisWrt(HeightMap::Tangent1, dim, p.x(), p.y());
        V3 dt2 = terrain_->GetDerivativeOfNormalizedBasisWrt(HeightMap::Tangent2, dim, p.x(), p.y());
        V3 dn = terrain_->GetDerivativeOfNormalizedBasisWrt(HeightMap::Normal, dim, p.x(), p.y());
        int idx = ee_motion_->GetOptIndex(NVI{ee_node_id, kPos, dim});
        int rr = row;
        jac.coeffRef(rr++, idx) = dot(tau, dt1);
        jac.coeffRef(rr++, idx) = dot(tau, dt2);
        jac.coeffRef(rr++, idx) = dot(tau, dn);
      }
      row += n_constraints_per_node_;
    }
  }
}

}  // namespace orc
