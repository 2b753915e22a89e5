// ORACLE — TEST INFRASTRUCTURE ONLY (see towr_oracle.h header).
// Time-discretised constraint sets.
#include <algorithm>
#include <cmath>

#include "constraints.h"

namespace orc {

namespace id {   // variables/variable_names.h:43-75
static const std::string base_lin_nodes = "base-lin";
static const std::string base_ang_nodes = "base-ang";
static std::string EEMotionNodes(int ee) { return "ee-motion_" + std::to_string(ee); }
static std::string EEForceNodes(int ee) { return "ee-force_" + std::to_string(ee); }
static std::string EETorqueNodes(int ee) { return "ee-torque_" + std::to_string(ee); }
static std::string EESchedule(int ee) { return "ee-schedule" + std::to_string(ee); }
}  // namespace id

// ===================================== time_discretization_constraint.cc
// :37-49   sample times by repeated addition, final T appended (possibly a duplicate)
TimeDiscretizationConstraint::TimeDiscretizationConstraint(double T, double dt, const std::string& name)
    : ConstraintSet(-1, name) {
  double t = 0.0;
  dts_ = {t};
  for (int i = 0; i < floor(T / dt); ++i) {
    t += dt;
    dts_.push_back(t);
  }
  dts_.push_back(T);
}
// :65-75
Vec TimeDiscretizationConstraint::GetValues() const {
  Vec g(GetRows(), 0.0);
  int k = 0;
  for (double t : dts_) UpdateConstraintAtInstance(t, k++, g);
  return g;
}
// :77-87
VecBound TimeDiscretizationConstraint::GetBounds() const {
  VecBound bounds(GetRows());
  int k = 0;
  for (double t : dts_) UpdateBoundsAtInstance(t, k++, bounds);
  return bounds;
}
// :89-96
void TimeDiscretizationConstraint::FillJacobianBlock(const std::string& var_set, SpMat& jac) const {
  int k = 0;
  for (double t : dts_) UpdateJacobianAtInstance(t, k++, var_set, jac);
}

// ====================================================== dynamic_constraint.cc
// :38-55
DynamicConstraint::DynamicConstraint(const std::shared_ptr<SingleRigidBodyDynamics>& m, double T, double dt, const SplineHolder& s)
    : TimeDiscretizationConstraint(T, dt, "dynamic") {
  model_ = m;
  base_linear_ = s.base_linear_;
  base_angular_ = s.angular_converter_ ? s.angular_converter_ : std::make_shared<EulerConverter>(s.base_angular_);
  ee_forces_ = s.ee_force_;
  ee_torques_ = s.ee_torque_;
  ee_motion_ = s.ee_motion_;
  SetRows(GetNumberOfNodes() * k6D);
}
// :63-68
void DynamicConstraint::UpdateConstraintAtInstance(double t, int k, Vec& g) const {
  UpdateModel(t);
  auto acc = model_->GetDynamicViolation();
  for (int d = 0; d < k6D; ++d) g[GetRow(k, AX) + d] = acc[d];
}
// :70-75
void DynamicConstraint::UpdateBoundsAtInstance(double, int k, VecBound& bounds) const {
  for (int dim = 0; dim < k6D; ++dim) bounds.at(GetRow(k, dim)) = BoundZero();
}
// :77-126
void DynamicConstraint::UpdateJacobianAtInstance(double t, int k, const std::string& var_set, SpMat& jac) const {
  UpdateModel(t);
  int n = jac.cols;
  SpMat jac_model(k6D, n);

  if (var_set == id::base_lin_nodes) {
    SpMat jac_base_lin_pos = base_linear_->GetJacobianWrtNodes(t, kPos);
    SpMat jac_base_lin_acc = base_linear_->GetJacobianWrtNodes(t, kAcc);
    jac_model = model_->GetJacobianWrtBaseLin(jac_base_lin_pos, jac_base_lin_acc);
  }
  if (var_set == id::base_ang_nodes) jac_model = model_->GetJacobianWrtBaseAng(*base_angular_, t);

  for (int ee = 0; ee < model_->GetEECount(); ++ee) {
    if (var_set == id::EEForceNodes(ee)) {
      SpMat jac_ee_force = ee_forces_.at(ee)->GetJacobianWrtNodes(t, kPos);
      jac_model = model_->GetJacobianWrtForce(jac_ee_force, ee);
    }
    if (var_set == id::EETorqueNodes(ee)) {
      SpMat jac_ee_torque = ee_torques_.at(ee)->GetJacobianWrtNodes(t, kPos);
      jac_model = model_->GetJacobianWrtTorque(jac_ee_torque, ee);
    }
    if (var_set == id::EEMotionNodes(ee)) {
      SpMat jac_ee_pos = ee_motion_.at(ee)->GetJacobianWrtNodes(t, kPos);
      jac_model = model_->GetJacobianWrtEEPos(jac_ee_pos, ee);
    }
    if (var_set == id::EESchedule(ee)) {
      SpMat jac_f_dT = ee_forces_.at(ee)->GetJacobianOfPosWrtDurations(t);
      jac_model += model_->GetJacobianWrtForce(jac_f_dT, ee);
      SpMat jac_x_dT = ee_motion_.at(ee)->GetJacobianOfPosWrtDurations(t);
      jac_model += model_->GetJacobianWrtEEPos(jac_x_dT, ee);
    }
  }
  jac.setMiddleRows(GetRow(k, AX), jac_model);
}
// :128-148
void DynamicConstraint::UpdateModel(double t) const {
  State3 com = base_linear_->GetPoint(t);
  M3 w_R_b = toDense(base_angular_->GetRotationMatrixBaseToWorld(t));
  V3 omega = base_angular_->GetAngularVelocityInWorld(t);
  V3 omega_dot = base_angular_->GetAngularAccelerationInWorld(t);
  int n_ee = model_->GetEECount();
  std::vector<V3> ee_pos, ee_force, ee_torque;
  for (int ee = 0; ee < n_ee; ++ee) {
    ee_force.push_back(ee_forces_.at(ee)->GetPoint(t).p);
    ee_torque.push_back(ee_torques_.at(ee)->GetPoint(t).p);
    ee_pos.push_back(ee_motion_.at(ee)->GetPoint(t).p);
  }
  model_->SetCurrent(com.p, com.a, w_R_b, omega, omega_dot, ee_force, ee_pos, ee_torque);
}

// ============================================== range_of_motion_constraint.cc
// :37-56
RangeOfMotionConstraint::RangeOfMotionConstraint(const std::shared_ptr<KinematicModel>& model, double T, double dt, int ee, const SplineHolder& s)
    : TimeDiscretizationConstraint(T, dt, "rangeofmotion-" + std::to_string(ee)) {
  base_linear_ = s.base_linear_;
  base_angular_ = s.angular_converter_ ? s.angular_converter_ : std::make_shared<EulerConverter>(s.base_angular_);
  ee_motion_ = s.ee_motion_.at(ee);
  max_deviation_from_nominal_ = model->max_dev_.at(ee);
  min_deviation_from_nominal_ = model->min_dev_.at(ee);
  nominal_ee_pos_B_ = model->nominal_stance_.at(ee);
  ee_ = ee;
  SetRows(GetNumberOfNodes() * k3D);
}
// :72-83
void RangeOfMotionConstraint::UpdateConstraintAtInstance(double t, int k, Vec& g) const {
  V3 base_W = base_linear_->GetPoint(t).p;
  V3 pos_ee_W = ee_motion_->GetPoint(t).p;
  SpMat b_R_w = base_angular_->GetRotationMatrixBaseToWorld(t).transpose();
  V3 vector_base_to_ee_W = pos_ee_W - base_W;
  V3 vector_base_to_ee_B = b_R_w * vector_base_to_ee_W;
  for (int d = 0; d < k3D; ++d) g[GetRow(k, X) + d] = vector_base_to_ee_B(d);
}
// :85-103 (swing relaxation is off unless rom_swing_relax_dims_ is set; not restated)
void RangeOfMotionConstraint::UpdateBoundsAtInstance(double, int k, VecBound& bounds) const {
  for (int dim = 0; dim < k3D; ++dim) {
    Bounds b;
    b += nominal_ee_pos_B_(dim);
    b.upper += max_deviation_from_nominal_(dim);
    b.lower += min_deviation_from_nominal_(dim);
    bounds.at(GetRow(k, dim)) = b;
  }
}
// :105-131
void RangeOfMotionConstraint::UpdateJacobianAtInstance(double t, int k, const std::string& var_set, SpMat& jac) const {
  SpMat b_R_w = base_angular_->GetRotationMatrixBaseToWorld(t).transpose();
  int row_start = GetRow(k, X);
  if (var_set == id::base_lin_nodes)
    jac.setMiddleRows(row_start, (-1 * b_R_w) * base_linear_->GetJacobianWrtNodes(t, kPos));
  if (var_set == id::base_ang_nodes) {
    V3 base_W = base_linear_->GetPoint(t).p;
    V3 ee_pos_W = ee_motion_->GetPoint(t).p;
    V3 r_W = ee_pos_W - base_W;
    jac.setMiddleRows(row_start, base_angular_->DerivOfRotVecMult(t, r_W, true));
  }
  if (var_set == id::EEMotionNodes(ee_))
    jac.setMiddleRows(row_start, b_R_w * ee_motion_->GetJacobianWrtNodes(t, kPos));
  if (var_set == id::EESchedule(ee_))
    jac.setMiddleRows(row_start, b_R_w * ee_motion_->GetJacobianOfPosWrtDurations(t));
}

// ========================================== force/torque_constraint_discretized.cc
// force_constraint_discretized.cc:38-50  out(dst_row, col) += b(r) * J(r, col) over all stored entries
static void AccumulateLinearFormJacobian(const SpMat& J3xN, const V3& b, int dst_row, SpMat& out) {
  for (int outer = 0; outer < J3xN.rows; ++outer)
    for (const auto& en : J3xN.r[outer]) out.coeffRef(dst_row, en.first) += b(outer) * en.second;
}
// force_constraint_discretized.cc:52-67  skipped entirely when scale == 0.0 (value-dependent pattern)
static void AccumulateScaledRowJacobian(const SpMat& J3xN, int src_row, double scale, int dst_row, SpMat& out) {
  if (scale == 0.0) return;
  for (int outer = 0; outer < J3xN.rows; ++outer)
    for (const auto& en : J3xN.r[outer])
      if (outer == src_row) out.coeffRef(dst_row, en.first) += scale * en.second;
}

// force_constraint_discretized.cc:71-89
ForceConstraintDiscretized::ForceConstraintDiscretized(const std::shared_ptr<HeightMap>& terrain, double T, double dt,
                                                       double force_limit, int ee, const SplineHolder& s)
    : TimeDiscretizationConstraint(T, dt, "force-disc-" + id::EEForceNodes(ee)) {
  terrain_ = terrain;
  fn_max_ = force_limit;
  mu_ = terrain->GetFrictionCoeff();
  ee_ = ee;
  ee_force_ = s.ee_force_.at(ee_);
  ee_motion_ = s.ee_motion_.at(ee_);
  n_constraints_per_instance_ = 1 + 2 * k2D;
  SetRows(GetNumberOfNodes() * n_constraints_per_instance_);
}
// :97-117
void ForceConstraintDiscretized::UpdateConstraintAtInstance(double t, int k, Vec& g) const {
  V3 p = ee_motion_->GetPoint(t).p;
  V3 f = ee_force_->GetPoint(t).p;
  V3 n = terrain_->GetNormalizedBasis(HeightMap::Normal, p.x(), p.y());
  V3 t1 = terrain_->GetNormalizedBasis(HeightMap::Tangent1, p.x(), p.y());
  V3 t2 = terrain_->GetNormalizedBasis(HeightMap::Tangent2, p.x(), p.y());
  int row = GetRow(k, 0);
  g[row++] = dot(f, n);
  g[row++] = dot(f, t1 - mu_ * n);
  g[row++] = dot(f, t1 + mu_ * n);
  g[row++] = dot(f, t2 - mu_ * n);
  g[row++] = dot(f, t2 + mu_ * n);
}
// :119-129
void ForceConstraintDiscretized::UpdateBoundsAtInstance(double, int k, VecBound& bounds) const {
  int row = GetRow(k, 0);
  bounds.at(row++) = Bounds(0.0, fn_max_);
  bounds.at(row++) = BoundSmallerZero();
  bounds.at(row++) = BoundGreaterZero();
  bounds.at(row++) = BoundSmallerZero();
  bounds.at(row++) = BoundGreaterZero();
}
// :131-221
void ForceConstraintDiscretized::UpdateJacobianAtInstance(double t, int k, const std::string& var_set, SpMat& jac) const {
  const V3 p = ee_motion_->GetPoint(t).p;
  const V3 f = ee_force_->GetPoint(t).p;
  const V3 n = terrain_->GetNormalizedBasis(HeightMap::Normal, p.x(), p.y());
  const V3 t1 = terrain_->GetNormalizedBasis(HeightMap::Tangent1, p.x(), p.y());
  const V3 t2 = terrain_->GetNormalizedBasis(HeightMap::Tangent2, p.x(), p.y());
  const V3 b[5] = {n, t1 - mu_ * n, t1 + mu_ * n, t2 - mu_ * n, t2 + mu_ * n};
  int r[5];
  for (int i = 0; i < 5; ++i) r[i] = GetRow(k, i);

  auto through_basis = [&](const SpMat& Jp) {
    for (int dim : {0, 1}) {
      V3 dn = terrain_->GetDerivativeOfNormalizedBasisWrt(HeightMap::Normal, dim, p.x(), p.y());
      V3 dt1 = terrain_->GetDerivativeOfNormalizedBasisWrt(HeightMap::Tangent1, dim, p.x(), p.y());
      V3 dt2 = terrain_->GetDerivativeOfNormalizedBasisWrt(HeightMap::Tangent2, dim, p.x(), p.y());
      const double s[5] = {dot(f, dn), dot(f, dt1 - mu_ * dn), dot(f, dt1 + mu_ * dn), dot(f, dt2 - mu_ * dn), dot(f, dt2 + mu_ * dn)};
      for (int i = 0; i < 5; ++i) AccumulateScaledRowJacobian(Jp, dim, s[i], r[i], jac);
    }
  };

  if (var_set == id::EEForceNodes(ee_)) {
    SpMat Jf = ee_force_->GetJacobianWrtNodes(t, kPos);
    for (int i = 0; i < 5; ++i) AccumulateLinearFormJacobian(Jf, b[i], r[i], jac);
  }
  if (var_set == id::EEMotionNodes(ee_)) through_basis(ee_motion_->GetJacobianWrtNodes(t, kPos));
  if (var_set == id::EESchedule(ee_)) {
    SpMat Jf_dT = ee_force_->GetJacobianOfPosWrtDurations(t);
    SpMat Jp_dT = ee_motion_->GetJacobianOfPosWrtDurations(t);
    for (int i = 0; i < 5; ++i) AccumulateLinearFormJacobian(Jf_dT, b[i], r[i], jac);
    through_basis(Jp_dT);
  }
}

// torque_constraint_discretized.cc:68-97
TorqueConstraintDiscretized::TorqueConstraintDiscretized(const std::shared_ptr<HeightMap>& terrain, double T, double dt,
                                                         double tx_min, double tx_max, double ty_min, double ty_max,
                                                         double k_friction, int ee, const SplineHolder& s)
    : TimeDiscretizationConstraint(T, dt, "torque-disc-" + id::EETorqueNodes(ee)) {
  terrain_ = terrain;
  tx_min_ = tx_min; tx_max_ = tx_max; ty_min_ = ty_min; ty_max_ = ty_max;
  k_friction_ = k_friction;
  mu_ = terrain->GetFrictionCoeff();
  ee_ = ee;
  ee_torque_ = s.ee_torque_.at(ee_);
  ee_force_ = s.ee_force_.at(ee_);
  ee_motion_ = s.ee_motion_.at(ee_);
  SetRows(GetNumberOfNodes() * 4);
}
// :105-129
void TorqueConstraintDiscretized::UpdateConstraintAtInstance(double t, int k, Vec& g) const {
  const V3 p = ee_motion_->GetPoint(t).p;
  const V3 f = ee_force_->GetPoint(t).p;
  const V3 tau = ee_torque_->GetPoint(t).p;
  const V3 n = terrain_->GetNormalizedBasis(HeightMap::Normal, p.x(), p.y());
  const V3 t1 = terrain_->GetNormalizedBasis(HeightMap::Tangent1, p.x(), p.y());
  const V3 t2 = terrain_->GetNormalizedBasis(HeightMap::Tangent2, p.x(), p.y());
  const double tau_t1 = dot(tau, t1), tau_t2 = dot(tau, t2), tau_n = dot(tau, n), f_n = dot(f, n);
  const double tz_lim = k_friction_ * mu_ * f_n;
  int row = GetRow(k, 0);
  g[row++] = tau_t1;
  g[row++] = tau_t2;
  g[row++] = tau_n - tz_lim;
  g[row++] = -tau_n - tz_lim;
}
// :131-139
void TorqueConstraintDiscretized::UpdateBoundsAtInstance(double, int k, VecBound& bounds) const {
  int row = GetRow(k, 0);
  bounds.at(row++) = Bounds(tx_min_, tx_max_);
  bounds.at(row++) = Bounds(ty_min_, ty_max_);
  bounds.at(row++) = BoundSmallerZero();
  bounds.at(row++) = BoundSmallerZero();
}
// :141-235
void TorqueConstraintDiscretized::UpdateJacobianAtInstance(double t, int k, const std::string& var_set, SpMat& jac) const {
  const V3 p = ee_motion_->GetPoint(t).p;
  const V3 f = ee_force_->GetPoint(t).p;
  const V3 tau = ee_torque_->GetPoint(t).p;
  const V3 n = terrain_->GetNormalizedBasis(HeightMap::Normal, p.x(), p.y());
  const V3 t1 = terrain_->GetNormalizedBasis(HeightMap::Tangent1, p.x(), p.y());
  const V3 t2 = terrain_->GetNormalizedBasis(HeightMap::Tangent2, p.x(), p.y());
  const int r0 = GetRow(k, 0), r1 = GetRow(k, 1), r2 = GetRow(k, 2), r3 = GetRow(k, 3);

  auto through_basis = [&](const SpMat& Jp) {
    for (int dim : {0, 1}) {
      const V3 dn = terrain_->GetDerivativeOfNormalizedBasisWrt(HeightMap::Normal, dim, p.x(), p.y());
      const V3 dt1 = terrain_->GetDerivativeOfNormalizedBasisWrt(HeightMap::Tangent1, dim, p.x(), p.y());
      const V3 dt2 = terrain_->GetDerivativeOfNormalizedBasisWrt(HeightMap::Tangent2, dim, p.x(), p.y());
      const double s_tx = dot(tau, dt1), s_ty = dot(tau, dt2);
      const double s_tau_n = dot(tau, dn), s_f_n = dot(f, dn);
      const double s_lim = k_friction_ * mu_ * s_f_n;
      AccumulateScaledRowJacobian(Jp, dim, s_tx, r0, jac);
      AccumulateScaledRowJacobian(Jp, dim, s_ty, r1, jac);
      AccumulateScaledRowJacobian(Jp, dim, (s_tau_n - s_lim), r2, jac);
      AccumulateScaledRowJacobian(Jp, dim, (-s_tau_n - s_lim), r3, jac);
    }
  };

  if (var_set == id::EETorqueNodes(ee_)) {
    SpMat Jtau = ee_torque_->GetJacobianWrtNodes(t, kPos);
    AccumulateLinearFormJacobian(Jtau, t1, r0, jac);
    AccumulateLinearFormJacobian(Jtau, t2, r1, jac);
    AccumulateLinearFormJacobian(Jtau, n, r2, jac);
    AccumulateLinearFormJacobian(Jtau, -n, r3, jac);
  }
  if (var_set == id::EEForceNodes(ee_)) {
    SpMat Jf = ee_force_->GetJacobianWrtNodes(t, kPos);
    const V3 b = (-k_friction_ * mu_) * n;
    AccumulateLinearFormJacobian(Jf, b, r2, jac);
    AccumulateLinearFormJacobian(Jf, b, r3, jac);
  }
  if (var_set == id::EEMotionNodes(ee_)) through_basis(ee_motion_->GetJacobianWrtNodes(t, kPos));
  if (var_set == id::EESchedule(ee_)) {
    SpMat Jtau_dT = ee_torque_->GetJacobianOfPosWrtDurations(t);
    SpMat Jf_dT = ee_force_->GetJacobianOfPosWrtDurations(t);
    SpMat Jp_dT = ee_motion_->GetJacobianOfPosWrtDurations(t);
    AccumulateLinearFormJacobian(Jtau_dT, t1, r0, jac);
    AccumulateLinearFormJacobian(Jtau_dT, t2, r1, jac);
    AccumulateLinearFormJacobian(Jtau_dT, n, r2, jac);
    AccumulateLinearFormJacobian(Jtau_dT, -n, r3, jac);
    const V3 b = (-k_friction_ * mu_) * n;
    AccumulateLinearFormJacobian(Jf_dT, b, r2, jac);
    AccumulateLinearFormJacobian(Jf_dT, b, r3, jac);
    through_basis(Jp_dT);
  }
}

// ================================================= terrain_constraint_hard.cc
// :37-49
TerrainConstraintHard::TerrainConstraintHard(const std::shared_ptr<HeightMap>& terrain, double T, double dt, int ee, const SplineHolder& s)
    : TimeDiscretizationConstraint(T, dt, "terrainhard-" + std::to_string(ee)) {
  ee_motion_ = s.ee_motion_.at(ee);
  terrain_ = terrain;
  ee_ = ee;
  k_coeff_ = 0.02;
  SetRows(GetNumberOfNodes());
}
// :51-75
void TerrainConstraintHard::UpdateConstraintAtInstance(double t, int k, Vec& g) const {
  V3 p = ee_motion_->GetPoint(t).p;
  V3 v = ee_motion_->GetPoint(t).v;
  V3 t1 = terrain_->GetNormalizedBasis(HeightMap::Tangent1, p.x(), p.y());
  V3 t2 = terrain_->GetNormalizedBasis(HeightMap::Tangent2, p.x(), p.y());
  double v_tangent1 = dot(v, t1), v_tangent2 = dot(v, t2);
  double v_tangent_magnitude = sqrt(v_tangent1 * v_tangent1 + v_tangent2 * v_tangent2);
  double delta_z = p.z() - terrain_->GetHeight(p.x(), p.y());
  double min_height = std::min(k_coeff_ * v_tangent_magnitude, k_coeff_);
  g[k] = delta_z - min_height;
}
// :77-85
void TerrainConstraintHard::UpdateBoundsAtInstance(double, int k, VecBound& bounds) const { bounds.at(k) = Bounds(0.0, 1e20); }
// :87-132
void TerrainConstraintHard::UpdateJacobianAtInstance(double t, int k, const std::string& var_set, SpMat& jac) const {
  if (var_set != id::EEMotionNodes(ee_)) return;
  V3 p = ee_motion_->GetPoint(t).p;
  V3 v = ee_motion_->GetPoint(t).v;
  V3 t1 = terrain_->GetNormalizedBasis(HeightMap::Tangent1, p.x(), p.y());
  V3 t2 = terrain_->GetNormalizedBasis(HeightMap::Tangent2, p.x(), p.y());
  double v_tangent1 = dot(v, t1), v_tangent2 = dot(v, t2);
  double v_tangent_magnitude = sqrt(v_tangent1 * v_tangent1 + v_tangent2 * v_tangent2);
  SpMat jac_pos = ee_motion_->GetJacobianWrtNodes(t, kPos);
  SpMat jac_vel = ee_motion_->GetJacobianWrtNodes(t, kVel);
  jac.setRow(k, jac_pos.row(Z));
  for (int dim : {X, Y}) {
    double terrain_deriv = terrain_->GetDerivativeOfHeightWrt(dim, p.x(), p.y());
    jac.setRow(k, jac.row(k) - terrain_deriv * jac_pos.row(dim));
  }
  if (v_tangent_magnitude > 1e-6) {
    if (k_coeff_ * v_tangent_magnitude < 0.05 - 1e-6) {
      V3 tangent_deriv = (v_tangent1 * t1 + v_tangent2 * t2) / v_tangent_magnitude;
      for (int dim : {X, Y, Z}) jac.setRow(k, jac.row(k) - (k_coeff_ * tangent_deriv(dim)) * jac_vel.row(dim));
    }
  }
}

// =================================================== base_motion_constraint.cc
// :38-58
BaseMotionConstraint::BaseMotionConstraint(double T, double dt, Bounds ax, Bounds ay, Bounds lz, const SplineHolder& s)
    : TimeDiscretizationConstraint(T, dt, "baseMotion") {
  base_linear_ = s.base_linear_;
  base_angular_ = s.base_angular_;
  node_bounds_.resize(k6D);
  node_bounds_.at(AX) = ax;
  node_bounds_.at(AY) = ay;
  node_bounds_.at(AZ) = NoBound();
  node_bounds_.at(LX) = NoBound();
  node_bounds_.at(LY) = NoBound();
  node_bounds_.at(LZ) = lz;
  SetRows(GetNumberOfNodes() * static_cast<int>(node_bounds_.size()));
}
// :60-66
void BaseMotionConstraint::UpdateConstraintAtInstance(double t, int k, Vec& g) const {
  V3 lin = base_linear_->GetPoint(t).p, ang = base_angular_->GetPoint(t).p;
  for (int d = 0; d < k3D; ++d) { g[GetRow(k, LX) + d] = lin(d); g[GetRow(k, AX) + d] = ang(d); }
}
// :68-73
void BaseMotionConstraint::UpdateBoundsAtInstance(double, int k, VecBound& bounds) const {
  for (int dim = 0; dim < static_cast<int>(node_bounds_.size()); ++dim) bounds.at(GetRow(k, dim)) = node_bounds_.at(dim);
}
// :75-85
void BaseMotionConstraint::UpdateJacobianAtInstance(double t, int k, const std::string& var_set, SpMat& jac) const {
  if (var_set == id::base_ang_nodes) jac.setMiddleRows(GetRow(k, AX), base_angular_->GetJacobianWrtNodes(t, kPos));
  if (var_set == id::base_lin_nodes) jac.setMiddleRows(GetRow(k, LX), base_linear_->GetJacobianWrtNodes(t, kPos));
}


// ---- constraints/ee_linear_constraint.cc:5-48
EELinearConstraint::EELinearConstraint(const std::vector<Term>& terms, const std::vector<std::shared_ptr<NodeSpline>>& splines,
                                       const std::vector<std::string>& var_names, Dx deriv, double tolerance, double T, double dt)
    : TimeDiscretizationConstraint(T, dt, "ee-linear"), terms_(terms), splines_(splines), var_names_(var_names), deriv_(deriv),
      tolerance_(tolerance) {
  SetRows(GetNumberOfNodes());
}
// :20-29
void EELinearConstraint::UpdateConstraintAtInstance(double t, int k, Vec& g) const {
  double val = 0.0;
  for (const auto& term : terms_) {
    auto pt = splines_.at(term.ee)->GetPoint(t);
    double v = (deriv_ == kPos) ? pt.p(term.dim) : pt.v(term.dim);
    val += term.coeff * v;
  }
  g[k] = val;
}
// :31-34
void EELinearConstraint::UpdateBoundsAtInstance(double, int k, VecBound& bounds) const { bounds.at(k) = Bounds(-tolerance_, tolerance_); }
// :36-48; jac.row(k) += coeff * jac_nodes.row(dim): Eigen's sparse row sum stores the union of both patterns
void EELinearConstraint::UpdateJacobianAtInstance(double t, int k, const std::string& var_set, SpMat& jac) const {
  for (const auto& term : terms_)
    if (var_set == var_names_.at(term.ee)) {
      SpMat jac_nodes = splines_.at(term.ee)->GetJacobianWrtNodes(t, deriv_);
      jac.setRow(k, jac.row(k) + term.coeff * jac_nodes.row(term.dim));
    }
}

}  // namespace orc
