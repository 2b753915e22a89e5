// ORACLE — TEST INFRASTRUCTURE ONLY (see towr_oracle.h).  The time-discretised cost terms of NlpFormulation::GetCost
// (nlp_formulation.cc:630-645): EnergyCost (costs/energy_cost.cc:38-163) and AngularMomentumCost
// (costs/angular_momentum_cost.cc:39-208).  Both sample the splines at t = 0, dt, 2 dt, ... <= T + 1e-9 and sum a
// squared norm weighted by weight * dt; the gradient is filled per variable set as ifopt::CostTerm::FillJacobianBlock
// does (a 1 x n block per set).
#include <cmath>
#include <numeric>

#include "constraints.h"

namespace orc {

// EnergyCost::GetSampleTimes (energy_cost.cc:52-66) == AngularMomentumCost::GetSampleTimes (angular_momentum_cost.cc:51-65)
std::vector<double> CostSampleTimes(const SplineHolder& s, double dt) {
  const Vec d = s.base_linear_->GetPolyDurations();
  const double T = std::accumulate(d.begin(), d.end(), 0.0);   // Spline::GetTotalTime, helpers/spline.cc:119-123
  std::vector<double> ts;
  if (dt <= 0.0) {
    ts.push_back(0.0);
    ts.push_back(T);
    return ts;
  }
  for (double t = 0.0; t <= T + 1e-9; t += dt) ts.push_back(t);
  return ts;
}

// EnergyCost::GetCost (energy_cost.cc:68-90)
double EnergyCostValue(const SplineHolder& s, double weight, double torque_weight, double dt) {
  if (weight <= 0.0) return 0.0;
  double cost = 0.0;
  const int n_ee = static_cast<int>(s.ee_force_.size());
  for (double t : CostSampleTimes(s, dt)) {
    double inst = 0.0;
    for (int ee = 0; ee < n_ee; ++ee) {
      const V3 f = s.ee_force_.at(ee)->GetPoint(t).p;
      const V3 tau = s.ee_torque_.at(ee)->GetPoint(t).p;
      inst += squaredNorm(f) + torque_weight * squaredNorm(tau);
    }
    cost += weight * inst * (dt > 0.0 ? dt : 1.0);
  }
  return cost;
}

static void AddRows(const SpMat& J, const double m[3], double* block) {
  for (int r = 0; r < 3; ++r)
    for (const Entry& en : J.r[r]) block[en.first] += m[r] * en.second;   // jac.coeffRef(0, it.col()) += m(r) * it.value()
}

// EnergyCost::FillJacobianBlock (energy_cost.cc:92-161); block: the set's columns, accumulated into
void EnergyCostFillBlock(const SplineHolder& s, double weight, double torque_weight, double dt, const std::string& var_set,
                         double* block) {
  if (weight <= 0.0) return;
  const double wdt = weight * (dt > 0.0 ? dt : 1.0);
  const int n_ee = static_cast<int>(s.ee_force_.size());
  for (double t : CostSampleTimes(s, dt)) {
    for (int ee = 0; ee < n_ee; ++ee) {
      const V3 f = s.ee_force_.at(ee)->GetPoint(t).p;
      const V3 tau = s.ee_torque_.at(ee)->GetPoint(t).p;
      const double mf[3] = {(2.0 * wdt) * f(0), (2.0 * wdt) * f(1), (2.0 * wdt) * f(2)};
      const double mt[3] = {(2.0 * wdt * torque_weight) * tau(0), (2.0 * wdt * torque_weight) * tau(1),
                            (2.0 * wdt * torque_weight) * tau(2)};
      if (var_set == "ee-force_" + std::to_string(ee)) AddRows(s.ee_force_.at(ee)->GetJacobianWrtNodes(t, kPos), mf, block);
      if (var_set == "ee-torque_" + std::to_string(ee) && torque_weight != 0.0)
        AddRows(s.ee_torque_.at(ee)->GetJacobianWrtNodes(t, kPos), mt, block);
      if (var_set == "ee-schedule" + std::to_string(ee)) {     // only exists when the durations are optimised
        AddRows(s.ee_force_.at(ee)->GetJacobianOfPosWrtDurations(t), mf, block);
        if (torque_weight != 0.0) AddRows(s.ee_torque_.at(ee)->GetJacobianOfPosWrtDurations(t), mt, block);
      }
    }
  }
}

// EEBasePosCost::GetCost (ee_base_pos_cost.cc:67-96): swing samples only, no dt factor
double EEBasePosCostValue(const SplineHolder& s, int ee, const V3& p_ref_B, double weight, double dt) {
  if (weight <= 0.0) return 0.0;
  double cost = 0.0;
  const EulerConverter& base_ang = *s.angular_converter_;
  for (double t : CostSampleTimes(s, dt)) {
    if (s.phase_durations_.at(ee)->IsContactPhase(t)) continue;
    const V3 p_base = s.base_linear_->GetPoint(t).p;
    const V3 p_ee = s.ee_motion_.at(ee)->GetPoint(t).p;
    const V3 r_W = p_ee - p_base;
    const M3 b_R_w = toDense(base_ang.GetRotationMatrixBaseToWorld(t)).transpose();
    const V3 p_B = b_R_w * r_W;
    const V3 e = p_B - p_ref_B;
    cost += weight * squaredNorm(e);
  }
  return cost;
}

// EEBasePosCost::FillJacobianBlock (ee_base_pos_cost.cc:98-160); the ee-schedule block is left out there too (:154-157)
void EEBasePosCostFillBlock(const SplineHolder& s, int ee, const V3& p_ref_B, double weight, double dt,
                            const std::string& var_set, double* block) {
  if (weight <= 0.0) return;
  const EulerConverter& base_ang = *s.angular_converter_;
  for (double t : CostSampleTimes(s, dt)) {
    if (s.phase_durations_.at(ee)->IsContactPhase(t)) continue;
    const V3 p_base = s.base_linear_->GetPoint(t).p;
    const V3 p_ee = s.ee_motion_.at(ee)->GetPoint(t).p;
    const V3 r_W = p_ee - p_base;
    const M3 b_R_w = toDense(base_ang.GetRotationMatrixBaseToWorld(t)).transpose();
    const V3 p_B = b_R_w * r_W;
    const V3 e = p_B - p_ref_B;
    const double m[3] = {(2.0 * weight) * e(0), (2.0 * weight) * e(1), (2.0 * weight) * e(2)};
    double mW[3], mWn[3];
    for (int c = 0; c < 3; ++c) {
      mW[c] = m[0] * b_R_w.m[0][c] + m[1] * b_R_w.m[1][c] + m[2] * b_R_w.m[2][c];            // m * b_R_w
      mWn[c] = (-m[0]) * b_R_w.m[0][c] + (-m[1]) * b_R_w.m[1][c] + (-m[2]) * b_R_w.m[2][c];   // -m * b_R_w
    }
    if (var_set == "ee-motion_" + std::to_string(ee)) AddRows(s.ee_motion_.at(ee)->GetJacobianWrtNodes(t, kPos), mW, block);
    if (var_set == "base-lin") AddRows(s.base_linear_->GetJacobianWrtNodes(t, kPos), mWn, block);
    if (var_set == "base-ang") AddRows(base_ang.DerivOfRotVecMult(t, r_W, true), m, block);
  }
}

static M3 Mul(const M3& A, const M3& B) {
  M3 C;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) C.m[i][j] = A.m[i][0] * B.m[0][j] + A.m[i][1] * B.m[1][j] + A.m[i][2] * B.m[2][j];
  return C;
}

// AngularMomentumCost::GetCost (angular_momentum_cost.cc:67-101); the dynamic model is always the SRBD here
double AngMomCostValue(const SplineHolder& s, const SingleRigidBodyDynamics& model, double weight, double dt) {
  if (weight <= 0.0) return 0.0;
  const EulerConverter& base_ang = *s.angular_converter_;
  const double wdt = weight * (dt > 0.0 ? dt : 1.0);
  const M3 I_b = model.GetInertiaB();
  double cost = 0.0;
  for (double t : CostSampleTimes(s, dt)) {
    const M3 R = toDense(base_ang.GetRotationMatrixBaseToWorld(t));
    const V3 omega = base_ang.GetAngularVelocityInWorld(t);
    const M3 I_w = Mul(Mul(R, I_b), R.transpose());
    const V3 v = I_w * omega;
    cost += wdt * squaredNorm(v);
  }
  return cost;
}

// AngularMomentumCost::FillJacobianBlock (angular_momentum_cost.cc:103-206), the SRBD branch
void AngMomCostFillBlock(const SplineHolder& s, const SingleRigidBodyDynamics& model, double weight, double dt,
                         const std::string& var_set, double* block) {
  if (weight <= 0.0) return;
  if (var_set != "base-ang") return;
  const EulerConverter& base_ang = *s.angular_converter_;
  const double wdt = weight * (dt > 0.0 ? dt : 1.0);
  const M3 I_b = model.GetInertiaB();
  const int n = s.base_angular_->GetNodeVariablesCount();
  for (double t : CostSampleTimes(s, dt)) {
    const M3 R = toDense(base_ang.GetRotationMatrixBaseToWorld(t));
    const M3 R_T = R.transpose();
    const V3 omega = base_ang.GetAngularVelocityInWorld(t);
    const M3 I_w = Mul(Mul(R, I_b), R.transpose());
    const V3 v = I_w * omega;
    const double m[3] = {(2.0 * wdt) * v(0), (2.0 * wdt) * v(1), (2.0 * wdt) * v(2)};
    const V3 u = R_T * omega;
    const V3 v2 = I_b * u;
    const SpMat J_R_v2 = base_ang.DerivOfRotVecMult(t, v2, false);
    const SpMat J_Rt_omega = base_ang.DerivOfRotVecMult(t, omega, true);
    const SpMat J_omega = base_ang.GetDerivOfAngVelWrtNodes(t);
    std::vector<double> term_a(3 * static_cast<size_t>(n), 0.0), term_b(term_a), term_u(term_a), J2(term_a);
    for (int r = 0; r < 3; ++r) for (const Entry& en : J_Rt_omega.r[r]) term_a[r * n + en.first] += en.second;
    for (int r = 0; r < 3; ++r) for (const Entry& en : J_omega.r[r]) term_b[r * n + en.first] += en.second;
    for (int c = 0; c < n; ++c) {
      double b[3], a[3], iu[3];
      for (int r = 0; r < 3; ++r) b[r] = R_T.m[r][0] * term_b[c] + R_T.m[r][1] * term_b[n + c] + R_T.m[r][2] * term_b[2 * n + c];
      for (int r = 0; r < 3; ++r) a[r] = term_a[r * n + c] + b[r];                                   // term_u
      for (int r = 0; r < 3; ++r) iu[r] = I_b.m[r][0] * a[0] + I_b.m[r][1] * a[1] + I_b.m[r][2] * a[2];
      for (int r = 0; r < 3; ++r) J2[r * n + c] = R.m[r][0] * iu[0] + R.m[r][1] * iu[1] + R.m[r][2] * iu[2];
    }
    AddRows(J_R_v2, m, block);
    for (int c = 0; c < n; ++c) {
      const double vcol = m[0] * J2[c] + m[1] * J2[n + c] + m[2] * J2[2 * n + c];    // grad2 = m * J2
      if (vcol != 0.0) block[c] += vcol;
    }
  }
}

}  // namespace orc
