// ORACLE — TEST INFRASTRUCTURE ONLY (see towr_oracle.h header).
// Variables, cubic-Hermite splines and the Euler converter.
#include <cassert>
#include <cmath>
#include <numeric>
#include <stdexcept>

#include "towr_oracle.h"

namespace orc {

// ---- test switch: which power function the run-time (PhaseSpline) polynomials use
static int g_pow_mode = 0;
void SetPowMode(int mode) { g_pow_mode = mode; }
int GetPowMode() { return g_pow_mode; }

double ExactPow(double x, int n) {
  switch (n) {
    case 0: return 1.0;
    case 1: return x;
    case 2: return x * x;
    case 3: {                      // x*x with its rounding error carried by FMA
      const double p = x * x;
      const double e = std::fma(x, x, -p);
      return std::fma(p, x, e * x);
    }
    case 4: {
      const double p = x * x;
      const double e = std::fma(x, x, -p);
      const double hi = p * p;
      return hi + (std::fma(p, p, -hi) + 2.0 * p * e);
    }
  }
  return std::pow(x, n);
}

double CubicHermitePolynomial::Pw(double x, int n) const {
  return (runtime_ && g_pow_mode == 1) ? ExactPow(x, n) : std::pow(x, n);
}

// ============================================================ polynomial.cc
// polynomial.cc:97-104
void CubicHermitePolynomial::UpdateCoeff() {
  coeff_[0] = n0_.p;
  coeff_[1] = n0_.v;
  coeff_[2] = -(3.0 * (n0_.p - n1_.p) + T_ * (2.0 * n0_.v + n1_.v)) / Pw(T_, 2);
  coeff_[3] = (2.0 * (n0_.p - n1_.p) + T_ * (n0_.v + n1_.v)) / Pw(T_, 3);
}

// polynomial.cc:63-72  d(t^c-derivative)/d(coeff c)
template <class PowFn>
static double DerivWrtCoeff(double t, Dx d, int c, PowFn pw) {
  switch (d) {
    case kPos: return pw(t, c);
    case kVel: return c >= 1 ? c * pw(t, c - 1) : 0.0;
    case kAcc: return c >= 2 ? c * (c - 1) * pw(t, c - 2) : 0.0;
  }
  return 0.0;
}

// polynomial.cc:47-61  out.at(d) += GetDerivativeWrtCoeff(t,d,c)*coeff_.at(c), c = A..D, from zero
State3 CubicHermitePolynomial::GetPoint(double t_local) const {
  State3 out;
  for (Dx d : {kPos, kVel, kAcc}) {
    V3 acc;
    for (int c = 0; c < 4; ++c) acc += DerivWrtCoeff(t_local, d, c, [this](double x, int n) { return Pw(x, n); }) * coeff_[c];
    (d == kPos ? out.p : d == kVel ? out.v : out.a) = acc;
  }
  return out;
}

// polynomial.cc:106-180
double CubicHermitePolynomial::GetDerivativeWrtStartNode(Dx dfdt, Dx node_value, double t) const {
  double t2 = Pw(t, 2), t3 = Pw(t, 3);
  double T = T_, T2 = Pw(T_, 2), T3 = Pw(T_, 3);
  switch (dfdt) {
    case kPos: return node_value == kPos ? (2*t3)/T3 - (3*t2)/T2 + 1 : t - (2*t2)/T + t3/T2;
    case kVel: return node_value == kPos ? (6*t2)/T3 - (6*t)/T2 : (3*t2)/T2 - (4*t)/T + 1;
    case kAcc: return node_value == kPos ? (12*t)/T3 - 6/T2 : (6*t)/T2 - 4/T;
  }
  return 0.0;
}

// polynomial.cc:122-234
double CubicHermitePolynomial::GetDerivativeWrtEndNode(Dx dfdt, Dx node_value, double t) const {
  double t2 = Pw(t, 2), t3 = Pw(t, 3);
  double T = T_, T2 = Pw(T_, 2), T3 = Pw(T_, 3);
  switch (dfdt) {
    case kPos: return node_value == kPos ? (3*t2)/T2 - (2*t3)/T3 : t3/T2 - t2/T;
    case kVel: return node_value == kPos ? (6*t)/T2 - (6*t2)/T3 : (3*t2)/T2 - (2*t)/T;
    case kAcc: return node_value == kPos ? 6/T2 - (12*t)/T3 : (6*t)/T2 - 2/T;
  }
  return 0.0;
}

// polynomial.cc:236-257
V3 CubicHermitePolynomial::GetDerivativeOfPosWrtDuration(double t) const {
  V3 x0 = n0_.p, x1 = n1_.p, v0 = n0_.v, v1 = n1_.v;
  double t2 = Pw(t, 2), t3 = Pw(t, 3);
  double T = T_, T2 = Pw(T_, 2), T3 = Pw(T_, 3), T4 = Pw(T_, 4);
  return (t3 * (v0 + v1)) / T3
       - (t2 * (2.0 * v0 + v1)) / T2
       - ((3 * t3) * (2.0 * x0 - 2.0 * x1 + T * v0 + T * v1)) / T4
       + ((2 * t2) * (3.0 * x0 - 3.0 * x1 + (2 * T) * v0 + T * v1)) / T3;
}

// ======================================================== nodes_variables.cc
// nodes_variables.cc:44-54
int NodesVariables::GetOptIndex(const NodeValueInfo& nvi_des) const {
  for (int idx = 0; idx < GetRows(); ++idx)
    for (const auto& nvi : GetNodeValuesInfo(idx))
      if (nvi == nvi_des) return idx;
  return NodeValueNotOptimized;
}

// nodes_variables.cc:56-66
Vec NodesVariables::GetValues() const {
  Vec x(GetRows());
  for (int idx = 0; idx < GetRows(); ++idx)
    for (const auto& nvi : GetNodeValuesInfo(idx)) x[idx] = nodes_.at(nvi.id).at(nvi.deriv)(nvi.dim);
  return x;
}

// nodes_variables.cc:68-76
void NodesVariables::SetVariables(const Vec& x) {
  for (int idx = 0; idx < static_cast<int>(x.size()); ++idx)
    for (const auto& nvi : GetNodeValuesInfo(idx)) nodes_.at(nvi.id).at(nvi.deriv)(nvi.dim) = x[idx];
  UpdateObservers();
}

void NodesVariables::UpdateObservers() const {
  for (auto* o : observers_) o->UpdateNodes();
}

// nodes_variables.cc:130-155
void NodesVariables::SetByLinearInterpolation(const V3& initial_val, const V3& final_val, double t_total) {
  V3 dp = final_val - initial_val;
  V3 average_velocity = dp / t_total;
  int num_nodes = static_cast<int>(nodes_.size());
  for (int idx = 0; idx < GetRows(); ++idx)
    for (const auto& nvi : GetNodeValuesInfo(idx)) {
      if (nvi.deriv == kPos) {
        V3 pos = initial_val + (nvi.id / static_cast<double>(num_nodes - 1)) * dp;
        nodes_.at(nvi.id).p(nvi.dim) = pos(nvi.dim);
      }
      if (nvi.deriv == kVel) nodes_.at(nvi.id).v(nvi.dim) = average_velocity(nvi.dim);
    }
}

// nodes_variables.cc:157-217
void NodesVariables::SetByLinearInterpolationRelativeToBase(const V3& ee_initial_W, const V3& ee_final_W,
                                                            const V3& base_pos_initial_W, const V3& base_pos_final_W,
                                                            const V3& base_rpy_initial_W, const V3& base_rpy_final_W,
                                                            double t_total) {
  if (t_total <= 0.0) throw std::runtime_error("SetByLinearInterpolationRelativeToBase requires t_total > 0.");
  const int num_nodes = static_cast<int>(nodes_.size());
  if (num_nodes < 2) return;

  const M3 w_R_b0 = EulerConverter::GetRotationMatrixBaseToWorldDense(base_rpy_initial_W);
  const M3 w_R_bT = EulerConverter::GetRotationMatrixBaseToWorldDense(base_rpy_final_W);
  const V3 r0_B = w_R_b0.transpose() * (ee_initial_W - base_pos_initial_W);
  const V3 rT_B = w_R_bT.transpose() * (ee_final_W - base_pos_final_W);
  const V3 dp_B = rT_B - r0_B;
  const V3 avg_vel_B = dp_B / t_total;
  const V3 base_avg_vel_W = (base_pos_final_W - base_pos_initial_W) / t_total;

  for (int idx = 0; idx < GetRows(); ++idx)
    for (const auto& nvi : GetNodeValuesInfo(idx)) {
      const double alpha = nvi.id / static_cast<double>(num_nodes - 1);
      const V3 base_pos_W = (1.0 - alpha) * base_pos_initial_W + alpha * base_pos_final_W;
      const V3 base_rpy_W = (1.0 - alpha) * base_rpy_initial_W + alpha * base_rpy_final_W;
      const M3 w_R_b = EulerConverter::GetRotationMatrixBaseToWorldDense(base_rpy_W);
      const V3 r_B = r0_B + alpha * dp_B;
      if (nvi.deriv == kPos) {
        const V3 ee_pos_W = base_pos_W + w_R_b * r_B;
        nodes_.at(nvi.id).p(nvi.dim) = ee_pos_W(nvi.dim);
      }
      if (nvi.deriv == kVel) {
        const V3 ee_vel_W = base_avg_vel_W + w_R_b * avg_vel_B;
        nodes_.at(nvi.id).v(nvi.dim) = ee_vel_W(nvi.dim);
      }
    }
  SetVariables(GetValues());
}

// nodes_variables.cc:219-263
void NodesVariables::AddBound(const NodeValueInfo& nvi_des, double lower, double upper) {
  for (int idx = 0; idx < GetRows(); ++idx)
    for (const auto& nvi : GetNodeValuesInfo(idx))
      if (nvi == nvi_des) bounds_.at(idx) = Bounds(lower, upper);
}
void NodesVariables::AddBounds(int node_id, Dx deriv, const std::vector<int>& dims, const V3& val) {
  for (int dim : dims) AddBound(NodeValueInfo{node_id, deriv, dim}, val(dim), val(dim));
}
void NodesVariables::AddBounds(int node_id, Dx deriv, const std::vector<int>& dims, const V3& lower, const V3& upper) {
  for (int dim : dims) AddBound(NodeValueInfo{node_id, deriv, dim}, lower(dim), upper(dim));
}

// nodes_variables_all.cc:34-61
NodesVariablesAll::NodesVariablesAll(int n_nodes, const std::string& name) : NodesVariables(name) {
  int n_opt = n_nodes * 2 * k3D;
  nodes_.assign(n_nodes, Node());
  bounds_.assign(n_opt, NoBound());
  SetRows(n_opt);
}
std::vector<NodesVariables::NodeValueInfo> NodesVariablesAll::GetNodeValuesInfo(int idx) const {
  int per_node = 2 * k3D;
  int internal_id = idx % per_node;
  NodeValueInfo nvi;
  nvi.deriv = internal_id < k3D ? kPos : kVel;
  nvi.dim = internal_id % k3D;
  nvi.id = static_cast<int>(std::floor(idx / per_node));
  return {nvi};
}

// ============================================ nodes_variables_phase_based.cc
// :39-59 BuildPolyInfos, :61-73 ctor
NodesVariablesPhaseBased::NodesVariablesPhaseBased(Kind kind, int phase_count, bool in_contact_at_start,
                                                   const std::string& name, int n_polys_in_changing_phase)
    : NodesVariables(name) {
  // motion/orientation: contact phase is the constant one (:203-213,:352-364);
  // force/torque: contact phase is the changing one (:260-272,:306-318)
  bool phase_constant = (kind == Motion || kind == Ang) ? in_contact_at_start : !in_contact_at_start;
  for (int i = 0; i < phase_count; ++i) {
    if (phase_constant) poly_info_.push_back(PolyInfo{i, 0, 1, true});
    else
      for (int j = 0; j < n_polys_in_changing_phase; ++j) poly_info_.push_back(PolyInfo{i, j, n_polys_in_changing_phase, false});
    phase_constant = !phase_constant;
  }
  nodes_.assign(poly_info_.size() + 1, Node());

  if (kind == Motion) BuildMotionLikeMap(true);
  else if (kind == Ang) BuildMotionLikeMap(false);
  else BuildForceLikeMap();

  int n = static_cast<int>(index_to_nvi_.size());   // :194-199 SetNumberOfVariables
  bounds_.assign(n, NoBound());
  SetRows(n);
}

// :215-258 (pin_swing_vz = true: swing z-velocity fixed to 0) and :366-396 (false)
void NodesVariablesPhaseBased::BuildMotionLikeMap(bool pin_swing_vz) {
  int idx = 0;
  for (int node_id = 0; node_id < static_cast<int>(nodes_.size()); ++node_id) {
    if (!IsConstantNode(node_id)) {
      for (int dim = 0; dim < k3D; ++dim) {
        index_to_nvi_[idx++].push_back(NodeValueInfo{node_id, kPos, dim});
        if (pin_swing_vz && dim == Z) nodes_.at(node_id).v(Z) = 0.0;
        else index_to_nvi_[idx++].push_back(NodeValueInfo{node_id, kVel, dim});
      }
    } else {
      nodes_.at(node_id).v = V3();
      nodes_.at(node_id + 1).v = V3();
      for (int dim = 0; dim < k3D; ++dim) {
        index_to_nvi_[idx].push_back(NodeValueInfo{node_id, kPos, dim});
        index_to_nvi_[idx].push_back(NodeValueInfo{node_id + 1, kPos, dim});
        idx++;
      }
      node_id += 1;
    }
  }
}

// :274-304 and :320-350
void NodesVariablesPhaseBased::BuildForceLikeMap() {
  int idx = 0;
  for (int id = 0; id < static_cast<int>(nodes_.size()); ++id) {
    if (!IsConstantNode(id)) {
      for (int dim = 0; dim < k3D; ++dim) {
        index_to_nvi_[idx++].push_back(NodeValueInfo{id, kPos, dim});
        index_to_nvi_[idx++].push_back(NodeValueInfo{id, kVel, dim});
      }
    } else {
      nodes_.at(id) = Node();
      nodes_.at(id + 1) = Node();
      id += 1;
    }
  }
}

// :75-86
Vec NodesVariablesPhaseBased::ConvertPhaseToPolyDurations(const Vec& phase_durations) const {
  Vec poly_durations;
  for (int i = 0; i < GetPolynomialCount(); ++i) {
    const auto& info = poly_info_.at(i);
    poly_durations.push_back(phase_durations.at(info.phase) / info.n_polys_in_phase);
  }
  return poly_durations;
}
// :88-99
double NodesVariablesPhaseBased::GetDerivativeOfPolyDurationWrtPhaseDuration(int poly_id) const {
  return 1. / poly_info_.at(poly_id).n_polys_in_phase;
}
int NodesVariablesPhaseBased::GetNumberOfPrevPolynomialsInPhase(int poly_id) const { return poly_info_.at(poly_id).poly_in_phase; }
// :101-113
bool NodesVariablesPhaseBased::IsConstantNode(int node_id) const {
  bool is_constant = false;
  for (int poly_id : GetAdjacentPolyIds(node_id))
    if (IsInConstantPhase(poly_id)) is_constant = true;
  return is_constant;
}
// :121-131
std::vector<int> NodesVariablesPhaseBased::GetIndicesOfNonConstantNodes() const {
  std::vector<int> ids;
  for (int id = 0; id < static_cast<int>(nodes_.size()); ++id)
    if (!IsConstantNode(id)) ids.push_back(id);
  return ids;
}
// :133-140
int NodesVariablesPhaseBased::GetPhase(int node_id) const {
  int poly_id = GetAdjacentPolyIds(node_id).front();
  return poly_info_.at(poly_id).phase;
}
// :142-151
int NodesVariablesPhaseBased::GetPolyIDAtStartOfPhase(int phase) const {
  for (int i = 0; i < static_cast<int>(poly_info_.size()); ++i)
    if (poly_info_.at(i).phase == phase) return i;
  throw std::runtime_error("phase not found in NodesVariablesPhaseBased::GetPolyIDAtStartOfPhase");
}
// :167-183
std::vector<int> NodesVariablesPhaseBased::GetAdjacentPolyIds(int node_id) const {
  int last_node_id = static_cast<int>(nodes_.size()) - 1;
  if (node_id == 0) return {0};
  if (node_id == last_node_id) return {last_node_id - 1};
  return {node_id - 1, node_id};
}

// ========================================================= phase_durations.cc
// :40-52
PhaseDurations::PhaseDurations(int ee, const Vec& timings, bool first_phase_in_contact, double min_d, double max_d)
    : VariableSet(static_cast<int>(timings.size()) - 1, "ee-schedule" + std::to_string(ee)) {
  durations_ = timings;
  t_total_ = std::accumulate(timings.begin(), timings.end(), 0.0);
  phase_duration_bounds_ = Bounds(min_d, max_d);
  initial_contact_state_ = first_phase_in_contact;
}
// :67-76
Vec PhaseDurations::GetValues() const {
  Vec x(GetRows());
  for (int i = 0; i < GetRows(); ++i) x[i] = durations_.at(i);
  return x;
}
// :78-100   (assert(t_total_>x.sum()) is a no-op in Release)
void PhaseDurations::SetVariables(const Vec& x) {
  for (int i = 0; i < GetRows(); ++i) durations_.at(i) = x[i];
  double sum = 0.0;
  for (int i = 0; i < GetRows(); ++i) sum += x[i];
  durations_.back() = t_total_ - sum;
  for (auto* s : observers_) s->UpdatePolynomialDurations();
}
// :102-111
VecBound PhaseDurations::GetBounds() const { return VecBound(GetRows(), phase_duration_bounds_); }
// :119-124
bool PhaseDurations::IsContactPhase(double t) const {
  int phase_id = NodeSpline::GetSegmentID(t, durations_);
  return phase_id % 2 == 0 ? initial_contact_state_ : !initial_contact_state_;
}
// :126-154  dense 3 x (P-1), converted with sparseView(1.0,-1.0): every coefficient stored
SpMat PhaseDurations::GetJacobianOfPos(int current_phase, const V3& dx_dT, const V3& xd) const {
  int n = GetRows();
  std::vector<V3> col(n);   // zero-initialised dense columns
  bool in_last_phase = (current_phase == static_cast<int>(durations_.size()) - 1);
  if (!in_last_phase) col[current_phase] = dx_dT;
  for (int phase = 0; phase < current_phase; ++phase) {
    col[phase] = -1.0 * xd;
    if (in_last_phase) col[phase] = col[phase] - dx_dT;
  }
  SpMat jac(k3D, n);
  for (int d = 0; d < k3D; ++d)
    for (int c = 0; c < n; ++c) jac.r[d].emplace_back(c, col[c](d));
  return jac;
}

// ================================================== spline.cc / node_spline.cc
// spline.cc:37-46, node_spline.cc:36-43
NodeSpline::NodeSpline(NodesVariables* nodes, const Vec& poly_durations) : nodes_(nodes) {
  polys_.assign(poly_durations.size(), CubicHermitePolynomial());
  for (size_t i = 0; i < polys_.size(); ++i) polys_[i].SetDuration(poly_durations[i]);
  UpdatePolynomialCoeff();
  nodes->AddObserver(this);            // helpers/nodes_observer.cc:36-43
  UpdateNodes();
  jac_wrt_nodes_structure_ = SpMat(k3D, nodes->GetRows());
}

// spline.cc:48-66
int NodeSpline::GetSegmentID(double t_global, const Vec& durations) {
  double eps = 1e-10;
  double t = 0;
  int i = 0;
  for (double d : durations) {
    t += d;
    if (t >= t_global - eps) return i;
    i++;
  }
  // reference: assert(false) — unreachable for t <= T; mirror Release behaviour (falls off the end)
  return static_cast<int>(durations.size()) - 1;
}
// spline.cc:68-78
std::pair<int, double> NodeSpline::GetLocalTime(double t_global, const Vec& durations) const {
  int id = GetSegmentID(t_global, durations);
  double t_local = t_global;
  for (int i = 0; i < id; i++) t_local -= durations.at(i);
  return {id, t_local};
}
// spline.cc:80-93
State3 NodeSpline::GetPoint(double t_global) const {
  auto lt = GetLocalTime(t_global, GetPolyDurations());
  return GetPoint(lt.first, lt.second);
}
// spline.cc:108-116
Vec NodeSpline::GetPolyDurations() const {
  Vec d;
  for (const auto& p : polys_) d.push_back(p.GetDuration());
  return d;
}
// node_spline.cc:45-54
void NodeSpline::UpdateNodes() {
  const auto& nodes = nodes_->GetNodes();
  for (size_t i = 0; i < polys_.size(); ++i)
    polys_[i].SetNodes(nodes.at(NodesVariables::GetNodeId(i, NodesVariables::Start)),
                       nodes.at(NodesVariables::GetNodeId(i, NodesVariables::End)));
  UpdatePolynomialCoeff();
}
// node_spline.cc:62-82
SpMat NodeSpline::GetJacobianWrtNodes(double t_global, Dx dxdt) const {
  auto lt = GetLocalTime(t_global, GetPolyDurations());
  return GetJacobianWrtNodes(lt.first, lt.second, dxdt);
}
SpMat NodeSpline::GetJacobianWrtNodes(int id, double t_local, Dx dxdt) const {
  SpMat jac = jac_wrt_nodes_structure_;
  FillJacobianWrtNodes(id, t_local, dxdt, jac, false);
  return jac;
}
// node_spline.cc:84-112
void NodeSpline::FillJacobianWrtNodes(int poly_id, double t_local, Dx dxdt, SpMat& jac, bool fill_with_zeros) const {
  for (int idx = 0; idx < jac.cols; ++idx)
    for (const auto& nvi : nodes_->GetNodeValuesInfo(idx))
      for (auto side : {NodesVariables::Start, NodesVariables::End}) {
        int node = NodesVariables::GetNodeId(poly_id, side);
        if (node == nvi.id) {
          double val = 0.0;
          if (side == NodesVariables::Start) val = polys_.at(poly_id).GetDerivativeWrtStartNode(dxdt, nvi.deriv, t_local);
          else val = polys_.at(poly_id).GetDerivativeWrtEndNode(dxdt, nvi.deriv, t_local);
          if (fill_with_zeros) val = 0.0;
          jac.coeffRef(nvi.dim, idx) += val;
        }
      }
}

// ============================================================ phase_spline.cc
// :35-52
PhaseSpline::PhaseSpline(const std::shared_ptr<NodesVariablesPhaseBased>& nodes, PhaseDurations* durations)
    : NodeSpline(nodes.get(), nodes->ConvertPhaseToPolyDurations(durations->GetPhaseDurations())),
      phase_nodes_(nodes), phase_durations_(durations) {
  durations->AddObserver(this);        // helpers/phase_durations_observer.cc:37-44
  for (auto& p : polys_) p.MarkRuntimeDurations();   // test switch only (SetPowMode), no effect in the default mode
  UpdatePolynomialDurations();
  // every global time can fall into every polynomial: pre-fill all columns
  for (int i = 0; i < nodes->GetPolynomialCount(); ++i) FillJacobianWrtNodes(i, 0.0, kPos, jac_wrt_nodes_structure_, true);
}
// :54-65
void PhaseSpline::UpdatePolynomialDurations() {
  auto poly_durations = phase_nodes_->ConvertPhaseToPolyDurations(phase_durations_->GetPhaseDurations());
  for (size_t i = 0; i < polys_.size(); ++i) polys_[i].SetDuration(poly_durations.at(i));
  UpdatePolynomialCoeff();
}
// :67-75
SpMat PhaseSpline::GetJacobianOfPosWrtDurations(double t_global) const {
  V3 dx_dT = GetDerivativeOfPosWrtPhaseDuration(t_global);
  V3 xd = GetPoint(t_global).v;
  int current_phase = GetSegmentID(t_global, phase_durations_->GetPhaseDurations());
  return phase_durations_->GetJacobianOfPos(current_phase, dx_dT, xd);
}
// :77-93
V3 PhaseSpline::GetDerivativeOfPosWrtPhaseDuration(double t_global) const {
  auto lt = GetLocalTime(t_global, GetPolyDurations());
  int poly_id = lt.first;
  double t_local = lt.second;
  V3 vel = GetPoint(t_global).v;
  V3 dxdT = polys_.at(poly_id).GetDerivativeOfPosWrtDuration(t_local);
  double inner_derivative = phase_nodes_->GetDerivativeOfPolyDurationWrtPhaseDuration(poly_id);
  double prev_polys_in_phase = phase_nodes_->GetNumberOfPrevPolynomialsInPhase(poly_id);
  return inner_derivative * (dxdT - prev_polys_in_phase * vel);
}

// ========================================================= euler_converter.cc
// :39-43
EulerConverter::EulerConverter(const std::shared_ptr<NodeSpline>& euler) : euler_(euler) {
  jac_wrt_nodes_structure_ = SpMat(k3D, euler->GetNodeVariablesCount());
}
// :58-70
V3 EulerConverter::GetAngularVelocityInWorld(double t) const {
  State3 ori = euler_->GetPoint(t);
  return GetM(ori.p) * ori.v;
}
// :72-83
V3 EulerConverter::GetAngularAccelerationInWorld(double t) const {
  State3 ori = euler_->GetPoint(t);
  return GetMdot(ori.p, ori.v) * ori.v + GetM(ori.p) * ori.a;
}
// :85-102
SpMat EulerConverter::GetDerivOfAngVelWrtNodes(double t) const {
  SpMat jac = jac_wrt_nodes_structure_;
  State3 ori = euler_->GetPoint(t);
  SpVec vel = sparseViewAll(ori.v);
  SpMat dVel_du = euler_->GetJacobianWrtNodes(t, kVel);
  for (Dim3D dim : {X, Y, Z}) {
    SpMat dM_du = GetDerivMwrtNodes(t, dim);
    jac.setRow(dim, vel * dM_du + GetM(ori.p).row(dim) * dVel_du);
  }
  return jac;
}
// :104-131
SpMat EulerConverter::GetDerivOfAngAccWrtNodes(double t) const {
  SpMat jac = jac_wrt_nodes_structure_;
  State3 ori = euler_->GetPoint(t);
  SpVec vel = sparseViewAll(ori.v);
  SpVec acc = sparseViewAll(ori.a);
  SpMat dVel_du = euler_->GetJacobianWrtNodes(t, kVel);
  SpMat dAcc_du = euler_->GetJacobianWrtNodes(t, kAcc);
  for (Dim3D dim : {X, Y, Z}) {
    SpMat dMdot_du = GetDerivMdotwrtNodes(t, dim);
    SpMat dM_du = GetDerivMwrtNodes(t, dim);
    jac.setRow(dim, vel * dMdot_du
                    + GetMdot(ori.p, ori.v).row(dim) * dVel_du
                    + acc * dM_du
                    + GetM(ori.p).row(dim) * dAcc_du);
  }
  return jac;
}
// :133-148   stored entries: (0,X),(0,Y),(1,X),(1,Y),(2,X),(2,Z)
SpMat EulerConverter::GetM(const V3& xyz) {
  double z = xyz(Z), y = xyz(Y);
  SpMat M(k3D, k3D);
  M.coeffRef(0, Y) = -sin(z); M.coeffRef(0, X) = cos(y) * cos(z);
  M.coeffRef(1, Y) = cos(z);  M.coeffRef(1, X) = cos(y) * sin(z);
  M.coeffRef(2, Z) = 1.0;     M.coeffRef(2, X) = -sin(y);
  return M;
}
// :150-166
SpMat EulerConverter::GetMdot(const V3& xyz, const V3& xyz_d) {
  double z = xyz(Z), zd = xyz_d(Z), y = xyz(Y), yd = xyz_d(Y);
  SpMat Mdot(k3D, k3D);
  Mdot.coeffRef(0, Y) = -cos(z) * zd; Mdot.coeffRef(0, X) = -cos(z) * sin(y) * yd - cos(y) * sin(z) * zd;
  Mdot.coeffRef(1, Y) = -sin(z) * zd; Mdot.coeffRef(1, X) = cos(y) * cos(z) * zd - sin(y) * sin(z) * yd;
  Mdot.coeffRef(2, X) = -cos(y) * yd;
  return Mdot;
}
// :168-198
SpMat EulerConverter::GetDerivMwrtNodes(double t, Dim3D ang_acc_dim) const {
  State3 ori = euler_->GetPoint(t);
  double z = ori.p(Z), y = ori.p(Y);
  SpVec jac_z = GetJac(t, kPos, Z);
  SpVec jac_y = GetJac(t, kPos, Y);
  SpMat jac = jac_wrt_nodes_structure_;
  switch (ang_acc_dim) {
    case X:
      jac.setRow(Y, -cos(z) * jac_z);
      jac.setRow(X, -cos(z) * sin(y) * jac_y - cos(y) * sin(z) * jac_z);
      break;
    case Y:
      jac.setRow(Y, -sin(z) * jac_z);
      jac.setRow(X, cos(y) * cos(z) * jac_z - sin(y) * sin(z) * jac_y);
      break;
    case Z:
      jac.setRow(X, -cos(y) * jac_y);
      break;
  }
  return jac;
}
// :200-221
SpMat EulerConverter::GetRotationMatrixBaseToWorld(double t) const {
  State3 ori = euler_->GetPoint(t);
  return GetRotationMatrixBaseToWorld(ori.p);
}
M3 EulerConverter::GetRotationMatrixBaseToWorldDense(const V3& xyz) {
  double x = xyz(X), y = xyz(Y), z = xyz(Z);
  M3 M;
  M(0,0) = cos(y)*cos(z); M(0,1) = cos(z)*sin(x)*sin(y) - cos(x)*sin(z); M(0,2) = sin(x)*sin(z) + cos(x)*cos(z)*sin(y);
  M(1,0) = cos(y)*sin(z); M(1,1) = cos(x)*cos(z) + sin(x)*sin(y)*sin(z); M(1,2) = cos(x)*sin(y)*sin(z) - cos(z)*sin(x);
  M(2,0) = -sin(y);       M(2,1) = cos(y)*sin(x);                        M(2,2) = cos(x)*cos(y);
  return M;
}
SpMat EulerConverter::GetRotationMatrixBaseToWorld(const V3& xyz) {
  return sparseView(GetRotationMatrixBaseToWorldDense(xyz), /*keep_zeros=*/true);   // M.sparseView(1.0,-1.0)
}
// :223-239
SpMat EulerConverter::DerivOfRotVecMult(double t, const V3& v, bool inverse) const {
  JacRowMatrix Rd = GetDerivativeOfRotationMatrixWrtNodes(t);
  SpMat jac = jac_wrt_nodes_structure_;
  for (int row : {X, Y, Z})
    for (int col : {X, Y, Z}) {
      const SpVec& jac_row = inverse ? Rd.at(col).at(row) : Rd.at(row).at(col);
      jac.setRow(row, jac.row(row) + v(col) * jac_row);
    }
  return jac;
}
// :241-268
EulerConverter::JacRowMatrix EulerConverter::GetDerivativeOfRotationMatrixWrtNodes(double t) const {
  JacRowMatrix jac;
  State3 ori = euler_->GetPoint(t);
  double x = ori.p(X), y = ori.p(Y), z = ori.p(Z);
  SpVec jac_x = GetJac(t, kPos, X);
  SpVec jac_y = GetJac(t, kPos, Y);
  SpVec jac_z = GetJac(t, kPos, Z);

  jac[X][X] = -cos(z)*sin(y)*jac_y - cos(y)*sin(z)*jac_z;
  jac[X][Y] = sin(x)*sin(z)*jac_x - cos(x)*cos(z)*jac_z - sin(x)*sin(y)*sin(z)*jac_z + cos(x)*cos(z)*sin(y)*jac_x + cos(y)*cos(z)*sin(x)*jac_y;
  jac[X][Z] = cos(x)*sin(z)*jac_x + cos(z)*sin(x)*jac_z - cos(z)*sin(x)*sin(y)*jac_x - cos(x)*sin(y)*sin(z)*jac_z + cos(x)*cos(y)*cos(z)*jac_y;

  jac[Y][X] = cos(y)*cos(z)*jac_z - sin(y)*sin(z)*jac_y;
  jac[Y][Y] = cos(x)*sin(y)*sin(z)*jac_x - cos(x)*sin(z)*jac_z - cos(z)*sin(x)*jac_x + cos(y)*sin(x)*sin(z)*jac_y + cos(z)*sin(x)*sin(y)*jac_z;
  jac[Y][Z] = sin(x)*sin(z)*jac_z - cos(x)*cos(z)*jac_x - sin(x)*sin(y)*sin(z)*jac_x + cos(x)*cos(y)*sin(z)*jac_y + cos(x)*cos(z)*sin(y)*jac_z;

  jac[Z][X] = -cos(y)*jac_y;
  jac[Z][Y] = cos(x)*cos(y)*jac_x - sin(x)*sin(y)*jac_y;
  jac[Z][Z] = -cos(y)*sin(x)*jac_x - cos(x)*sin(y)*jac_y;
  return jac;
}
// :270-304
SpMat EulerConverter::GetDerivMdotwrtNodes(double t, Dim3D ang_acc_dim) const {
  State3 ori = euler_->GetPoint(t);
  double z = ori.p(Z), zd = ori.v(Z), y = ori.p(Y), yd = ori.v(Y);
  SpVec jac_z = GetJac(t, kPos, Z);
  SpVec jac_y = GetJac(t, kPos, Y);
  SpVec jac_zd = GetJac(t, kVel, Z);
  SpVec jac_yd = GetJac(t, kVel, Y);
  SpMat jac = jac_wrt_nodes_structure_;
  switch (ang_acc_dim) {
    case X:
      jac.setRow(Y, sin(z)*zd*jac_z - cos(z)*jac_zd);
      jac.setRow(X, sin(y)*sin(z)*yd*jac_z - cos(y)*sin(z)*jac_zd - cos(y)*cos(z)*yd*jac_y - cos(y)*cos(z)*zd*jac_z - cos(z)*sin(y)*jac_yd + sin(y)*sin(z)*jac_y*zd);
      break;
    case Y:
      jac.setRow(Y, -sin(z)*jac_zd - cos(z)*zd*jac_z);
      jac.setRow(X, cos(y)*cos(z)*jac_zd - sin(y)*sin(z)*jac_yd - cos(y)*sin(z)*yd*jac_y - cos(z)*sin(y)*yd*jac_z - cos(z)*sin(y)*jac_y*zd - cos(y)*sin(z)*zd*jac_z);
      break;
    case Z:
      jac.setRow(X, sin(y)*yd*jac_y - cos(y)*jac_yd);
      break;
  }
  return jac;
}

}  // namespace orc
