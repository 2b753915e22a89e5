// ORACLE — TEST INFRASTRUCTURE ONLY (see towr_oracle.h header).
// Single-rigid-body dynamics, robot constants and analytic terrains.
#include <cmath>
#include <stdexcept>

#include "towr_oracle.h"

namespace orc {

// ========================================== single_rigid_body_dynamics.cc
// :36-44 (off-diagonals negated)
static M3 BuildInertiaTensor(double Ixx, double Iyy, double Izz, double Ixy, double Ixz, double Iyz) {
  M3 I;
  I(0,0) = Ixx;  I(0,1) = -Ixy; I(0,2) = -Ixz;
  I(1,0) = -Ixy; I(1,1) = Iyy;  I(1,2) = -Iyz;
  I(2,0) = -Ixz; I(2,1) = -Iyz; I(2,2) = Izz;
  return I;
}

// :47-57  in x v = Cross(in)*v ; six explicitly stored off-diagonal entries
static SpMat Cross(const V3& in) {
  SpMat out(3, 3);
  out.coeffRef(0, 1) = -in(2); out.coeffRef(0, 2) = in(1);
  out.coeffRef(1, 0) = in(2);  out.coeffRef(1, 2) = -in(0);
  out.coeffRef(2, 0) = -in(1); out.coeffRef(2, 1) = in(0);
  return out;
}

// :59-74 ; dynamic_model.cc:34-49
SingleRigidBodyDynamics::SingleRigidBodyDynamics(double mass, double Ixx, double Iyy, double Izz,
                                                 double Ixy, double Ixz, double Iyz, int ee_count)
    : m_(mass), g_(9.80665) {
  I_b = sparseView(BuildInertiaTensor(Ixx, Iyy, Izz, Ixy, Ixz, Iyz), /*keep_zeros=*/false);  // inertia_b.sparseView()
  w_R_b_(0,0) = w_R_b_(1,1) = w_R_b_(2,2) = 1.0;
  ee_force_.assign(ee_count, V3());
  ee_pos_.assign(ee_count, V3());
  ee_torque_.assign(ee_count, V3());
}

// dynamic_model.cc:51-66
void SingleRigidBodyDynamics::SetCurrent(const V3& com_W, const V3& com_acc_W, const M3& w_R_b,
                                         const V3& omega_W, const V3& omega_dot_W,
                                         const std::vector<V3>& force_W, const std::vector<V3>& pos_W,
                                         const std::vector<V3>& torque_W) {
  com_pos_ = com_W; com_acc_ = com_acc_W;
  w_R_b_ = w_R_b; omega_ = omega_W; omega_dot_ = omega_dot_W;
  ee_force_ = force_W; ee_pos_ = pos_W; ee_torque_ = torque_W;
}

// :76-102
std::array<double, 6> SingleRigidBodyDynamics::GetDynamicViolation() const {
  V3 f_sum, tau_sum;
  for (size_t ee = 0; ee < ee_pos_.size(); ++ee) {
    V3 f = ee_force_.at(ee);
    V3 tau_contact = ee_torque_.at(ee);
    tau_sum += cross(f, com_pos_ - ee_pos_.at(ee)) + tau_contact;
    f_sum += f;
  }
  SpMat I_w = sparseView(w_R_b_, false) * I_b * sparseView(w_R_b_.transpose(), false);
  V3 ang = I_w * omega_dot_ + Cross(omega_) * (I_w * omega_) - tau_sum;
  V3 lin = m() * com_acc_ - f_sum - V3(0.0, 0.0, -m() * g());
  return {ang(0), ang(1), ang(2), lin(0), lin(1), lin(2)};
}

// :104-122
SpMat SingleRigidBodyDynamics::GetJacobianWrtBaseLin(const SpMat& jac_pos_base_lin, const SpMat& jac_acc_base_lin) const {
  int n = jac_pos_base_lin.cols;
  SpMat jac_tau_sum(k3D, n);
  for (const V3& f : ee_force_) {
    SpMat jac_tau = Cross(f) * jac_pos_base_lin;
    jac_tau_sum += jac_tau;
  }
  SpMat jac(k6D, n);
  jac.setMiddleRows(AX, -jac_tau_sum);
  jac.setMiddleRows(LX, m() * jac_acc_base_lin);
  return jac;
}

// :124-166
SpMat SingleRigidBodyDynamics::GetJacobianWrtBaseAng(const EulerConverter& base_euler, double t) const {
  SpMat R_sv = sparseView(w_R_b_, false);
  SpMat I_w = R_sv * I_b * sparseView(w_R_b_.transpose(), false);

  // derivative of R*I_b*R^T * wd : product rule over R, R^T, wd
  V3 v11 = (I_b * w_R_b_.transpose()) * omega_dot_;
  SpMat jac11 = base_euler.DerivOfRotVecMult(t, v11, false);
  SpMat jac12 = R_sv * I_b * base_euler.DerivOfRotVecMult(t, omega_dot_, true);
  SpMat jac_ang_acc = base_euler.GetDerivOfAngAccWrtNodes(t);
  SpMat jac13 = I_w * jac_ang_acc;
  SpMat jac1 = jac11 + jac12 + jac13;

  // derivative of w x (I_w w)
  V3 v21 = (I_b * w_R_b_.transpose()) * omega_;
  SpMat jac21 = base_euler.DerivOfRotVecMult(t, v21, false);
  SpMat jac22 = R_sv * I_b * base_euler.DerivOfRotVecMult(t, omega_, true);
  SpMat jac_ang_vel = base_euler.GetDerivOfAngVelWrtNodes(t);
  SpMat jac23 = I_w * jac_ang_vel;
  SpMat jac2 = Cross(omega_) * (jac21 + jac22 + jac23) - Cross(I_w * omega_) * jac_ang_vel;

  int n = jac_ang_vel.cols;
  SpMat jac(k6D, n);
  jac.setMiddleRows(AX, jac1 + jac2);
  return jac;
}

// :168-180
SpMat SingleRigidBodyDynamics::GetJacobianWrtForce(const SpMat& jac_force, int ee) const {
  V3 r = com_pos_ - ee_pos_.at(ee);
  SpMat jac_tau = (-Cross(r)) * jac_force;
  SpMat jac(k6D, jac_force.cols);
  jac.setMiddleRows(AX, -jac_tau);
  jac.setMiddleRows(LX, -jac_force);
  return jac;
}

// :182-191
SpMat SingleRigidBodyDynamics::GetJacobianWrtTorque(const SpMat& jac_torque, int) const {
  SpMat jac(k6D, jac_torque.cols);
  jac.setMiddleRows(AX, -jac_torque);
  return jac;
}

// :193-204
SpMat SingleRigidBodyDynamics::GetJacobianWrtEEPos(const SpMat& jac_ee_pos, int ee) const {
  V3 f = ee_force_.at(ee);
  SpMat jac_tau = Cross(f) * (-jac_ee_pos);
  SpMat jac(k6D, jac_tau.cols);
  jac.setMiddleRows(AX, -jac_tau);
  return jac;
}

// ================================================================ robot_model.cc
// models/examples/{monoped,biped,hyq,anymal}_model.h ; foot order LF,RF,LH,RH (endeffector_mappings.h:43-44)
RobotModel RobotModel::Make(Robot r) {
  RobotModel rm;
  auto kin = std::make_shared<KinematicModel>();
  auto quad = [&](double xn, double yn, double zn, const V3& dev) {
    kin->nominal_stance_ = {V3(xn, yn, zn), V3(xn, -yn, zn), V3(-xn, yn, zn), V3(-xn, -yn, zn)};
    kin->max_dev_.assign(4, dev);
    kin->min_dev_.assign(4, -dev);
  };
  switch (r) {
    case Monoped:
      rm.dynamic_model_ = std::make_shared<SingleRigidBodyDynamics>(20, 1.2, 5.5, 6.0, 0.0, -0.2, -0.01, 1);
      kin->nominal_stance_ = {V3(0.0, 0.0, -0.58)};
      kin->max_dev_ = {V3(0.30, 0.15, 0.30)};
      kin->min_dev_ = {V3(-0.30, -0.15, -0.30)};
      break;
    case Biped:
      rm.dynamic_model_ = std::make_shared<SingleRigidBodyDynamics>(20, 1.209, 5.583, 6.056, 0.005, -0.190, -0.012, 2);
      kin->nominal_stance_ = {V3(0.0, 0.20, -0.65), V3(0.0, -0.20, -0.65)};
      kin->max_dev_.assign(2, V3(0.25, 0.15, 0.40));
      kin->min_dev_.assign(2, V3(-0.25, -0.15, -0.40));
      break;
    case Hyq:
      rm.dynamic_model_ = std::make_shared<SingleRigidBodyDynamics>(83, 4.26, 8.97, 9.88, -0.0063, 0.193, 0.0126, 4);
      quad(0.31, 0.29, -0.58, V3(0.25, 0.20, 0.10));
      break;
    case Anymal:
      rm.dynamic_model_ = std::make_shared<SingleRigidBodyDynamics>(29.5, 0.946438, 1.94478, 2.01835, 0.000938112, -0.00595386, -0.00146328, 4);
      quad(0.34, 0.19, -0.42, V3(0.15, 0.1, 0.10));
      break;
    default: throw std::runtime_error("robot model not implemented");
  }
  rm.kinematic_model_ = kin;
  return rm;
}

// ================================================================= height_map.cc
// :76-90
V3 HeightMap::GetDerivativeOfNormalizedBasisWrt(Direction basis, int dim, double x, double y) const {
  V3 dv_wrt_dim = GetBasis(basis, x, y, dim);       // inner derivative
  V3 v = GetBasis(basis, x, y, -1);                 // outer derivative
  // :136-143  1/v.squaredNorm()*(v.norm()*Unit(idx) - v(idx)*v.normalized())
  V3 dn_norm_wrt_n = (1 / squaredNorm(v)) * (norm(v) * V3::Unit(dim) - v(dim) * normalized(v));
  return cwiseProduct(dn_norm_wrt_n, dv_wrt_dim);
}

// :62-74 dispatch, :92-134 normal / tangents (deriv_dim >= 0: derivative of the un-normalised vector)
V3 HeightMap::GetBasis(Direction basis, double x, double y, int deriv_dim) const {
  bool basis_requested = deriv_dim < 0;
  V3 out;
  switch (basis) {
    case Normal:
      for (int dim : {0, 1})
        out(dim) = basis_requested ? -GetDerivativeOfHeightWrt(dim, x, y)
                                   : -GetSecondDerivativeOfHeightWrt(dim, deriv_dim, x, y);
      out(2) = basis_requested ? 1.0 : 0.0;
      break;
    case Tangent1:
      out(0) = basis_requested ? 1.0 : 0.0;
      out(1) = 0.0;
      out(2) = basis_requested ? GetDerivativeOfHeightWrt(0, x, y) : GetSecondDerivativeOfHeightWrt(0, deriv_dim, x, y);
      break;
    case Tangent2:
      out(0) = 0.0;
      out(1) = basis_requested ? 1.0 : 0.0;
      out(2) = basis_requested ? GetDerivativeOfHeightWrt(1, x, y) : GetSecondDerivativeOfHeightWrt(1, deriv_dim, x, y);
      break;
  }
  return out;
}

// :145-161
double HeightMap::GetSecondDerivativeOfHeightWrt(int dim1, int dim2, double x, double y) const {
  if (dim1 == 0) return dim2 == 0 ? GetHeightDerivWrtXX(x, y) : GetHeightDerivWrtXY(x, y);
  return dim2 == 0 ? GetHeightDerivWrtYX(x, y) : GetHeightDerivWrtYY(x, y);
}

// ======================================================= height_map_examples.cc
namespace {
struct FlatGround : HeightMap {       // examples.h:45-52
  double GetHeight(double, double) const override { return 0.0; }
};
struct Block : HeightMap {            // examples.h:57-70, .cc:43-68
  double block_start = 0.7, length_ = 3.5, height_ = 0.5, eps_ = 0.03;
  const double slope_ = height_ / eps_;
  double GetHeight(double x, double) const override {
    double h = 0.0;
    if (block_start <= x && x <= block_start + eps_) h = slope_ * (x - block_start);
    if (block_start + eps_ <= x && x <= block_start + length_) h = height_;
    return h;
  }
  double GetHeightDerivWrtX(double x, double) const override {
    double dhdx = 0.0;
    if (block_start <= x && x <= block_start + eps_) dhdx = slope_;
    return dhdx;
  }
};
struct Stairs : HeightMap {           // examples.h:75-86, .cc:72-88
  double first_step_start_ = 1.0, first_step_width_ = 0.4, height_first_step = 0.2, height_second_step = 0.4, width_top = 1.0;
  double GetHeight(double x, double) const override {
    double h = 0.0;
    if (x >= first_step_start_) h = height_first_step;
    if (x >= first_step_start_ + first_step_width_) h = height_second_step;
    if (x >= first_step_start_ + first_step_width_ + width_top) h = 0.0;
    return h;
  }
};
struct Gap : HeightMap {              // examples.h:91-114, .cc:92-124
  const double gap_start_ = 1.0, w = 0.5, h = 1.5;
  const double slope_ = h / w, dx = w / 2.0, xc = gap_start_ + dx, gap_end_x = gap_start_ + w;
  const double a = (4 * h) / (w * w);
  const double b = -(8 * h * xc) / (w * w);
  const double c = -(h * (w - 2 * xc) * (w + 2 * xc)) / (w * w);
  double GetHeight(double x, double) const override {
    double hh = 0.0;
    if (gap_start_ <= x && x <= gap_end_x) hh = a * x * x + b * x + c;
    return hh;
  }
  double GetHeightDerivWrtX(double x, double) const override {
    double dhdx = 0.0;
    if (gap_start_ <= x && x <= gap_end_x) dhdx = 2 * a * x + b;
    return dhdx;
  }
  double GetHeightDerivWrtXX(double x, double) const override {
    double dzdxx = 0.0;
    if (gap_start_ <= x && x <= gap_end_x) dzdxx = 2 * a;
    return dzdxx;
  }
};
struct Slope : HeightMap {            // examples.h:119-133, .cc:128-160
  const double slope_start_ = 1.0, up_length_ = 1.0, down_length_ = 1.0, height_center = 0.7;
  const double x_down_start_ = slope_start_ + up_length_;
  const double x_flat_start_ = x_down_start_ + down_length_;
  const double slope_ = height_center / up_length_;
  double GetHeight(double x, double) const override {
    double z = 0.0;
    if (x >= slope_start_) z = slope_ * (x - slope_start_);
    if (x >= x_down_start_) z = height_center - slope_ * (x - x_down_start_);
    if (x >= x_flat_start_) z = 0.0;
    return z;
  }
  double GetHeightDerivWrtX(double x, double) const override {
    double dzdx = 0.0;
    if (x >= slope_start_) dzdx = slope_;
    if (x >= x_down_start_) dzdx = -slope_;
    if (x >= x_flat_start_) dzdx = 0.0;
    return dzdx;
  }
};
struct Chimney : HeightMap {          // examples.h:138-150, .cc:164-184
  const double x_start_ = 1.0, length_ = 1.5, y_start_ = 0.5, slope_ = 3.0;
  const double x_end_ = x_start_ + length_;
  double GetHeight(double x, double y) const override {
    double z = 0.0;
    if (x_start_ <= x && x <= x_end_) z = slope_ * (y - y_start_);
    return z;
  }
  double GetHeightDerivWrtY(double x, double) const override {
    double dzdy = 0.0;
    if (x_start_ <= x && x <= x_end_) dzdy = slope_;
    return dzdy;
  }
};
struct ChimneyLR : HeightMap {        // examples.h:155-168, .cc:188-211
  const double x_start_ = 0.5, length_ = 1.0, y_start_ = 0.5, slope_ = 2;
  const double x_end1_ = x_start_ + length_, x_end2_ = x_start_ + 2 * length_;
  double GetHeight(double x, double y) const override {
    double z = 0.0;
    if (x_start_ <= x && x <= x_end1_) z = slope_ * (y - y_start_);
    if (x_end1_ <= x && x <= x_end2_) z = -slope_ * (y + y_start_);
    return z;
  }
  double GetHeightDerivWrtY(double x, double) const override {
    double dzdy = 0.0;
    if (x_start_ <= x && x <= x_end1_) dzdy = slope_;
    if (x_end1_ <= x && x <= x_end2_) dzdy = -slope_;
    return dzdy;
  }
};
struct FiveStepStairs : HeightMap {   // test/hopper_example.cc:53-86
  double step_depth_ = 0.3, step_height_ = 0.15, stairs_start_ = 0.5;
  int num_steps_ = 5;
  double GetHeight(double x, double) const override {
    if (x < stairs_start_) return 0.0;
    double relative_x = x - stairs_start_;
    int step_number = static_cast<int>(relative_x / step_depth_);
    if (step_number >= num_steps_) return num_steps_ * step_height_;
    return (step_number + 1) * step_height_;
  }
};
// User height map as data: knots x_[0] < ... < x_[n-1], heights h_.  steps_: h = 0 before the first knot, h_[i] on
// [x_[i], x_[i+1]), h_.back() from the last knot on, no slope.  Otherwise ramps: h_[0] before the first knot, linear on
// [x_[i], x_[i+1]) with that segment's slope as GetHeightDerivWrtX, h_.back() and no slope from the last knot on.
struct TableTerrain : HeightMap {
  bool steps_;
  std::vector<double> x_, h_;
  TableTerrain(bool steps, std::vector<double> x, std::vector<double> h) : steps_(steps), x_(std::move(x)), h_(std::move(h)) {}
  int Interval(double x) const {      // -1: before the first knot; i: x_[i] <= x (< x_[i+1] when there is one)
    int i = -1;
    for (size_t k = 0; k < x_.size(); ++k)
      if (x >= x_[k]) i = static_cast<int>(k);
    return i;
  }
  double GetHeight(double x, double) const override {
    const int i = Interval(x);
    if (i < 0) return steps_ ? 0.0 : h_.front();
    if (steps_ || i + 1 == static_cast<int>(x_.size())) return h_[i];
    return h_[i] + GetHeightDerivWrtX(x, 0.0) * (x - x_[i]);
  }
  double GetHeightDerivWrtX(double x, double) const override {
    const int i = Interval(x);
    if (steps_ || i < 0 || i + 1 == static_cast<int>(x_.size())) return 0.0;
    return (h_[i + 1] - h_[i]) / (x_[i + 1] - x_[i]);
  }
};
}  // namespace

std::shared_ptr<HeightMap> HeightMap::MakeTableTerrain(int mode, const std::vector<double>& x, const std::vector<double>& h) {
  if (x.empty() || x.size() != h.size()) throw std::runtime_error("terrain table: bad knot count");
  return std::make_shared<TableTerrain>(mode == 0, x, h);
}

// height_map.cc:38-51
std::shared_ptr<HeightMap> HeightMap::MakeTerrain(int id) {
  switch (id) {
    case FlatID: return std::make_shared<FlatGround>();
    case BlockID: return std::make_shared<Block>();
    case StairsID: return std::make_shared<Stairs>();
    case GapID: return std::make_shared<Gap>();
    case SlopeID: return std::make_shared<Slope>();
    case ChimneyID: return std::make_shared<Chimney>();
    case ChimneyLRID: return std::make_shared<ChimneyLR>();
    case FiveStepStairsID: return std::make_shared<FiveStepStairs>();
    default: throw std::runtime_error("unknown terrain id");
  }
}

}  // namespace orc
