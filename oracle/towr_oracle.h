// ORACLE — TEST INFRASTRUCTURE ONLY (see sparse_shim.h header).
//
// CPU restatement of the hot path of hexb66/towr2025 (constraint values and
// sparse-Jacobian values of TOWR's NLP), written without Eigen/ifopt/Ipopt,
// which are not installed here and not vendored by the reference.
// Every function cites the reference file:line it follows (paths relative to
// /root/reference/towr/).
//
// PARITY UNPINNED: the reference's own tests for this path are empty stubs
// (test/dynamic_constraint_test.cc:40-43, test/dynamic_model_test.cc:36-49) and
// it cannot be compiled here, so there is no golden vector or reference output
// to pin this restatement against.  Its authority is line-by-line fidelity to
// the cited source, plus the substitutes tests/ run: Hermite identities,
// finite-difference Jacobian checks, physical known answers and the structure
// sizes of SURVEY.md §8 (1042 / 2732 / 23992 for ANYmal fly-trot 2 s), and the
// second, independently written restatement oracle_np.py (numpy/scipy.sparse, from
// the reference's Eigen expressions): tests/test_oracle_cross.py holds the two to
// 1e-12 / 1e-14 on values and to an identical stored-entry pattern on 29 classes.
#pragma once
#include <array>
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "sparse_shim.h"

namespace orc {

// ---------------------------------------------------------------- basic types
enum Dx { kPos = 0, kVel, kAcc };
enum Dim3D { X = 0, Y, Z };
enum Dim6D { AX = 0, AY, AZ, LX, LY, LZ };
constexpr int k2D = 2, k3D = 3, k6D = 6;

struct Bounds {                       // ifopt::Bounds
  double lower, upper;
  Bounds(double lo = 0.0, double up = 0.0) : lower(lo), upper(up) {}
  void operator+=(double s) { lower += s; upper += s; }
};
constexpr double kInf = 1.0e20;      // ifopt::inf
inline Bounds NoBound() { return Bounds(-kInf, +kInf); }
inline Bounds BoundZero() { return Bounds(0.0, 0.0); }
inline Bounds BoundGreaterZero() { return Bounds(0.0, +kInf); }
inline Bounds BoundSmallerZero() { return Bounds(-kInf, 0.0); }
using VecBound = std::vector<Bounds>;
using Vec = std::vector<double>;

struct State3 { V3 p, v, a; };        // towr::State with 3 derivatives, dim 3
struct Node { V3 p, v; const V3& at(Dx d) const { return d == kPos ? p : v; } V3& at(Dx d) { return d == kPos ? p : v; } };
struct BaseState { Node lin, ang; };

// ------------------------------------------------------------------ polynomial
// helpers/polynomial.cc:47-257
class CubicHermitePolynomial {
 public:
  void SetNodes(const Node& n0, const Node& n1) { n0_ = n0; n1_ = n1; }
  void SetDuration(double T) { T_ = T; }
  double GetDuration() const { return T_; }
  void UpdateCoeff();
  State3 GetPoint(double t_local) const;
  double GetDerivativeWrtStartNode(Dx dfdt, Dx node_deriv, double t) const;
  double GetDerivativeWrtEndNode(Dx dfdt, Dx node_deriv, double t) const;
  V3 GetDerivativeOfPosWrtDuration(double t) const;
  // TEST SWITCH (not in the reference): polynomials of a PhaseSpline are marked; with SetPowMode(1) their integer powers
  // are the correctly rounded ones the device computes at run time instead of glibc's std::pow (see Pw in splines.cc)
  void MarkRuntimeDurations() { runtime_ = true; }
 private:
  double Pw(double x, int n) const;
  bool runtime_ = false;
  double T_ = 0.0;
  Node n0_, n1_;
  V3 coeff_[4];
};

// 0 (default): std::pow everywhere, as the reference.  1: marked polynomials use correctly rounded x^2, x^3, x^4.
void SetPowMode(int mode);
int GetPowMode();
// x^n for n = 0..4, correctly rounded (n = 3, 4: up to the rare double-rounding case): the device's Cube / Pow4
// (towr2025_b200/csrc/device_math.cuh) restated
double ExactPow(double x, int n);

// ------------------------------------------------------------- variable sets
class VariableSet {
 public:
  VariableSet(int n, std::string name) : rows_(n), name_(std::move(name)) {}
  virtual ~VariableSet() = default;
  virtual Vec GetValues() const = 0;
  virtual void SetVariables(const Vec& x) = 0;
  virtual VecBound GetBounds() const = 0;
  int GetRows() const { return rows_; }
  const std::string& GetName() const { return name_; }
 protected:
  void SetRows(int n) { rows_ = n; }
 private:
  int rows_;
  std::string name_;
};

class NodeSpline;

// variables/nodes_variables.cc:38-278
class NodesVariables : public VariableSet {
 public:
  struct NodeValueInfo {
    int id; Dx deriv; int dim;
    bool operator==(const NodeValueInfo& o) const { return id == o.id && deriv == o.deriv && dim == o.dim; }
  };
  enum Side { Start = 0, End };
  static constexpr int NodeValueNotOptimized = -1;

  explicit NodesVariables(const std::string& name) : VariableSet(-1, name) {}
  virtual std::vector<NodeValueInfo> GetNodeValuesInfo(int opt_idx) const = 0;
  int GetOptIndex(const NodeValueInfo& nvi) const;
  Vec GetValues() const override;
  void SetVariables(const Vec& x) override;
  VecBound GetBounds() const override { return bounds_; }
  const std::vector<Node>& GetNodes() const { return nodes_; }
  int GetPolynomialCount() const { return static_cast<int>(nodes_.size()) - 1; }
  static int GetNodeId(int poly_id, Side side) { return poly_id + side; }
  void AddObserver(NodeSpline* s) { observers_.push_back(s); }
  void SetByLinearInterpolation(const V3& initial, const V3& final, double t_total);
  void SetByLinearInterpolationRelativeToBase(const V3& ee_initial_W, const V3& ee_final_W,
                                              const V3& base_pos_initial_W, const V3& base_pos_final_W,
                                              const V3& base_rpy_initial_W, const V3& base_rpy_final_W,
                                              double t_total);
  void AddBounds(int node_id, Dx deriv, const std::vector<int>& dims, const V3& val);
  void AddBounds(int node_id, Dx deriv, const std::vector<int>& dims, const V3& lower, const V3& upper);
  void AddStartBound(Dx d, const std::vector<int>& dims, const V3& val) { AddBounds(0, d, dims, val); }
  void AddFinalBound(Dx d, const std::vector<int>& dims, const V3& val) { AddBounds(static_cast<int>(nodes_.size()) - 1, d, dims, val); }
 protected:
  void AddBound(const NodeValueInfo& nvi, double lower, double upper);
  void UpdateObservers() const;
  std::vector<Node> nodes_;
  VecBound bounds_;
  std::vector<NodeSpline*> observers_;
};

// variables/nodes_variables_all.cc:34-61
class NodesVariablesAll : public NodesVariables {
 public:
  NodesVariablesAll(int n_nodes, const std::string& name);
  std::vector<NodeValueInfo> GetNodeValuesInfo(int idx) const override;
};

// variables/nodes_variables_phase_based.cc:39-396
class NodesVariablesPhaseBased : public NodesVariables {
 public:
  struct PolyInfo { int phase, poly_in_phase, n_polys_in_phase; bool is_constant; };
  enum Kind { Motion, Ang, Force, Torque };
  NodesVariablesPhaseBased(Kind kind, int phase_count, bool in_contact_at_start,
                           const std::string& name, int n_polys_in_changing_phase);
  std::vector<NodeValueInfo> GetNodeValuesInfo(int idx) const override { return index_to_nvi_.at(idx); }
  Vec ConvertPhaseToPolyDurations(const Vec& phase_durations) const;
  double GetDerivativeOfPolyDurationWrtPhaseDuration(int poly_id) const;
  int GetNumberOfPrevPolynomialsInPhase(int poly_id) const;
  bool IsConstantNode(int node_id) const;
  bool IsInConstantPhase(int poly_id) const { return poly_info_.at(poly_id).is_constant; }
  std::vector<int> GetIndicesOfNonConstantNodes() const;
  int GetPhase(int node_id) const;
  int GetPolyIDAtStartOfPhase(int phase) const;
  int GetNodeIDAtStartOfPhase(int phase) const { return GetNodeId(GetPolyIDAtStartOfPhase(phase), Start); }
  V3 GetValueAtStartOfPhase(int phase) const { return nodes_.at(GetNodeIDAtStartOfPhase(phase)).p; }
 private:
  std::vector<int> GetAdjacentPolyIds(int node_id) const;
  void BuildMotionLikeMap(bool pin_swing_vz);   // :215-258 (EEMotion) and :366-396 (EEAng)
  void BuildForceLikeMap();                     // :274-304 (EEForce) and :320-350 (EETorque)
  std::vector<PolyInfo> poly_info_;
  std::map<int, std::vector<NodeValueInfo>> index_to_nvi_;
};

class PhaseSpline;

// variables/phase_durations.cc:40-154
class PhaseDurations : public VariableSet {
 public:
  PhaseDurations(int ee, const Vec& timings, bool first_phase_in_contact, double min_d, double max_d);
  Vec GetValues() const override;
  void SetVariables(const Vec& x) override;
  VecBound GetBounds() const override;
  const Vec& GetPhaseDurations() const { return durations_; }
  bool IsContactPhase(double t) const;
  SpMat GetJacobianOfPos(int current_phase, const V3& dx_dT, const V3& xd) const;
  void AddObserver(PhaseSpline* s) { observers_.push_back(s); }
 private:
  Vec durations_;
  double t_total_;
  bool initial_contact_state_;
  Bounds phase_duration_bounds_;
  std::vector<PhaseSpline*> observers_;
};

// --------------------------------------------------------------------- splines
// helpers/spline.cc:37-124, helpers/node_spline.cc:36-112
class NodeSpline {
 public:
  NodeSpline(NodesVariables* nodes, const Vec& poly_durations);
  virtual ~NodeSpline() = default;
  static int GetSegmentID(double t_global, const Vec& durations);
  std::pair<int, double> GetLocalTime(double t_global, const Vec& durations) const;
  State3 GetPoint(double t_global) const;
  State3 GetPoint(int poly_id, double t_local) const { return polys_.at(poly_id).GetPoint(t_local); }
  Vec GetPolyDurations() const;
  int GetPolynomialCount() const { return static_cast<int>(polys_.size()); }
  int GetNodeVariablesCount() const { return nodes_->GetRows(); }
  void UpdateNodes();
  SpMat GetJacobianWrtNodes(double t_global, Dx dxdt) const;
  SpMat GetJacobianWrtNodes(int poly_id, double t_local, Dx dxdt) const;
  virtual SpMat GetJacobianOfPosWrtDurations(double) const { return SpMat(); }  // node_spline.h:102 assert(false)
 protected:
  void FillJacobianWrtNodes(int poly_id, double t_local, Dx dxdt, SpMat& jac, bool fill_with_zeros) const;
  void UpdatePolynomialCoeff() { for (auto& p : polys_) p.UpdateCoeff(); }
  std::vector<CubicHermitePolynomial> polys_;
  NodesVariables* nodes_;
  SpMat jac_wrt_nodes_structure_;
};

// helpers/phase_spline.cc:35-93
class PhaseSpline : public NodeSpline {
 public:
  PhaseSpline(const std::shared_ptr<NodesVariablesPhaseBased>& nodes, PhaseDurations* durations);
  void UpdatePolynomialDurations();
  SpMat GetJacobianOfPosWrtDurations(double t_global) const override;
 private:
  V3 GetDerivativeOfPosWrtPhaseDuration(double t_global) const;
  std::shared_ptr<NodesVariablesPhaseBased> phase_nodes_;
  PhaseDurations* phase_durations_;
};

// helpers/euler_converter.cc:39-310
class EulerConverter {
 public:
  using JacRowMatrix = std::array<std::array<SpVec, 3>, 3>;
  explicit EulerConverter(const std::shared_ptr<NodeSpline>& euler);
  SpMat GetRotationMatrixBaseToWorld(double t) const;
  static SpMat GetRotationMatrixBaseToWorld(const V3& xyz);
  static M3 GetRotationMatrixBaseToWorldDense(const V3& xyz);
  V3 GetAngularVelocityInWorld(double t) const;
  V3 GetAngularAccelerationInWorld(double t) const;
  SpMat GetDerivOfAngVelWrtNodes(double t) const;
  SpMat GetDerivOfAngAccWrtNodes(double t) const;
  SpMat DerivOfRotVecMult(double t, const V3& v, bool inverse) const;
  static SpMat GetM(const V3& xyz);
  static SpMat GetMdot(const V3& xyz, const V3& xyz_d);
 private:
  SpMat GetDerivMwrtNodes(double t, Dim3D dim) const;
  SpMat GetDerivMdotwrtNodes(double t, Dim3D dim) const;
  JacRowMatrix GetDerivativeOfRotationMatrixWrtNodes(double t) const;
  SpVec GetJac(double t, Dx deriv, Dim3D dim) const { return euler_->GetJacobianWrtNodes(t, deriv).row(dim); }
  std::shared_ptr<NodeSpline> euler_;
  SpMat jac_wrt_nodes_structure_;
};

struct SplineHolder {                 // variables/spline_holder.h:48-84, helpers/spline_holder.cc:35-69
  std::shared_ptr<NodeSpline> base_linear_, base_angular_;
  std::shared_ptr<EulerConverter> angular_converter_;
  std::vector<std::shared_ptr<NodeSpline>> ee_motion_, ee_ang_, ee_force_, ee_torque_;
  std::vector<std::shared_ptr<PhaseDurations>> phase_durations_;
};

// ---------------------------------------------------------------------- models
// models/dynamic_model.cc:34-66, models/single_rigid_body_dynamics.cc:36-204
class SingleRigidBodyDynamics {
 public:
  SingleRigidBodyDynamics(double mass, double Ixx, double Iyy, double Izz, double Ixy, double Ixz, double Iyz, int ee_count);
  void SetCurrent(const V3& com_W, const V3& com_acc_W, const M3& w_R_b, const V3& omega_W, const V3& omega_dot_W,
                  const std::vector<V3>& force_W, const std::vector<V3>& pos_W, const std::vector<V3>& torque_W);
  std::array<double, 6> GetDynamicViolation() const;
  SpMat GetJacobianWrtBaseLin(const SpMat& jac_pos_base_lin, const SpMat& jac_acc_base_lin) const;
  SpMat GetJacobianWrtBaseAng(const EulerConverter& base_euler, double t) const;
  SpMat GetJacobianWrtForce(const SpMat& jac_force, int ee) const;
  SpMat GetJacobianWrtTorque(const SpMat& jac_torque, int ee) const;
  SpMat GetJacobianWrtEEPos(const SpMat& jac_ee_pos, int ee) const;
  double m() const { return m_; }
  double g() const { return g_; }
  M3 GetInertiaB() const { return toDense(I_b); }   // single_rigid_body_dynamics.h:97
  int GetEECount() const { return static_cast<int>(ee_pos_.size()); }
 private:
  double m_, g_;
  SpMat I_b;
  V3 com_pos_, com_acc_, omega_, omega_dot_;
  M3 w_R_b_;
  std::vector<V3> ee_force_, ee_pos_, ee_torque_;
};

struct KinematicModel {               // models/kinematic_model.h:49-104 + models/examples/*_model.h
  std::vector<V3> nominal_stance_, max_dev_, min_dev_;
};
struct RobotModel {                   // models/robot_model.cc:36-64
  enum Robot { Monoped = 0, Biped, Hyq, Anymal };
  std::shared_ptr<SingleRigidBodyDynamics> dynamic_model_;
  std::shared_ptr<KinematicModel> kinematic_model_;
  static RobotModel Make(Robot r);
};

// --------------------------------------------------------------------- terrain
// terrain/height_map.cc:38-163, terrain/height_map_examples.cc:35-211, examples/height_map_examples.h:45-166
class HeightMap {
 public:
  enum TerrainID { FlatID = 0, BlockID, StairsID, GapID, SlopeID, ChimneyID, ChimneyLRID, FiveStepStairsID /* test/hopper_example.cc:53-86 */ };
  enum Direction { Normal, Tangent1, Tangent2 };
  virtual ~HeightMap() = default;
  static std::shared_ptr<HeightMap> MakeTerrain(int id);
  // a user-defined height map given as data (orc_spec::terrain_table_*): what deriving from HeightMap and overriding
  // GetHeight / GetHeightDerivWrtX is in the reference (test/hopper_example.cc:53-91), for a piecewise profile along x
  static std::shared_ptr<HeightMap> MakeTableTerrain(int mode, const std::vector<double>& x, const std::vector<double>& h);
  virtual double GetHeight(double x, double y) const = 0;
  double GetDerivativeOfHeightWrt(int dim, double x, double y) const { return dim == 0 ? GetHeightDerivWrtX(x, y) : GetHeightDerivWrtY(x, y); }
  V3 GetNormalizedBasis(Direction basis, double x, double y) const { return normalized(GetBasis(basis, x, y, -1)); }
  V3 GetDerivativeOfNormalizedBasisWrt(Direction basis, int dim, double x, double y) const;
  double GetFrictionCoeff() const { return friction_coeff_; }
 protected:
  double friction_coeff_ = 0.5;
  virtual double GetHeightDerivWrtX(double, double) const { return 0.0; }
  virtual double GetHeightDerivWrtY(double, double) const { return 0.0; }
  virtual double GetHeightDerivWrtXX(double, double) const { return 0.0; }
  virtual double GetHeightDerivWrtXY(double, double) const { return 0.0; }
  virtual double GetHeightDerivWrtYX(double, double) const { return 0.0; }
  virtual double GetHeightDerivWrtYY(double, double) const { return 0.0; }
 private:
  V3 GetBasis(Direction basis, double x, double y, int deriv_dim /* -1 = basis itself */) const;
  double GetSecondDerivativeOfHeightWrt(int dim1, int dim2, double x, double y) const;
};

// ----------------------------------------------------------------- constraints
class Variables {                     // ifopt::Composite of variable sets (insertion order = column order)
 public:
  void Add(const std::shared_ptr<VariableSet>& v) { sets_.push_back(v); }
  const std::vector<std::shared_ptr<VariableSet>>& GetComponents() const { return sets_; }
  std::shared_ptr<VariableSet> GetComponent(const std::string& name) const;
  template <class T> std::shared_ptr<T> GetComponent(const std::string& name) const { return std::dynamic_pointer_cast<T>(GetComponent(name)); }
  int GetRows() const { int n = 0; for (auto& s : sets_) n += s->GetRows(); return n; }
 private:
  std::vector<std::shared_ptr<VariableSet>> sets_;
};

class ConstraintSet {                 // ifopt::ConstraintSet surface
 public:
  ConstraintSet(int rows, std::string name) : rows_(rows), name_(std::move(name)) {}
  virtual ~ConstraintSet() = default;
  virtual Vec GetValues() const = 0;
  virtual VecBound GetBounds() const = 0;
  virtual void FillJacobianBlock(const std::string& var_set, SpMat& jac) const = 0;
  virtual void InitVariableDependedQuantities(const std::shared_ptr<Variables>&) {}
  void LinkWithVariables(const std::shared_ptr<Variables>& x) { variables_ = x; InitVariableDependedQuantities(x); }
  int GetRows() const { return rows_; }
  const std::string& GetName() const { return name_; }
 protected:
  void SetRows(int n) { rows_ = n; }
  std::shared_ptr<Variables> variables_;
 private:
  int rows_;
  std::string name_;
};

// constraints/time_discretization_constraint.cc:37-96
class TimeDiscretizationConstraint : public ConstraintSet {
 public:
  TimeDiscretizationConstraint(double T, double dt, const std::string& name);
  Vec GetValues() const override;
  VecBound GetBounds() const override;
  void FillJacobianBlock(const std::string& var_set, SpMat& jac) const override;
  const Vec& SampleTimes() const { return dts_; }
 protected:
  int GetNumberOfNodes() const { return static_cast<int>(dts_.size()); }
  virtual void UpdateConstraintAtInstance(double t, int k, Vec& g) const = 0;
  virtual void UpdateBoundsAtInstance(double t, int k, VecBound& b) const = 0;
  virtual void UpdateJacobianAtInstance(double t, int k, const std::string& var_set, SpMat& jac) const = 0;
  Vec dts_;
};

struct Parameters;
struct NlpFormulation;

// ---------------------------------------------------------------- parameters
// parameters.cc:40-167, parameters.h:135-336
struct Parameters {
  enum ConstraintName { Dynamic = 0, EndeffectorRom, TotalTime, Terrain, TerrainHard, Force, Torque, Swing, BaseRom, BaseAcc, BaseHeight };
  Parameters();
  std::vector<Vec> ee_phase_durations_;
  std::vector<bool> ee_in_contact_at_start_;
  std::vector<ConstraintName> constraints_;
  double dt_constraint_dynamic_, dt_constraint_range_of_motion_, dt_constraint_base_motion_;
  double dt_constraint_force_, dt_constraint_torque_;
  double dt_cost_energy_, energy_cost_torque_weight_, dt_cost_ang_mom_;   // parameters.h:245-247
  bool enable_swing_ee_base_pos_tracking = false;                          // parameters.h:233-235
  double swing_ee_base_pos_tracking_weight_ = 1e-2, dt_cost_swing_ee_base_pos_tracking_ = 0.05;
  double duration_base_polynomial_;
  int ee_polynomials_per_swing_phase_, force_polynomials_per_stance_phase_, torque_polynomials_per_stance_phase_;
  double force_limit_in_normal_direction_;
  bool enable_stance_tracking = true;
  double torque_tx_min_, torque_tx_max_, torque_ty_min_, torque_ty_max_, torque_k_friction_;
  std::vector<double> ee_swing_height_min_, ee_swing_height_max_;
  std::vector<int> bounds_final_lin_pos_, bounds_final_lin_vel_, bounds_final_ang_pos_, bounds_final_ang_vel_;
  std::pair<double, double> bound_phase_duration_;
  Bounds base_rom_ax, base_rom_ay, base_rom_lz;
  std::vector<std::vector<std::vector<double>>> ee_stance_position_;
  struct EELinearConstraintDef {        // parameters.h:315-327
    struct Term { int ee; int dim; double coeff; };
    enum Target { Motion, Angle };
    Target target = Motion;
    Dx deriv = kPos;
    std::vector<Term> terms;
    double tolerance = 0.0;
    double dt = 0.1;
  };
  std::vector<EELinearConstraintDef> ee_linear_constraints_;
  void OptimizePhaseDurations() { constraints_.push_back(TotalTime); }
  Vec GetBasePolyDurations() const;
  int GetPhaseCount(int ee) const { return static_cast<int>(ee_phase_durations_.at(ee).size()); }
  bool IsOptimizeTimings() const;
  int GetEECount() const { return static_cast<int>(ee_in_contact_at_start_.size()); }
  double GetTotalTime() const;
};

// nlp_formulation.cc:76-602
struct NlpFormulation {
  BaseState initial_base_, final_base_;
  std::vector<V3> initial_ee_W_;
  RobotModel model_;
  std::shared_ptr<HeightMap> terrain_;
  Parameters params_;
  std::vector<std::shared_ptr<VariableSet>> GetVariableSets(SplineHolder& spline_holder);
  std::vector<std::shared_ptr<ConstraintSet>> GetConstraints(const SplineHolder& s) const;
 private:
  std::vector<std::shared_ptr<ConstraintSet>> GetConstraint(Parameters::ConstraintName name, const SplineHolder& s) const;
  mutable SplineHolder spline_holder_for_costs_;
  std::vector<std::shared_ptr<VariableSet>> keep_alive_;   // splines observe variable sets through raw pointers
};

// ifopt::Problem restated from its call sites (SURVEY.md §8b "Who calls it")
class Problem {
 public:
  Problem() : variables_(std::make_shared<Variables>()) {}
  void AddVariableSet(const std::shared_ptr<VariableSet>& v) { variables_->Add(v); }
  void AddConstraintSet(const std::shared_ptr<ConstraintSet>& c) { c->LinkWithVariables(variables_); constraints_.push_back(c); }
  int GetNumberOfOptimizationVariables() const { return variables_->GetRows(); }
  int GetNumberOfConstraints() const { int n = 0; for (auto& c : constraints_) n += c->GetRows(); return n; }
  Vec GetVariableValues() const;
  void SetVariables(const double* x);
  Vec EvaluateConstraints(const double* x);
  // row-major sorted triplets of the full constraint Jacobian at x (values != NULL and values == NULL paths)
  void EvalJacobian(const double* x, std::vector<int>* rows, std::vector<int>* cols, std::vector<double>* vals);
  VecBound GetBoundsOnOptimizationVariables() const;
  VecBound GetBoundsOnConstraints() const;
  const std::vector<std::shared_ptr<ConstraintSet>>& Constraints() const { return constraints_; }
  const std::shared_ptr<Variables>& Vars() const { return variables_; }
 private:
  std::shared_ptr<Variables> variables_;
  std::vector<std::shared_ptr<ConstraintSet>> constraints_;
};

// ------------------------------------------------------------ gait generators
// initialization/gait_generator.cc:54-144 and {monoped,biped,quadruped}_gait_generator.cc
class GaitGenerator {
 public:
  using ContactState = std::vector<bool>;
  using GaitInfo = std::pair<Vec, std::vector<ContactState>>;
  enum Combos { C0 = 0, C1, C2, C3, C4 };
  enum Gaits { Stand = 0, Flight, Walk1, Walk2, Walk2E, Run1, Run2, Run2E, Run3, Run3E, Hop1, Hop1E, Hop2, Hop3, Hop3E, Hop5 };
  virtual ~GaitGenerator() = default;
  static std::shared_ptr<GaitGenerator> MakeGaitGenerator(int leg_count);
  virtual void SetCombo(Combos c) = 0;
  Vec GetPhaseDurations(double T, int ee) const;
  bool IsInContactAtStart(int ee) const { return contacts_.front().at(ee); }
  void SetGaits(const std::vector<Gaits>& gaits);
 protected:
  virtual GaitInfo GetGait(Gaits g) const = 0;
  GaitInfo RemoveTransition(const GaitInfo& g) const;
  std::vector<Vec> GetPhaseDurations() const;
  Vec times_;
  std::vector<ContactState> contacts_;
};

}  // namespace orc
