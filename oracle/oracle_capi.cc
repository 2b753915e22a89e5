// ORACLE — TEST INFRASTRUCTURE ONLY (see towr_oracle.h header).
#include "oracle_capi.h"

#include <cmath>
#include <cstring>
#include <numeric>
#include <exception>
#include <stdexcept>
#include <string>

#include "constraints.h"
#ifdef _OPENMP
#include <omp.h>
#endif

using namespace orc;

static thread_local std::string g_err;

struct orc_problem {
  std::vector<std::pair<int, double>> costs;   // Parameters::costs_ (CostName, weight)
  NlpFormulation formulation;
  SplineHolder splines;
  Problem nlp;
  Vec x0;
  std::vector<V3> swing_ref_B;                 // EEBasePosCost reference positions, one per foot when the cost is on
  struct NodeCostSpec { int set, ee, deriv, dim; double weight; };
  std::vector<NodeCostSpec> node_costs;        // individual NodeCost terms of the spec
};

static NlpFormulation MakeFormulation(const orc_spec& s) {
  NlpFormulation f;
  const int E = s.n_ee;
  f.model_.dynamic_model_ = std::make_shared<SingleRigidBodyDynamics>(s.mass, s.inertia[0], s.inertia[1], s.inertia[2],
                                                                      s.inertia[3], s.inertia[4], s.inertia[5], E);
  auto kin = std::make_shared<KinematicModel>();
  for (int e = 0; e < E; ++e) {
    kin->nominal_stance_.push_back(V3(s.nominal_stance[e][0], s.nominal_stance[e][1], s.nominal_stance[e][2]));
    kin->max_dev_.push_back(V3(s.max_dev[e][0], s.max_dev[e][1], s.max_dev[e][2]));
    kin->min_dev_.push_back(V3(s.min_dev[e][0], s.min_dev[e][1], s.min_dev[e][2]));
  }
  f.model_.kinematic_model_ = kin;
  if (s.terrain_id == 8)
    f.terrain_ = HeightMap::MakeTableTerrain(s.terrain_table_mode,
                                             std::vector<double>(s.terrain_table_x, s.terrain_table_x + s.terrain_table_n),
                                             std::vector<double>(s.terrain_table_h, s.terrain_table_h + s.terrain_table_n));
  else
    f.terrain_ = HeightMap::MakeTerrain(s.terrain_id);

  Parameters& p = f.params_;
  p.constraints_.clear();
  for (int i = 0; i < s.n_constraints; ++i) p.constraints_.push_back(static_cast<Parameters::ConstraintName>(s.constraints[i]));
  for (int e = 0; e < E; ++e) {
    p.ee_phase_durations_.push_back(Vec(s.phase_durations[e], s.phase_durations[e] + s.n_phases[e]));
    p.ee_in_contact_at_start_.push_back(s.in_contact_at_start[e] != 0);
  }
  p.dt_constraint_dynamic_ = s.dt_constraint_dynamic;
  p.dt_constraint_range_of_motion_ = s.dt_constraint_range_of_motion;
  p.dt_constraint_base_motion_ = s.dt_constraint_base_motion;
  p.dt_constraint_force_ = s.dt_constraint_force;
  p.dt_constraint_torque_ = s.dt_constraint_torque;
  p.duration_base_polynomial_ = s.duration_base_polynomial;
  p.ee_polynomials_per_swing_phase_ = s.ee_polynomials_per_swing_phase;
  p.force_polynomials_per_stance_phase_ = s.force_polynomials_per_stance_phase;
  p.torque_polynomials_per_stance_phase_ = s.torque_polynomials_per_stance_phase;
  p.force_limit_in_normal_direction_ = s.force_limit_in_normal_direction;
  p.torque_tx_min_ = s.torque_tx_min; p.torque_tx_max_ = s.torque_tx_max;
  p.torque_ty_min_ = s.torque_ty_min; p.torque_ty_max_ = s.torque_ty_max;
  p.torque_k_friction_ = s.torque_k_friction;
  p.bound_phase_duration_ = {s.bound_phase_duration[0], s.bound_phase_duration[1]};
  p.dt_cost_energy_ = s.dt_cost_energy;
  p.energy_cost_torque_weight_ = s.energy_cost_torque_weight;
  p.dt_cost_ang_mom_ = s.dt_cost_ang_mom;
  p.enable_swing_ee_base_pos_tracking = s.enable_swing_ee_base_pos_tracking != 0;
  p.swing_ee_base_pos_tracking_weight_ = s.swing_ee_base_pos_tracking_weight;
  p.dt_cost_swing_ee_base_pos_tracking_ = s.dt_cost_swing_ee_base_pos_tracking;
  p.ee_linear_constraints_.clear();
  for (int i = 0; i < s.n_ee_linear && i < ORC_MAX_EE_LINEAR; ++i) {
    Parameters::EELinearConstraintDef def;
    def.target = s.ee_linear_target[i] ? Parameters::EELinearConstraintDef::Angle : Parameters::EELinearConstraintDef::Motion;
    def.deriv = s.ee_linear_deriv[i] ? kVel : kPos;
    for (int j = 0; j < s.ee_linear_n_terms[i] && j < ORC_MAX_EE_LINEAR_TERMS; ++j)
      def.terms.push_back({s.ee_linear_term_ee[i][j], s.ee_linear_term_dim[i][j], s.ee_linear_term_coeff[i][j]});
    def.tolerance = s.ee_linear_tolerance[i];
    def.dt = s.ee_linear_dt[i];
    p.ee_linear_constraints_.push_back(def);
  }
  p.enable_stance_tracking = s.enable_stance_tracking != 0;   // orc_spec_defaults: off
  p.ee_stance_position_.clear();
  for (int e = 0; e < E; ++e) {
    std::vector<std::vector<double>> pos;
    for (int k = 0; k < s.n_stance_positions[e] && k < ORC_MAX_STANCES; ++k)
      pos.push_back({s.ee_stance_position[e][k][0], s.ee_stance_position[e][k][1]});
    p.ee_stance_position_.push_back(pos);
  }
  auto dims = [](const int* flags) { std::vector<int> d; for (int i = 0; i < 3; ++i) if (flags[i]) d.push_back(i); return d; };
  p.bounds_final_lin_pos_ = dims(s.bounds_final_lin_pos);
  p.bounds_final_lin_vel_ = dims(s.bounds_final_lin_vel);
  p.bounds_final_ang_pos_ = dims(s.bounds_final_ang_pos);
  p.bounds_final_ang_vel_ = dims(s.bounds_final_ang_vel);

  auto node = [](const double* a) { Node n; n.p = V3(a[0], a[1], a[2]); n.v = V3(a[3], a[4], a[5]); return n; };
  f.initial_base_.lin = node(s.initial_base_lin);
  f.initial_base_.ang = node(s.initial_base_ang);
  f.final_base_.lin = node(s.final_base_lin);
  f.final_base_.ang = node(s.final_base_ang);
  for (int e = 0; e < E; ++e) f.initial_ee_W_.push_back(V3(s.initial_ee_W[e][0], s.initial_ee_W[e][1], s.initial_ee_W[e][2]));
  return f;
}

static void Build(orc_problem* p, const orc_spec& s) {
  p->formulation = MakeFormulation(s);
  for (int c = 0; c < s.n_costs && c < ORC_MAX_COSTS; ++c) p->costs.emplace_back(s.cost_ids[c], s.cost_weights[c]);
  for (int c = 0; c < s.n_node_costs && c < ORC_MAX_NODE_COSTS; ++c) {
    if (s.node_cost_set[c] < 0 || s.node_cost_set[c] > 5 || s.node_cost_deriv[c] < 0 || s.node_cost_deriv[c] > 1 ||
        s.node_cost_dim[c] < 0 || s.node_cost_dim[c] > 2 || (s.node_cost_set[c] >= 2 && (s.node_cost_ee[c] < 0 || s.node_cost_ee[c] >= s.n_ee)))
      throw std::runtime_error("bad NodeCost term");
    p->node_costs.push_back({s.node_cost_set[c], s.node_cost_ee[c], s.node_cost_deriv[c], s.node_cost_dim[c], s.node_cost_weight[c]});
  }
  for (auto& v : p->formulation.GetVariableSets(p->splines)) p->nlp.AddVariableSet(v);
  for (auto& c : p->formulation.GetConstraints(p->splines)) p->nlp.AddConstraintSet(c);
  p->x0 = p->nlp.GetVariableValues();
  // GetCosts (nlp_formulation.cc:612-626): r_B0 = b_R_w0 (initial_ee_W - initial_base.lin.p)
  const NlpFormulation& f = p->formulation;
  if (f.params_.enable_swing_ee_base_pos_tracking && f.params_.swing_ee_base_pos_tracking_weight_ > 0.0) {
    const M3 b_R_w0 = EulerConverter::GetRotationMatrixBaseToWorldDense(f.initial_base_.ang.p).transpose();
    for (int ee = 0; ee < f.params_.GetEECount(); ++ee) p->swing_ref_B.push_back(b_R_w0 * (f.initial_ee_W_.at(ee) - f.initial_base_.lin.p));
  }
}

extern "C" {

void orc_spec_defaults(orc_spec* s) {
  std::memset(s, 0, sizeof(*s));
  Parameters d;
  s->n_costs = 0;   // parameters.cc:90: no cost by default
  s->n_constraints = static_cast<int>(d.constraints_.size());
  for (int i = 0; i < s->n_constraints; ++i) s->constraints[i] = d.constraints_[i];
  s->dt_constraint_dynamic = d.dt_constraint_dynamic_;
  s->dt_constraint_range_of_motion = d.dt_constraint_range_of_motion_;
  s->dt_constraint_base_motion = d.dt_constraint_base_motion_;
  s->dt_constraint_force = d.dt_constraint_force_;
  s->dt_constraint_torque = d.dt_constraint_torque_;
  s->duration_base_polynomial = d.duration_base_polynomial_;
  s->ee_polynomials_per_swing_phase = d.ee_polynomials_per_swing_phase_;
  s->force_polynomials_per_stance_phase = d.force_polynomials_per_stance_phase_;
  s->torque_polynomials_per_stance_phase = d.torque_polynomials_per_stance_phase_;
  s->force_limit_in_normal_direction = d.force_limit_in_normal_direction_;
  s->torque_tx_min = d.torque_tx_min_; s->torque_tx_max = d.torque_tx_max_;
  s->torque_ty_min = d.torque_ty_min_; s->torque_ty_max = d.torque_ty_max_;
  s->torque_k_friction = d.torque_k_friction_;
  s->bound_phase_duration[0] = d.bound_phase_duration_.first;
  s->bound_phase_duration[1] = d.bound_phase_duration_.second;
  s->dt_cost_energy = d.dt_cost_energy_;
  s->energy_cost_torque_weight = d.energy_cost_torque_weight_;
  s->dt_cost_ang_mom = d.dt_cost_ang_mom_;
  s->enable_swing_ee_base_pos_tracking = d.enable_swing_ee_base_pos_tracking ? 1 : 0;
  s->swing_ee_base_pos_tracking_weight = d.swing_ee_base_pos_tracking_weight_;
  s->dt_cost_swing_ee_base_pos_tracking = d.dt_cost_swing_ee_base_pos_tracking_;
  for (int i = 0; i < 3; ++i) s->bounds_final_lin_pos[i] = s->bounds_final_lin_vel[i] = s->bounds_final_ang_pos[i] = s->bounds_final_ang_vel[i] = 1;
}

int orc_spec_robot(orc_spec* s, int robot) {
  try {
    RobotModel rm = RobotModel::Make(static_cast<RobotModel::Robot>(robot));
    static const double inertia[4][7] = {   // models/examples/*_model.h
        {20, 1.2, 5.5, 6.0, 0.0, -0.2, -0.01},
        {20, 1.209, 5.583, 6.056, 0.005, -0.190, -0.012},
        {83, 4.26, 8.97, 9.88, -0.0063, 0.193, 0.0126},
        {29.5, 0.946438, 1.94478, 2.01835, 0.000938112, -0.00595386, -0.00146328}};
    s->mass = inertia[robot][0];
    for (int i = 0; i < 6; ++i) s->inertia[i] = inertia[robot][1 + i];
    s->n_ee = static_cast<int>(rm.kinematic_model_->nominal_stance_.size());
    for (int e = 0; e < s->n_ee; ++e)
      for (int d = 0; d < 3; ++d) {
        s->nominal_stance[e][d] = rm.kinematic_model_->nominal_stance_[e](d);
        s->max_dev[e][d] = rm.kinematic_model_->max_dev_[e](d);
        s->min_dev[e][d] = rm.kinematic_model_->min_dev_[e](d);
      }
    return 0;
  } catch (const std::exception& e) { g_err = e.what(); return -1; }
}

int orc_spec_gait(orc_spec* s, int combo, double T) {
  try {
    auto gen = GaitGenerator::MakeGaitGenerator(s->n_ee);
    gen->SetCombo(static_cast<GaitGenerator::Combos>(combo));
    for (int e = 0; e < s->n_ee; ++e) {
      Vec d = gen->GetPhaseDurations(T, e);
      if (d.size() > ORC_MAX_PHASES) throw std::runtime_error("too many phases");
      s->n_phases[e] = static_cast<int>(d.size());
      for (size_t i = 0; i < d.size(); ++i) s->phase_durations[e][i] = d[i];
      s->in_contact_at_start[e] = gen->IsInContactAtStart(e) ? 1 : 0;
    }
    return 0;
  } catch (const std::exception& e) { g_err = e.what(); return -1; }
}

void orc_spec_nominal_states(orc_spec* s, double goal_x, double goal_y, double goal_yaw) {
  double z_ground = 0.0;
  for (int e = 0; e < s->n_ee; ++e) {
    s->initial_ee_W[e][0] = s->nominal_stance[e][0];
    s->initial_ee_W[e][1] = s->nominal_stance[e][1];
    s->initial_ee_W[e][2] = z_ground;
  }
  for (int i = 0; i < 6; ++i) s->initial_base_lin[i] = s->initial_base_ang[i] = s->final_base_lin[i] = s->final_base_ang[i] = 0.0;
  s->initial_base_lin[2] = -s->nominal_stance[0][2] + z_ground;
  s->final_base_lin[0] = goal_x;
  s->final_base_lin[1] = goal_y;
  s->final_base_lin[2] = s->initial_base_lin[2];
  s->final_base_ang[2] = goal_yaw;
}

orc_problem* orc_create(const orc_spec* s) {
  orc_problem* p = nullptr;
  try {
    p = new orc_problem();
    Build(p, *s);
    return p;
  } catch (const std::exception& e) {
    g_err = e.what();
    delete p;
    return nullptr;
  }
}
void orc_destroy(orc_problem* p) { delete p; }
const char* orc_last_error(void) { return g_err.c_str(); }

int orc_n_vars(const orc_problem* p) { return p->nlp.GetNumberOfOptimizationVariables(); }
int orc_n_rows(const orc_problem* p) { return p->nlp.GetNumberOfConstraints(); }
int orc_n_var_sets(const orc_problem* p) { return static_cast<int>(p->nlp.Vars()->GetComponents().size()); }
int orc_n_con_sets(const orc_problem* p) { return static_cast<int>(p->nlp.Constraints().size()); }
const char* orc_var_set_name(const orc_problem* p, int i) { return p->nlp.Vars()->GetComponents().at(i)->GetName().c_str(); }
int orc_var_set_rows(const orc_problem* p, int i) { return p->nlp.Vars()->GetComponents().at(i)->GetRows(); }
const char* orc_con_set_name(const orc_problem* p, int i) { return p->nlp.Constraints().at(i)->GetName().c_str(); }
int orc_con_set_rows(const orc_problem* p, int i) { return p->nlp.Constraints().at(i)->GetRows(); }

void orc_x0(const orc_problem* p, double* x) { std::memcpy(x, p->x0.data(), p->x0.size() * sizeof(double)); }
void orc_var_bounds(const orc_problem* p, double* lower, double* upper) {
  VecBound b = p->nlp.GetBoundsOnOptimizationVariables();
  for (size_t i = 0; i < b.size(); ++i) { lower[i] = b[i].lower; upper[i] = b[i].upper; }
}
void orc_con_bounds(const orc_problem* p, double* lower, double* upper) {
  VecBound b = p->nlp.GetBoundsOnConstraints();
  for (size_t i = 0; i < b.size(); ++i) { lower[i] = b[i].lower; upper[i] = b[i].upper; }
}
void orc_eval_g(orc_problem* p, const double* x, double* g) {
  Vec v = p->nlp.EvaluateConstraints(x);
  std::memcpy(g, v.data(), v.size() * sizeof(double));
}
int orc_eval_jac(orc_problem* p, const double* x, int* rows, int* cols, double* vals) {
  std::vector<int> r, c;
  std::vector<double> v;
  Vec cur;
  if (!x) { cur = p->nlp.GetVariableValues(); x = cur.data(); }
  p->nlp.EvalJacobian(x, &r, &c, &v);
  if (rows) std::memcpy(rows, r.data(), r.size() * sizeof(int));
  if (cols) std::memcpy(cols, c.data(), c.size() * sizeof(int));
  if (vals) std::memcpy(vals, v.data(), v.size() * sizeof(double));
  return static_cast<int>(v.size());
}
int orc_nnz(orc_problem* p, const double* x) { return orc_eval_jac(p, x, nullptr, nullptr, nullptr); }

}  // extern "C"

extern "C" {

int orc_batch_eval(const orc_spec* s, int n, const double* x, double* g, double* vals, int n_threads) {
  int status = 0;
  int nv = 0, nr = 0, nnz = 0;
  {
    orc_problem probe;
    Build(&probe, *s);
    nv = probe.nlp.GetNumberOfOptimizationVariables();
    nr = probe.nlp.GetNumberOfConstraints();
    nnz = orc_nnz(&probe, nullptr);
  }
#ifdef _OPENMP
  if (n_threads > 0) omp_set_num_threads(n_threads);
#endif
#pragma omp parallel
  {
    orc_problem local;              // one object graph per thread (instances are not thread-safe)
    Build(&local, *s);
    std::vector<double> v;
#pragma omp for schedule(dynamic)
    for (int i = 0; i < n; ++i) {
      const double* xi = x + static_cast<size_t>(i) * nv;
      if (g) {
        Vec gi = local.nlp.EvaluateConstraints(xi);
        std::memcpy(g + static_cast<size_t>(i) * nr, gi.data(), nr * sizeof(double));
      }
      if (vals) {
        local.nlp.EvalJacobian(xi, nullptr, nullptr, &v);
        if (static_cast<int>(v.size()) != nnz) {
#pragma omp atomic write
          status = -1;
        } else {
          std::memcpy(vals + static_cast<size_t>(i) * nnz, v.data(), nnz * sizeof(double));
        }
      }
    }
  }
  return status;
}

// TEST SWITCH: 0 = std::pow everywhere (the reference, default); 1 = PhaseSpline polynomials use the correctly rounded
// integer powers the device computes at run time.  Global, not thread-safe: set it before evaluating.
void orc_set_pow_mode(int mode) { SetPowMode(mode); }
int orc_get_pow_mode(void) { return GetPowMode(); }
// x^n (n = 2..4) three ways: std::pow, the device's compensated product (ExactPow), and the __float128 product rounded
// to double (correct rounding up to a 2^-113 relative error before the final rounding)
void orc_pow_compare(double x, int n, double* std_pow, double* exact_pow, double* quad_rounded) {
  *std_pow = std::pow(x, n);
  *exact_pow = ExactPow(x, n);
  __float128 q = 1;
  for (int i = 0; i < n; ++i) q *= static_cast<__float128>(x);
  *quad_rounded = static_cast<double>(q);
}

// Debug hook: state of one spline at global time t after SetVariables(x).  kind: 0 ee-motion, 1 ee-force, 2 ee-torque,
// 3 base-lin, 4 base-ang.  out[0] polynomial id, out[1] local time, out[2] polynomial duration, out[3..5] pos,
// out[6..8] vel, out[9..11] d pos / d phase durations column 0 (PhaseSpline only, else 0), out[12] number of polynomials
void orc_debug_spline(orc_problem* p, const double* x, int kind, int ee, double t, double* out) {
  p->nlp.SetVariables(x);
  std::shared_ptr<NodeSpline> sp = kind == 0 ? p->splines.ee_motion_.at(ee) : kind == 1 ? p->splines.ee_force_.at(ee)
                                   : kind == 2 ? p->splines.ee_torque_.at(ee) : kind == 3 ? p->splines.base_linear_
                                                                                          : p->splines.base_angular_;
  const Vec d = sp->GetPolyDurations();
  auto lt = sp->GetLocalTime(t, d);
  out[0] = lt.first; out[1] = lt.second; out[2] = d.at(lt.first);
  State3 pt = sp->GetPoint(t);
  for (int i = 0; i < 3; ++i) { out[3 + i] = pt.p(i); out[6 + i] = pt.v(i); out[9 + i] = 0.0; }
  SpMat J = sp->GetJacobianOfPosWrtDurations(t);
  for (int r = 0; r < 3 && r < static_cast<int>(J.r.size()); ++r)
    for (auto& en : J.r[r]) if (en.first == 0) out[9 + r] = en.second;
  out[12] = static_cast<double>(d.size());
}
// node values (p0, v0, p1, v1: 12 doubles) of polynomial `poly` of the same spline
void orc_debug_poly_nodes(orc_problem* p, int kind, int ee, int poly, double* out12) {
  std::shared_ptr<NodesVariables> nv;
  for (auto& v : p->nlp.Vars()->GetComponents()) {
    const std::string want = kind == 0 ? "ee-motion_" + std::to_string(ee) : kind == 1 ? "ee-force_" + std::to_string(ee)
                             : kind == 2 ? "ee-torque_" + std::to_string(ee) : kind == 3 ? "base-lin" : "base-ang";
    if (v->GetName() == want) nv = std::dynamic_pointer_cast<NodesVariables>(v);
  }
  const auto& nodes = nv->GetNodes();
  for (int i = 0; i < 3; ++i) {
    out12[i] = nodes.at(poly).p(i); out12[3 + i] = nodes.at(poly).v(i);
    out12[6 + i] = nodes.at(poly + 1).p(i); out12[9 + i] = nodes.at(poly + 1).v(i);
  }
}

// Persistent pool for the timed CPU baseline: one object graph per thread built ONCE (set-up, as a solver thread would
// hold its own NlpFormulation), so that the timed region holds evaluations only.
struct orc_pool {
  std::vector<std::unique_ptr<orc_problem>> per_thread;
  int nv = 0, nr = 0, nnz = 0;
};

orc_pool* orc_pool_create(const orc_spec* s, int n_threads) {
  try {
    auto pool = std::make_unique<orc_pool>();
    if (n_threads < 1) n_threads = 1;
    for (int t = 0; t < n_threads; ++t) {
      pool->per_thread.emplace_back(new orc_problem);
      Build(pool->per_thread.back().get(), *s);
    }
    orc_problem* p0 = pool->per_thread[0].get();
    pool->nv = p0->nlp.GetNumberOfOptimizationVariables();
    pool->nr = p0->nlp.GetNumberOfConstraints();
    pool->nnz = orc_nnz(p0, nullptr);
    return pool.release();
  } catch (const std::exception& e) {
    g_err = e.what();
    return nullptr;
  }
}
void orc_pool_destroy(orc_pool* p) { delete p; }
int orc_pool_threads(const orc_pool* p) { return static_cast<int>(p->per_thread.size()); }
int orc_pool_nnz(const orc_pool* p) { return p->nnz; }

int orc_pool_eval(orc_pool* pool, int n, const double* x, double* g, double* vals) {
  int status = 0;
  const int nv = pool->nv, nr = pool->nr, nnz = pool->nnz;
  const int nt = static_cast<int>(pool->per_thread.size());
#pragma omp parallel num_threads(nt)
  {
    int me = 0;
#ifdef _OPENMP
    me = omp_get_thread_num();
#endif
    orc_problem& local = *pool->per_thread[me % nt];
    std::vector<double> v;
#pragma omp for schedule(dynamic)
    for (int i = 0; i < n; ++i) {
      const double* xi = x + static_cast<size_t>(i) * nv;
      if (g) {
        Vec gi = local.nlp.EvaluateConstraints(xi);
        std::memcpy(g + static_cast<size_t>(i) * nr, gi.data(), nr * sizeof(double));
      }
      if (vals) {
        local.nlp.EvalJacobian(xi, nullptr, nullptr, &v);
        if (static_cast<int>(v.size()) != nnz) {
#pragma omp atomic write
          status = -1;
        } else {
          std::memcpy(vals + static_cast<size_t>(i) * nnz, v.data(), nnz * sizeof(double));
        }
      }
    }
  }
  return status;
}

static Node MakeNode(const double* p, const double* v) { Node n; n.p = V3(p[0], p[1], p[2]); n.v = V3(v[0], v[1], v[2]); return n; }

void orc_hermite_point(const double* p0, const double* v0, const double* p1, const double* v1, double T, double t,
                       double* pos, double* vel, double* acc) {
  CubicHermitePolynomial poly;
  poly.SetNodes(MakeNode(p0, v0), MakeNode(p1, v1));
  poly.SetDuration(T);
  poly.UpdateCoeff();
  State3 s = poly.GetPoint(t);
  for (int d = 0; d < 3; ++d) { pos[d] = s.p(d); vel[d] = s.v(d); acc[d] = s.a(d); }
}
void orc_hermite_basis(double T, double t, double* out12) {
  CubicHermitePolynomial poly;
  poly.SetDuration(T);
  int k = 0;
  for (Dx dxdt : {kPos, kVel, kAcc}) {
    out12[k++] = poly.GetDerivativeWrtStartNode(dxdt, kPos, t);
    out12[k++] = poly.GetDerivativeWrtStartNode(dxdt, kVel, t);
    out12[k++] = poly.GetDerivativeWrtEndNode(dxdt, kPos, t);
    out12[k++] = poly.GetDerivativeWrtEndNode(dxdt, kVel, t);
  }
}
void orc_hermite_dpos_dT(const double* p0, const double* v0, const double* p1, const double* v1, double T, double t, double* out3) {
  CubicHermitePolynomial poly;
  poly.SetNodes(MakeNode(p0, v0), MakeNode(p1, v1));
  poly.SetDuration(T);
  poly.UpdateCoeff();
  V3 d = poly.GetDerivativeOfPosWrtDuration(t);
  for (int i = 0; i < 3; ++i) out3[i] = d(i);
}
void orc_terrain(int terrain_id, double x, double y, double* height, double* dh, double* basis, double* dbasis) {
  auto t = HeightMap::MakeTerrain(terrain_id);
  *height = t->GetHeight(x, y);
  dh[0] = t->GetDerivativeOfHeightWrt(0, x, y);
  dh[1] = t->GetDerivativeOfHeightWrt(1, x, y);
  HeightMap::Direction dirs[3] = {HeightMap::Normal, HeightMap::Tangent1, HeightMap::Tangent2};
  for (int b = 0; b < 3; ++b) {
    V3 v = t->GetNormalizedBasis(dirs[b], x, y);
    for (int i = 0; i < 3; ++i) basis[b * 3 + i] = v(i);
    for (int dim = 0; dim < 2; ++dim) {
      V3 dv = t->GetDerivativeOfNormalizedBasisWrt(dirs[b], dim, x, y);
      for (int i = 0; i < 3; ++i) dbasis[(b * 2 + dim) * 3 + i] = dv(i);
    }
  }
}
void orc_sample_times(double T, double dt, double* out, int* n) {
  struct Probe : TimeDiscretizationConstraint {
    Probe(double T_, double dt_) : TimeDiscretizationConstraint(T_, dt_, "probe") {}
    void UpdateConstraintAtInstance(double, int, Vec&) const override {}
    void UpdateBoundsAtInstance(double, int, VecBound&) const override {}
    void UpdateJacobianAtInstance(double, int, const std::string&, SpMat&) const override {}
  } probe(T, dt);
  *n = static_cast<int>(probe.SampleTimes().size());
  if (out) for (int i = 0; i < *n; ++i) out[i] = probe.SampleTimes()[i];
}

// ---- costs: NlpFormulation::GetCosts / GetCost / MakeForcesCost / MakeEEMotionCost (nlp_formulation.cc:605-681)
}  // extern "C"
namespace {
struct NodeCostDef { std::string nodes_id; Dx deriv; int dim; double weight; };   // costs/node_cost.cc:36-45
std::vector<NodeCostDef> CostTerms(const orc_problem* p) {
  std::vector<NodeCostDef> terms;
  const int E = static_cast<int>(p->splines.ee_motion_.size());
  for (const auto& pr : p->costs) {
    const double weight = pr.second;
    if (pr.first == 0) {            // ForcesCostID, :648-666
      for (int ee = 0; ee < E; ee++)
        for (int dim : {0, 1, 2}) {
          terms.push_back({"ee-force_" + std::to_string(ee), kPos, dim, weight});
          terms.push_back({"ee-torque_" + std::to_string(ee), kPos, dim, weight});
          terms.push_back({"ee-force_" + std::to_string(ee), kVel, dim, 0.1 * weight});
          terms.push_back({"ee-torque_" + std::to_string(ee), kVel, dim, 0.1 * weight});
        }
    } else if (pr.first == 1) {     // EEMotionCostID, :668-681
      for (int ee = 0; ee < E; ee++) {
        terms.push_back({"ee-motion_" + std::to_string(ee), kVel, 0, weight});
        terms.push_back({"ee-motion_" + std::to_string(ee), kVel, 1, weight});
        terms.push_back({"ee-motion_" + std::to_string(ee), kVel, 2, 0.5 * weight});
      }
    } else if (pr.first == 2 || pr.first == 3) {
      // EnergyCostID / AngMomCostID (:636-643): time-discretised terms, summed by SampledCost* below
    } else {
      throw std::runtime_error("cost not defined!");     // :644
    }
  }
  static const char* const kSetNames[6] = {"base-lin", "base-ang", "ee-motion_", "ee-ang_", "ee-force_", "ee-torque_"};
  for (const auto& nc : p->node_costs) {          // nlp.AddCostSet(std::make_shared<NodeCost>(...)), squat_jump_example.cc:205-211
    const std::string id = nc.set < 2 ? std::string(kSetNames[nc.set]) : kSetNames[nc.set] + std::to_string(nc.ee);
    terms.push_back({id, nc.deriv == 0 ? kPos : kVel, nc.dim, nc.weight});
  }
  return terms;
}
}  // namespace
extern "C" {

// ifopt::Problem::EvaluateCostFunction: sum of the cost terms' values.  NodeCost::GetCost (node_cost.cc:53-62); the
// reference leaves `cost` uninitialised there (:55), restated with the intended 0.
double orc_eval_f(orc_problem* p, const double* x) {
  p->nlp.SetVariables(x);
  double total = 0.0;
  for (const NodeCostDef& t : CostTerms(p)) {
    auto nodes = p->nlp.Vars()->GetComponent<NodesVariables>(t.nodes_id);
    double cost = 0.0;
    for (const Node& n : nodes->GetNodes()) {
      double val = n.at(t.deriv)(t.dim);
      cost += t.weight * std::pow(val, 2);
    }
    total += cost;
  }
  const Parameters& pa = p->formulation.params_;
  for (const auto& pr : p->costs) {
    if (pr.first == 2) total += EnergyCostValue(p->splines, pr.second, pa.energy_cost_torque_weight_, pa.dt_cost_energy_);
    if (pr.first == 3) total += AngMomCostValue(p->splines, *p->formulation.model_.dynamic_model_, pr.second, pa.dt_cost_ang_mom_);
  }
  for (size_t ee = 0; ee < p->swing_ref_B.size(); ++ee)     // GetCosts, nlp_formulation.cc:612-626: after the listed costs
    total += EEBasePosCostValue(p->splines, static_cast<int>(ee), p->swing_ref_B[ee], pa.swing_ee_base_pos_tracking_weight_,
                                pa.dt_cost_swing_ee_base_pos_tracking_);
  return total;
}

// ifopt::Problem::EvaluateCostFunctionGradient: NodeCost::FillJacobianBlock (node_cost.cc:64-76) of every term, summed
void orc_eval_grad_f(orc_problem* p, const double* x, double* grad) {
  p->nlp.SetVariables(x);
  const int n = p->nlp.GetNumberOfOptimizationVariables();
  for (int i = 0; i < n; ++i) grad[i] = 0.0;
  for (const NodeCostDef& t : CostTerms(p)) {
    int col0 = 0;
    for (auto& vs : p->nlp.Vars()->GetComponents()) {
      if (vs->GetName() == t.nodes_id) {
        auto nodes = std::dynamic_pointer_cast<NodesVariables>(vs);
        for (int i = 0; i < nodes->GetRows(); ++i)
          for (const auto& nvi : nodes->GetNodeValuesInfo(i))
            if (nvi.deriv == t.deriv && nvi.dim == t.dim) {
              double val = nodes->GetNodes().at(nvi.id).at(t.deriv)(t.dim);
              grad[col0 + i] += t.weight * 2.0 * val;
            }
      }
      col0 += vs->GetRows();
    }
  }
  const Parameters& pa = p->formulation.params_;
  for (const auto& pr : p->costs) {
    if (pr.first != 2 && pr.first != 3) continue;
    int col0 = 0;
    for (auto& vs : p->nlp.Vars()->GetComponents()) {     // ifopt::Composite::GetJacobian: one block per variable set
      if (pr.first == 2)
        EnergyCostFillBlock(p->splines, pr.second, pa.energy_cost_torque_weight_, pa.dt_cost_energy_, vs->GetName(), grad + col0);
      else
        AngMomCostFillBlock(p->splines, *p->formulation.model_.dynamic_model_, pr.second, pa.dt_cost_ang_mom_, vs->GetName(),
                            grad + col0);
      col0 += vs->GetRows();
    }
  }
  for (size_t ee = 0; ee < p->swing_ref_B.size(); ++ee) {
    int col0 = 0;
    for (auto& vs : p->nlp.Vars()->GetComponents()) {
      EEBasePosCostFillBlock(p->splines, static_cast<int>(ee), p->swing_ref_B[ee], pa.swing_ee_base_pos_tracking_weight_,
                             pa.dt_cost_swing_ee_base_pos_tracking_, vs->GetName(), grad + col0);
      col0 += vs->GetRows();
    }
  }
}

int orc_trajectory_width(const orc_problem* p) { return 19 + 25 * static_cast<int>(p->splines.ee_motion_.size()); }

// utils/save_data.cpp:54-119, column order :68-112
int orc_sample_trajectory(orc_problem* p, const double* x, double dt, double* rows) {
  p->nlp.SetVariables(x);
  const SplineHolder& sol = p->splines;
  Vec durs = sol.base_linear_->GetPolyDurations();     // Spline::GetTotalTime (spline.cc:118-123)
  const double T_total = std::accumulate(durs.begin(), durs.end(), 0.0);
  const int E = static_cast<int>(sol.ee_motion_.size()), W = 19 + 25 * E;
  int n = 0;
  double t = 0.0;
  while (t <= T_total + 1e-9) {
    if (rows) {
      double* r = rows + static_cast<size_t>(n) * W;
      auto put9 = [&](double* o, const State3& st) { for (int d = 0; d < 3; ++d) { o[d] = st.p(d); o[3 + d] = st.v(d); o[6 + d] = st.a(d); } };
      r[0] = t;
      put9(r + 1, sol.base_linear_->GetPoint(t));
      put9(r + 10, sol.base_angular_->GetPoint(t));
      for (int i = 0; i < E; ++i) {
        double* o = r + 19 + 25 * i;
        put9(o, sol.ee_motion_.at(i)->GetPoint(t));
        put9(o + 9, sol.ee_ang_.at(i)->GetPoint(t));
        const State3 f = sol.ee_force_.at(i)->GetPoint(t), tq = sol.ee_torque_.at(i)->GetPoint(t);
        for (int d = 0; d < 3; ++d) { o[18 + d] = f.p(d); o[21 + d] = tq.p(d); }
        o[24] = sol.phase_durations_.at(i)->IsContactPhase(t) ? 1.0 : 0.0;
      }
    }
    t += dt;
    ++n;
  }
  return n;
}

}  // extern "C"
