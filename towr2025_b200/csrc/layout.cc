// layout.cc — variable sets, node maps and spline timing of one structure class (host).
//
// Mirrors, as flat tables, what the reference builds as object graphs:
//   NodesVariablesAll::GetNodeValuesInfo           towr/src/variables/nodes_variables_all.cc:45-61
//   NodesVariablesPhaseBased (BuildPolyInfos, maps) towr/src/variables/nodes_variables_phase_based.cc:39-59,215-396
//   Parameters::GetBasePolyDurations               towr/src/parameters.cc:114-130
//   column order of the NLP                        towr/src/nlp_formulation.cc:81-101
#include "structure.h"
#include "terrain.h"

#include <cmath>
#include <numeric>
#include <stdexcept>

namespace tb {

bool VarSet::IsConstantNode(int node) const {
  // a node is constant when a polynomial next to it belongs to a constant phase
  // (nodes_variables_phase_based.cc:101-113,167-183)
  if (poly_const.empty()) return false;
  const int last = n_nodes - 1;
  if (node > 0 && poly_const[node - 1]) return true;
  if (node < last && poly_const[node]) return true;
  return false;
}

std::vector<int> VarSet::NonConstantNodes() const {
  std::vector<int> ids;
  for (int n = 0; n < n_nodes; ++n)
    if (!IsConstantNode(n)) ids.push_back(n);
  return ids;
}

int VarSet::PolyAtStartOfPhase(int phase) const {
  for (int p = 0; p < n_polys(); ++p)
    if (poly_phase[p] == phase) return p;
  throw std::runtime_error("phase not present in spline " + name);
}

int VarSet::PhaseOfNode(int node) const {
  // first adjacent polynomial decides (nodes_variables_phase_based.cc:133-140)
  const int last = n_nodes - 1;
  int poly = node == 0 ? 0 : (node == last ? last - 1 : node - 1);
  return poly_phase[poly];
}

int VarSet::OptIndex(int node, int deriv, int dim) const {
  for (int idx = 0; idx < n_vars; ++idx)
    for (const NodeRef& r : refs[idx])
      if (r.node == node && r.deriv == deriv && r.dim == dim) return idx;
  return -1;
}

const VarSet* HostProgram::Find(int kind, int ee) const {
  for (const VarSet& v : var_sets)
    if (v.kind == kind && (ee < 0 || v.ee == ee)) return &v;
  return nullptr;
}

// every node value is a variable: idx -> (node idx/6, pos|vel, dim)
static VarSet MakeDenseNodeSet(const std::string& name, int kind, int n_nodes) {
  VarSet v;
  v.name = name;
  v.kind = kind;
  v.n_nodes = n_nodes;
  v.n_vars = n_nodes * 6;
  v.refs.resize(v.n_vars);
  for (int idx = 0; idx < v.n_vars; ++idx) {
    int in_node = idx % 6;
    v.refs[idx].push_back(NodeRef{idx / 6, in_node < 3 ? 0 : 1, in_node % 3});
  }
  return v;
}

// phase-based parameterisation.  changing_in_contact: true for force/torque (stance phases carry the
// polynomials), false for motion/orientation (swing phases do).
static VarSet MakePhaseNodeSet(const std::string& name, int kind, int ee, int n_phases, bool contact_at_start,
                               int polys_per_changing_phase) {
  VarSet v;
  v.name = name;
  v.kind = kind;
  v.ee = ee;
  const bool motion_like = (kind == kEEMotion || kind == kEEAng);
  bool constant = motion_like ? contact_at_start : !contact_at_start;
  for (int ph = 0; ph < n_phases; ++ph) {
    int n = constant ? 1 : polys_per_changing_phase;
    for (int j = 0; j < n; ++j) {
      v.poly_phase.push_back(ph);
      v.poly_in_phase.push_back(j);
      v.polys_in_phase.push_back(n);
      v.poly_const.push_back(constant ? 1 : 0);
    }
    constant = !constant;
  }
  v.n_nodes = static_cast<int>(v.poly_phase.size()) + 1;

  if (motion_like) {
    // swing node: px,vx,py,vy,pz(,vz) ; a stance phase shares one position triple between its two nodes
    const bool pin_vz = (kind == kEEMotion);
    for (int node = 0; node < v.n_nodes; ++node) {
      if (!v.IsConstantNode(node)) {
        for (int dim = 0; dim < 3; ++dim) {
          v.refs.push_back({NodeRef{node, 0, dim}});
          if (!(pin_vz && dim == 2)) v.refs.push_back({NodeRef{node, 1, dim}});
        }
      } else {
        for (int dim = 0; dim < 3; ++dim) v.refs.push_back({NodeRef{node, 0, dim}, NodeRef{node + 1, 0, dim}});
        node += 1;
      }
    }
  } else {
    // stance node: p,v interleaved per dim ; nodes touching a swing phase stay zero
    for (int node = 0; node < v.n_nodes; ++node) {
      if (!v.IsConstantNode(node)) {
        for (int dim = 0; dim < 3; ++dim) {
          v.refs.push_back({NodeRef{node, 0, dim}});
          v.refs.push_back({NodeRef{node, 1, dim}});
        }
      } else {
        node += 1;
      }
    }
  }
  v.n_vars = static_cast<int>(v.refs.size());
  return v;
}

double TotalTime(const towr_b200_spec& s) {
  // Parameters::GetTotalTime (parameters.cc:144-158): duration of foot 0's schedule
  double T = 0.0;
  for (int i = 0; i < s.n_phases[0]; ++i) T += s.phase_durations[0][i];
  return T;
}

std::vector<double> BasePolyDurations(const towr_b200_spec& s) {
  std::vector<double> d;
  const double dt = s.duration_base_polynomial;
  double left = TotalTime(s);
  while (left > 1e-10) {
    d.push_back(left > dt ? dt : left);
    left -= dt;
  }
  return d;
}

bool OptimizesDurations(const towr_b200_spec& s) {
  for (int i = 0; i < s.n_constraints; ++i)
    if (s.constraints[i] == TOWR_B200_C_TOTAL_TIME) return true;
  return false;
}

void ValidateSpec(const towr_b200_spec& s) {
  if (s.n_ee < 1 || s.n_ee > TOWR_B200_MAX_EE) throw std::runtime_error("n_ee must be in 1..4");
  if (s.n_constraints < 0 || s.n_constraints > TOWR_B200_MAX_CONSTRAINTS) throw std::runtime_error("bad n_constraints");
  for (int e = 0; e < s.n_ee; ++e) {
    if (s.n_phases[e] < 1 || s.n_phases[e] > TOWR_B200_MAX_PHASES) throw std::runtime_error("bad phase count");
    for (int i = 0; i < s.n_phases[e]; ++i)
      if (!(s.phase_durations[e][i] > 0.0)) throw std::runtime_error("phase durations must be positive");
  }
  const double T = TotalTime(s);
  for (int e = 1; e < s.n_ee; ++e) {
    double Te = 0.0;
    for (int i = 0; i < s.n_phases[e]; ++i) Te += s.phase_durations[e][i];
    if (std::fabs(Te - T) >= 1e-6) throw std::runtime_error("feet must share one total duration (parameters.cc:155)");
  }
  if (!(s.duration_base_polynomial > 0.0)) throw std::runtime_error("duration_base_polynomial must be positive");
  if (s.ee_polynomials_per_swing_phase < 1 || s.force_polynomials_per_stance_phase < 1 ||
      s.torque_polynomials_per_stance_phase < 1)
    throw std::runtime_error("polynomials per phase must be >= 1");
  if (s.terrain_id < 0 || s.terrain_id >= TOWR_B200_TERRAIN_COUNT) throw std::runtime_error("unknown terrain id");
  if (s.terrain_id == TOWR_B200_TERRAIN_TABLE) {
    if (s.terrain_table_n < 1 || s.terrain_table_n > TOWR_B200_MAX_TERRAIN_KNOTS) throw std::runtime_error("terrain table: 1..16 knots");
    if (s.terrain_table_mode != TOWR_B200_TABLE_STEPS && s.terrain_table_mode != TOWR_B200_TABLE_RAMPS)
      throw std::runtime_error("terrain table: mode must be TABLE_STEPS or TABLE_RAMPS");
    for (int i = 0; i < s.terrain_table_n; ++i) {
      if (!std::isfinite(s.terrain_table_x[i]) || !std::isfinite(s.terrain_table_h[i])) throw std::runtime_error("terrain table: non-finite knot");
      if (i > 0 && !(s.terrain_table_x[i] > s.terrain_table_x[i - 1])) throw std::runtime_error("terrain table: knots must increase in x");
    }
  }
  if (s.n_costs < 0 || s.n_costs > TOWR_B200_MAX_COSTS) throw std::runtime_error("bad n_costs");
  if (s.n_ee_linear < 0 || s.n_ee_linear > TOWR_B200_MAX_EE_LINEAR) throw std::runtime_error("bad n_ee_linear");
  for (int i = 0; i < s.n_ee_linear; ++i) {
    if (s.ee_linear_n_terms[i] < 1 || s.ee_linear_n_terms[i] > TOWR_B200_MAX_EE_LINEAR_TERMS) throw std::runtime_error("EELinearConstraint: 1..4 terms");
    if (s.ee_linear_deriv[i] < 0 || s.ee_linear_deriv[i] > 1) throw std::runtime_error("EELinearConstraint: deriv must be kPos or kVel");
    if (!(s.ee_linear_dt[i] > 0.0) || !std::isfinite(s.ee_linear_dt[i]) || !std::isfinite(s.ee_linear_tolerance[i]))
      throw std::runtime_error("EELinearConstraint: bad dt / tolerance");
    for (int j = 0; j < s.ee_linear_n_terms[i]; ++j)
      if (s.ee_linear_term_ee[i][j] < 0 || s.ee_linear_term_ee[i][j] >= s.n_ee || s.ee_linear_term_dim[i][j] < 0 ||
          s.ee_linear_term_dim[i][j] > 2 || !std::isfinite(s.ee_linear_term_coeff[i][j]))
        throw std::runtime_error("EELinearConstraint: bad term (foot, dimension or coefficient)");
    if (s.ee_linear_target[i] != 0 && s.ee_linear_target[i] != 1) throw std::runtime_error("EELinearConstraint: target must be Motion or Angle");
  }
  if (s.n_node_costs < 0 || s.n_node_costs > TOWR_B200_MAX_NODE_COSTS) throw std::runtime_error("bad n_node_costs");
  for (int c = 0; c < s.n_node_costs; ++c) {
    const int set = s.node_cost_set[c];
    if (set < TOWR_B200_NODES_BASE_LIN || set > TOWR_B200_NODES_EE_TORQUE || s.node_cost_deriv[c] < 0 || s.node_cost_deriv[c] > 1 ||
        s.node_cost_dim[c] < 0 || s.node_cost_dim[c] > 2 || !std::isfinite(s.node_cost_weight[c]) ||
        (set >= TOWR_B200_NODES_EE_MOTION && (s.node_cost_ee[c] < 0 || s.node_cost_ee[c] >= s.n_ee)))
      throw std::runtime_error("bad NodeCost term (set, foot, derivative, dimension or weight)");
  }
  if (s.enable_swing_ee_base_pos_tracking &&
      (!std::isfinite(s.swing_ee_base_pos_tracking_weight) || !std::isfinite(s.dt_cost_swing_ee_base_pos_tracking)))
    throw std::runtime_error("EEBasePosCost: non-finite weight / sampling period");
  for (int c = 0; c < s.n_costs; ++c) {
    if (s.cost_ids[c] < TOWR_B200_COST_FORCES || s.cost_ids[c] > TOWR_B200_COST_ANG_MOM)
      throw std::runtime_error("cost not defined!");   // nlp_formulation.cc:644
    if (s.cost_ids[c] == TOWR_B200_COST_ENERGY && (!std::isfinite(s.dt_cost_energy) || !std::isfinite(s.energy_cost_torque_weight)))
      throw std::runtime_error("EnergyCost: non-finite dt_cost_energy / energy_cost_torque_weight");
    if (s.cost_ids[c] == TOWR_B200_COST_ANG_MOM && !std::isfinite(s.dt_cost_ang_mom))
      throw std::runtime_error("AngularMomentumCost: non-finite dt_cost_ang_mom");
  }
}

void BuildLayout(const towr_b200_spec& s, HostProgram& hp) {
  ValidateSpec(s);
  hp.n_ee = s.n_ee;
  hp.total_time = TotalTime(s);
  hp.optimize_durations = OptimizesDurations(s);
  hp.curvature_dims = TerrainClassBits(s.terrain_id);   // curvature dims | kClassTableTerrain | kClassTerrainHard
  for (int i = 0; i < s.n_constraints; ++i)
    if (s.constraints[i] == TOWR_B200_C_TERRAIN_HARD || (s.constraints[i] == TOWR_B200_C_TORQUE && !(s.dt_constraint_torque > 0.0)))
      hp.curvature_dims |= kClassTerrainHard;
  if (s.n_ee_linear > 0) hp.curvature_dims |= kClassTerrainHard;   // ItemEELinear lives in the extended instantiations too
  const int E = s.n_ee;
  const std::vector<double> base_T = BasePolyDurations(s);
  const int n_base_nodes = static_cast<int>(base_T.size()) + 1;

  std::vector<VarSet>& vs = hp.var_sets;
  vs.clear();
  vs.push_back(MakeDenseNodeSet("base-lin", kBaseLin, n_base_nodes));
  vs.push_back(MakeDenseNodeSet("base-ang", kBaseAng, n_base_nodes));
  vs[0].poly_T = base_T;
  vs[1].poly_T = base_T;
  struct Fam { const char* prefix; int kind; int polys; };
  const Fam fams[4] = {{"ee-motion_", kEEMotion, s.ee_polynomials_per_swing_phase},
                       {"ee-ang_", kEEAng, s.ee_polynomials_per_swing_phase},
                       {"ee-force_", kEEForce, s.force_polynomials_per_stance_phase},
                       {"ee-torque_", kEETorque, s.torque_polynomials_per_stance_phase}};
  for (const Fam& f : fams)
    for (int e = 0; e < E; ++e) {
      VarSet v = MakePhaseNodeSet(f.prefix + std::to_string(e), f.kind, e, s.n_phases[e],
                                  s.in_contact_at_start[e] != 0, f.polys);
      for (int p = 0; p < v.n_polys(); ++p)   // ConvertPhaseToPolyDurations, nodes_variables_phase_based.cc:75-86
        v.poly_T.push_back(s.phase_durations[e][v.poly_phase[p]] / v.polys_in_phase[p]);
      vs.push_back(std::move(v));
    }
  if (hp.optimize_durations)
    for (int e = 0; e < E; ++e) {
      VarSet v;
      v.name = "ee-schedule" + std::to_string(e);
      v.kind = kEESchedule;
      v.ee = e;
      v.n_vars = s.n_phases[e] - 1;
      vs.push_back(std::move(v));
    }

  // columns, node slots, polynomial records.  The ee-ang splines feed no constraint on this path unless an
  // EELinearConstraint definition targets them (parameters.h:319-320): staged only then.
  bool needs_ee_ang = false;
  for (int i = 0; i < s.n_ee_linear; ++i) needs_ee_ang = needs_ee_ang || s.ee_linear_target[i] == 1;
  int col = 0, node_slot = 0;
  hp.polys.clear();
  hp.node_src.clear();
  for (VarSet& v : vs) {
    v.col0 = col;
    col += v.n_vars;
    if (v.kind == kEESchedule || (v.kind == kEEAng && !needs_ee_ang)) continue;
    v.node_slot0 = node_slot;
    v.poly0 = static_cast<int>(hp.polys.size());
    std::vector<int> src(static_cast<size_t>(v.n_nodes) * 6, -1);
    for (int idx = 0; idx < v.n_vars; ++idx)
      for (const NodeRef& r : v.refs[idx]) src[r.node * 6 + r.deriv * 3 + r.dim] = v.col0 + idx;
    hp.node_src.insert(hp.node_src.end(), src.begin(), src.end());
    for (int p = 0; p < v.n_polys(); ++p) {
      PolyRec pr;
      pr.node0 = node_slot + p;
      pr.pad = 0;
      pr.T = v.poly_T[p];
      pr.T2 = RuntimePow(pr.T, 2);
      pr.T3 = RuntimePow(pr.T, 3);
      hp.polys.push_back(pr);
    }
    node_slot += v.n_nodes;
  }
  hp.n_vars = col;
  hp.n_node_values = node_slot * 6;

  // runtime tables of the PhaseSpline path
  hp.rt_splines.clear();
  hp.rt_feet.clear();
  hp.poly_meta.assign(hp.polys.size(), kPolyFixed);
  if (hp.optimize_durations) {
    const int kinds[4] = {kEEMotion, kEEForce, kEETorque, kEEAng};
    const int n_kinds = needs_ee_ang ? 4 : 3;
    hp.rt_splines.resize(n_kinds * static_cast<size_t>(E));
    for (int k = 0; k < n_kinds; ++k)
      for (int e = 0; e < E; ++e) {
        const VarSet& v = *hp.Find(kinds[k], e);
        hp.rt_splines[k * E + e] = RtSpline{v.poly0, v.n_polys(), v.col0, v.n_vars};
        for (int pl = 0; pl < v.n_polys(); ++pl)
          hp.poly_meta[v.poly0 + pl] = static_cast<uint32_t>(v.poly_phase[pl]) | (static_cast<uint32_t>(v.polys_in_phase[pl]) << 8) |
                                       (static_cast<uint32_t>(v.poly_in_phase[pl]) << 16) | (static_cast<uint32_t>(e) << 24);
      }
    int phase0 = 0;
    for (int e = 0; e < E; ++e) {
      double total = 0.0;                       // std::accumulate(timings.begin(), timings.end(), 0.0)
      for (int i = 0; i < s.n_phases[e]; ++i) total += s.phase_durations[e][i];
      hp.rt_feet.push_back(RtFoot{static_cast<short>(phase0), static_cast<short>(s.n_phases[e]), hp.Find(kEESchedule, e)->col0, total});
      phase0 += s.n_phases[e];
    }
  }
}

}  // namespace tb
