// formulation_init.cc — per-problem host quantities: initial guess x0, variable bounds, constraint bounds,
// and the constants the kernels read.  Restates NlpFormulation::Make*Variables
// (towr/src/nlp_formulation.cc:121-346), NodesVariables::SetByLinearInterpolation[RelativeToBase]
// (towr/src/variables/nodes_variables.cc:130-217), AddBounds (:219-263) and each set's GetBounds.
#include <cmath>
#include <limits>
#include <stdexcept>

#include "structure.h"
#include "cost.h"
#include "terrain.h"

namespace tb {

namespace {
struct V3 { double v[3]; };
inline V3 mk(const double* p) { return V3{{p[0], p[1], p[2]}}; }

// ZYX Euler -> base-to-world rotation (euler_converter.cc:200-221)
void EulerToR(const V3& e, double R[3][3]) {
  const double x = e.v[0], y = e.v[1], z = e.v[2];
  R[0][0] = cos(y) * cos(z); R[0][1] = cos(z) * sin(x) * sin(y) - cos(x) * sin(z); R[0][2] = sin(x) * sin(z) + cos(x) * cos(z) * sin(y);
  R[1][0] = cos(y) * sin(z); R[1][1] = cos(x) * cos(z) + sin(x) * sin(y) * sin(z); R[1][2] = cos(x) * sin(y) * sin(z) - cos(z) * sin(x);
  R[2][0] = -sin(y);         R[2][1] = cos(y) * sin(x);                            R[2][2] = cos(x) * cos(y);
}
V3 MatVec(const double R[3][3], const V3& a, bool transpose) {
  V3 o;
  for (int i = 0; i < 3; ++i)
    o.v[i] = transpose ? R[0][i] * a.v[0] + R[1][i] * a.v[1] + R[2][i] * a.v[2]
                       : R[i][0] * a.v[0] + R[i][1] * a.v[1] + R[i][2] * a.v[2];
  return o;
}

constexpr double kNoBound = 1.0e20;   // ifopt::inf

struct Derived {           // quantities several variable sets share
  double T;
  V3 base_final_pos;       // final base position used for the initial guess (terrain height - nominal z)
  V3 ee_final[kMaxEE];     // final foothold, on the terrain
};

TerrainTable SpecTable(const towr_b200_spec& s) {
  TerrainTable t{};
  if (s.terrain_id != TOWR_B200_TERRAIN_TABLE) return t;
  t.n = s.terrain_table_n;
  t.mode = s.terrain_table_mode;
  for (int i = 0; i < t.n && i < kMaxTerrainKnots; ++i) { t.x[i] = s.terrain_table_x[i]; t.h[i] = s.terrain_table_h[i]; }
  return t;
}
// terrain_->GetHeight etc. of the problem's own height map
TerrainSample SpecTerrain(const towr_b200_spec& s, double x, double y) {
  return s.terrain_id == TOWR_B200_TERRAIN_TABLE ? EvalTerrainTable(SpecTable(s), x) : EvalTerrain(s.terrain_id, x, y);
}

Derived Derive(const towr_b200_spec& s) {
  Derived d;
  d.T = TotalTime(s);
  const double x = s.final_base_lin[0], y = s.final_base_lin[1];
  d.base_final_pos = V3{{x, y, SpecTerrain(s, x, y).h - s.nominal_stance[0][2]}};
  double R[3][3];
  EulerToR(V3{{0.0, 0.0, s.final_base_ang[2]}}, R);
  for (int e = 0; e < s.n_ee; ++e) {
    V3 off = MatVec(R, mk(s.nominal_stance[e]), false);
    double ex = s.final_base_lin[0] + off.v[0], ey = s.final_base_lin[1] + off.v[1];
    d.ee_final[e] = V3{{ex, ey, SpecTerrain(s, ex, ey).h}};
  }
  return d;
}

// value a linear interpolation assigns to node value (node, deriv, dim)
double LinearInterp(const V3& a, const V3& b, double T, int n_nodes, const NodeRef& r) {
  double dp = b.v[r.dim] - a.v[r.dim];
  if (r.deriv == 0) return a.v[r.dim] + (r.node / static_cast<double>(n_nodes - 1)) * dp;
  return dp / T;
}

// SetByLinearInterpolationRelativeToBase: foot fixed in the (interpolated) base frame
double RelativeToBase(const towr_b200_spec& s, const Derived& d, int ee, int n_nodes, const NodeRef& r) {
  if (d.T <= 0.0) throw std::runtime_error("SetByLinearInterpolationRelativeToBase requires t_total > 0.");
  const V3 b0 = mk(s.initial_base_lin), bT = mk(s.final_base_lin);
  const V3 a0 = mk(s.initial_base_ang), aT = mk(s.final_base_ang);
  double R0[3][3], RT[3][3], R[3][3];
  EulerToR(a0, R0);
  EulerToR(aT, RT);
  V3 d0, dT;
  for (int i = 0; i < 3; ++i) { d0.v[i] = s.initial_ee_W[ee][i] - b0.v[i]; dT.v[i] = d.ee_final[ee].v[i] - bT.v[i]; }
  const V3 r0_B = MatVec(R0, d0, true), rT_B = MatVec(RT, dT, true);
  V3 dp_B, avg_vel_B, base_avg_vel;
  for (int i = 0; i < 3; ++i) {
    dp_B.v[i] = rT_B.v[i] - r0_B.v[i];
    avg_vel_B.v[i] = dp_B.v[i] / d.T;
    base_avg_vel.v[i] = (bT.v[i] - b0.v[i]) / d.T;
  }
  const double alpha = r.node / static_cast<double>(n_nodes - 1);
  V3 base_pos, base_rpy, r_B;
  for (int i = 0; i < 3; ++i) {
    base_pos.v[i] = (1.0 - alpha) * b0.v[i] + alpha * bT.v[i];
    base_rpy.v[i] = (1.0 - alpha) * a0.v[i] + alpha * aT.v[i];
    r_B.v[i] = r0_B.v[i] + alpha * dp_B.v[i];
  }
  EulerToR(base_rpy, R);
  if (r.deriv == 0) return base_pos.v[r.dim] + MatVec(R, r_B, false).v[r.dim];
  return base_avg_vel.v[r.dim] + MatVec(R, avg_vel_B, false).v[r.dim];
}

void SetBound(const VarSet& v, int node, int deriv, int dim, double lo, double up, double* L, double* U) {
  for (int idx = 0; idx < v.n_vars; ++idx)
    for (const NodeRef& r : v.refs[idx])
      if (r.node == node && r.deriv == deriv && r.dim == dim) { L[v.col0 + idx] = lo; U[v.col0 + idx] = up; }
}
}  // namespace

void ProblemX0(const HostProgram& hp, const towr_b200_spec& s, double* x) {
  const Derived d = Derive(s);
  const double g = 9.80665;   // dynamic_model.cc:37
  const V3 f_stance{{0.0, 0.0, s.mass * g / s.n_ee}};
  const V3 zero{{0.0, 0.0, 0.0}};
  for (const VarSet& v : hp.var_sets) {
    for (int idx = 0; idx < v.n_vars; ++idx) {
      double val = 0.0;
      if (v.kind == kEESchedule) {
        val = s.phase_durations[v.ee][idx];
      } else {
        const NodeRef& r = v.refs[idx].back();   // GetValues(): the last node value mapped to idx wins
        switch (v.kind) {
          case kBaseLin: val = LinearInterp(mk(s.initial_base_lin), d.base_final_pos, d.T, v.n_nodes, r); break;
          case kBaseAng:
          case kEEAng: val = LinearInterp(mk(s.initial_base_ang), mk(s.final_base_ang), d.T, v.n_nodes, r); break;
          case kEEMotion: val = RelativeToBase(s, d, v.ee, v.n_nodes, r); break;
          case kEEForce: val = LinearInterp(f_stance, f_stance, d.T, v.n_nodes, r); break;
          case kEETorque: val = LinearInterp(zero, zero, d.T, v.n_nodes, r); break;
        }
      }
      x[v.col0 + idx] = val;
    }
  }
}

void ProblemVarBounds(const HostProgram& hp, const towr_b200_spec& s, double* L, double* U) {
  const Derived d = Derive(s);
  const double g = 9.80665;
  for (int i = 0; i < hp.n_vars; ++i) { L[i] = -kNoBound; U[i] = +kNoBound; }
  auto fix = [&](const VarSet& v, int node, int deriv, const int* dim_flags, const double* val) {
    for (int dim = 0; dim < 3; ++dim)
      if (!dim_flags || dim_flags[dim]) SetBound(v, node, deriv, dim, val[dim], val[dim], L, U);
  };
  const double zero3[3] = {0.0, 0.0, 0.0};
  for (const VarSet& v : hp.var_sets) {
    const int last = v.n_nodes - 1;
    switch (v.kind) {
      case kBaseLin:
        fix(v, 0, 0, nullptr, s.initial_base_lin);
        fix(v, 0, 1, nullptr, s.initial_base_lin + 3);
        fix(v, last, 0, s.bounds_final_lin_pos, s.final_base_lin);
        fix(v, last, 1, s.bounds_final_lin_vel, s.final_base_lin + 3);
        break;
      case kBaseAng:
        fix(v, 0, 0, nullptr, s.initial_base_ang);
        fix(v, 0, 1, nullptr, s.initial_base_ang + 3);
        fix(v, last, 0, s.bounds_final_ang_pos, s.final_base_ang);
        fix(v, last, 1, s.bounds_final_ang_vel, s.final_base_ang + 3);
        break;
      case kEEMotion:
        if (s.enable_stance_tracking) {                    // nlp_formulation.cc:218-231
          int stance_id = 0;
          for (int phase = 0; phase < s.n_phases[v.ee]; ++phase) {
            const int node = v.PolyAtStartOfPhase(phase);  // GetNodeIDAtStartOfPhase: start node of that polynomial
            if (!v.IsConstantNode(node)) continue;
            if (stance_id >= s.n_stance_positions[v.ee] || stance_id >= TOWR_B200_MAX_STANCES)
              throw std::out_of_range("ee_stance_position_ has fewer entries than the foot has stance phases (enable_stance_tracking)");
            const double pos[3] = {s.ee_stance_position[v.ee][stance_id][0], s.ee_stance_position[v.ee][stance_id][1], 0.0};
            const int xy[3] = {1, 1, 0};
            fix(v, node, 0, xy, pos);
            ++stance_id;
            ++phase;                                       // the reference skips the phase after a bounded one (:229)
          }
        } else {                                           // :232-236
          fix(v, 0, 0, nullptr, s.initial_ee_W[v.ee]);
          fix(v, last, 0, nullptr, d.ee_final[v.ee].v);
        }
        break;
      case kEEAng:
        fix(v, 0, 0, nullptr, s.initial_base_ang);
        fix(v, last, 0, nullptr, s.final_base_ang);
        break;
      case kEEForce: {
        const double f[3] = {0.0, 0.0, s.mass * g / s.n_ee};
        fix(v, 0, 0, nullptr, f);
        fix(v, last, 0, nullptr, f);
        fix(v, 0, 1, nullptr, zero3);
        fix(v, last, 1, nullptr, zero3);
        break;
      }
      case kEETorque:
        fix(v, 0, 0, nullptr, zero3);
        fix(v, last, 0, nullptr, zero3);
        fix(v, 0, 1, nullptr, zero3);
        fix(v, last, 1, nullptr, zero3);
        break;
      case kEESchedule:
        for (int i = 0; i < v.n_vars; ++i) { L[v.col0 + i] = s.bound_phase_duration[0]; U[v.col0 + i] = s.bound_phase_duration[1]; }
        break;
    }
  }
}

void ProblemConBounds(const HostProgram& hp, const towr_b200_spec& s, double* L, double* U) {
  std::vector<double> x0;                                // the initial guess, built when a bound depends on it
  auto node_pos = [&](const VarSet& v, int node, int dim) {   // a node's position in the initial guess (0: not optimised)
    if (x0.empty()) { x0.resize(hp.n_vars); ProblemX0(hp, s, x0.data()); }
    const int idx = v.OptIndex(node, 0, dim);
    return idx < 0 ? 0.0 : x0[v.col0 + idx];
  };
  for (int r = 0; r < hp.n_rows; ++r) {
    const RowBound& b = hp.row_bounds[r];
    double lo = 0.0, up = 0.0;
    switch (b.kind) {
      case kBoundZero: break;
      case kBoundTerrainSwing: lo = 0.02; up = std::numeric_limits<double>::infinity(); break;   // nlp_formulation.cc:464-473
      case kBoundRom: {
        double base = 0.0;
        base += s.nominal_stance[b.ee][b.dim];
        up = base + s.max_dev[b.ee][b.dim];
        lo = base + s.min_dev[b.ee][b.dim];
        break;
      }
      case kBoundForceNormal: lo = 0.0; up = s.force_limit_in_normal_direction; break;
      case kBoundLE0: lo = -kNoBound; up = 0.0; break;
      case kBoundGE0: lo = 0.0; up = +kNoBound; break;
      case kBoundBaseHeight: lo = 0.0; up = 1e20; break;
      case kBoundTotalDuration: lo = 0.1; up = hp.total_time - 0.2; break;
      case kBoundNone: lo = -kNoBound; up = +kNoBound; break;
      case kBoundTorqueX: lo = s.torque_tx_min; up = s.torque_tx_max; break;     // torque_constraint_discretized.cc:131-139
      case kBoundTorqueY: lo = s.torque_ty_min; up = s.torque_ty_max; break;
      case kBoundEELinear: lo = -s.ee_linear_tolerance[b.dim]; up = s.ee_linear_tolerance[b.dim]; break;   // ee_linear_constraint.cc:31-34
      case kBoundTorqueNodeZ: {                          // torque_constraint.cc:94-112, evaluated on the initial guess
        const VarSet& fo = *hp.Find(kEEForce, b.ee);
        const VarSet& mo = *hp.Find(kEEMotion, b.ee);
        const int fnode = fo.PolyAtStartOfPhase(b.dim), pnode = mo.PolyAtStartOfPhase(b.dim);
        const double f[3] = {node_pos(fo, fnode, 0), node_pos(fo, fnode, 1), node_pos(fo, fnode, 2)};
        const TerrainSample ts = SpecTerrain(s, node_pos(mo, pnode, 0), node_pos(mo, pnode, 1));
        double n[3], t1[3], t2[3];
        TerrainBasis(ts, n, t1, t2);
        const double f_n = (f[0] * n[0] + f[1] * n[1]) + f[2] * n[2];
        const double tz_limit = s.torque_k_friction * kFrictionCoeff * f_n;
        lo = -tz_limit;
        up = tz_limit;
        break;
      }
    }
    L[r] = lo;
    U[r] = up;
  }
}

// NlpFormulation::GetCosts (nlp_formulation.cc:616-620): the foot's initial position in the initial base frame
void SwingReferenceB(const towr_b200_spec& s, double out[12]) {
  double R[3][3];
  EulerToR(V3{{s.initial_base_ang[0], s.initial_base_ang[1], s.initial_base_ang[2]}}, R);
  for (int e = 0; e < 4; ++e)
    for (int i = 0; i < 3; ++i) {
      const double r[3] = {s.initial_ee_W[e][0] - s.initial_base_lin[0], s.initial_ee_W[e][1] - s.initial_base_lin[1],
                           s.initial_ee_W[e][2] - s.initial_base_lin[2]};
      out[3 * e + i] = e < s.n_ee ? (R[0][i] * r[0] + R[1][i] * r[1]) + R[2][i] * r[2] : 0.0;   // row i of R^T
    }
}

void FillProblemConst(const towr_b200_spec& s, ProblemConst& pc) {
  pc.mass = s.mass;
  const double Ixx = s.inertia[0], Iyy = s.inertia[1], Izz = s.inertia[2], Ixy = s.inertia[3], Ixz = s.inertia[4], Iyz = s.inertia[5];
  const double I[9] = {Ixx, -Ixy, -Ixz, -Ixy, Iyy, -Iyz, -Ixz, -Iyz, Izz};
  for (int i = 0; i < 9; ++i) pc.I[i] = I[i];
  pc.terrain_id = s.terrain_id;
  pc.pad = 0;
  pc.table = SpecTable(s);
}

void ProblemConstantValues(const HostProgram& hp, double mass, double torque_k_friction, bool flat_basis, double* vals) {
  // the terrain-basis components as ItemForce / ItemTorque / ItemHeightNode compute them where h_x = h_y = 0
  double code_value[kCodeCount];
  {
    TerrainSample ts = EvalTerrain(0, 0.0, 0.0);
    double n[3], t1[3], t2[3];
    TerrainBasis(ts, n, t1, t2);
    const double mu = kFrictionCoeff, kf = torque_k_friction;
    for (int d = 0; d < 3; ++d) {
      code_value[kCodeForce + d] = n[d];
      code_value[kCodeForce + 3 + d] = t1[d] - mu * n[d];
      code_value[kCodeForce + 6 + d] = t1[d] + mu * n[d];
      code_value[kCodeForce + 9 + d] = t2[d] - mu * n[d];
      code_value[kCodeForce + 12 + d] = t2[d] + mu * n[d];
      code_value[kCodeTorque + d] = t1[d];
      code_value[kCodeTorque + 3 + d] = t2[d];
      code_value[kCodeTorque + 6 + d] = n[d];
      code_value[kCodeTorque + 9 + d] = -n[d];
      code_value[kCodeTorque + 12 + d] = (-kf * mu) * n[d];
    }
    code_value[kCodeHeight] = -ts.hx;
    code_value[kCodeHeight + 1] = -ts.hy;
  }
  for (size_t e = 0; e < hp.entries.size(); ++e) {
    const uint32_t w = hp.entries[e];
    const int g = static_cast<int>(w & 0xFFFF);
    const double bv = hp.B[(w >> 16) & 0x7FFF];
    if (!(w & kEntryGG) && g < kGReserved) {
      const double gv = g == kGOne ? 1.0 : g == kGMinusOne ? -1.0 : mass;       // what stage A puts into these slots
      vals[e] = gv * bv;                                                          // the same product stage B forms
    } else if (flat_basis && hp.entry_code[e] >= 0) {
      vals[e] = code_value[hp.entry_code[e]] * bv;
    } else {
      vals[e] = 0.0;
    }
  }
}

double TerrainHeight(int terrain_id, double x, double y) { return EvalTerrain(terrain_id, x, y).h; }

}  // namespace tb
