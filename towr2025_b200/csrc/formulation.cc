// formulation.cc — host-side set-up: robot constants, gait tables, parameter defaults, terrain heights,
// initial guess and bounds.  Runs once per problem; nothing here is on the per-iteration path.
//
// Reference behaviour restated (paths relative to towr/):
//   Parameters::Parameters                    src/parameters.cc:40-105
//   RobotModel                                include/towr/models/examples/{monoped,biped,hyq,anymal}_model.h
//   GaitGenerator                             src/initialization/gait_generator.cc:54-144 + per-robot stride tables
//   NlpFormulation::Make*Variables            src/nlp_formulation.cc:121-346 (initial guess + variable bounds)
//   ConstraintSet::GetBounds of each set      src/constraints/*.cc
#include <cmath>
#include <cstring>
#include <stdexcept>

#include "structure.h"

namespace tb {

// ------------------------------------------------------------------------------------------------ robots
namespace {
struct RobotDef {
  int n_ee;
  double mass, I[6];           // Ixx Iyy Izz Ixy Ixz Iyz
  double nominal[4][3];
  double dev[3];               // symmetric box: max = +dev, min = -dev
};
const RobotDef kRobots[4] = {
    {1, 20.0, {1.2, 5.5, 6.0, 0.0, -0.2, -0.01}, {{0.0, 0.0, -0.58}}, {0.30, 0.15, 0.30}},
    {2, 20.0, {1.209, 5.583, 6.056, 0.005, -0.190, -0.012}, {{0.0, 0.20, -0.65}, {0.0, -0.20, -0.65}}, {0.25, 0.15, 0.40}},
    {4, 83.0, {4.26, 8.97, 9.88, -0.0063, 0.193, 0.0126},
     {{0.31, 0.29, -0.58}, {0.31, -0.29, -0.58}, {-0.31, 0.29, -0.58}, {-0.31, -0.29, -0.58}}, {0.25, 0.20, 0.10}},
    {4, 29.5, {0.946438, 1.94478, 2.01835, 0.000938112, -0.00595386, -0.00146328},
     {{0.34, 0.19, -0.42}, {0.34, -0.19, -0.42}, {-0.34, 0.19, -0.42}, {-0.34, -0.19, -0.42}}, {0.15, 0.1, 0.10}},
};

// ------------------------------------------------------------------------------------------------- gaits
// A stride is a list of (duration, contact mask); bit e of the mask = foot e on the ground.
struct Step { double t; unsigned mask; };
using Stride = std::vector<Step>;

Stride WithoutTransition(Stride s) {   // gait_generator.cc:131-144: last phase merged into the one before
  double last = s.back().t;
  s.pop_back();
  s.back().t += last;
  return s;
}

enum Gait { Stand, Flight, Walk1, Walk2, Walk2E, Run1, Run2, Run2E, Run3, Run3E, Hop1, Hop1E, Hop2, Hop3, Hop3E, Hop5 };

Stride MonopedStride(int g) {
  switch (g) {
    case Stand: return {{0.5, 1}};
    case Flight: return {{0.5, 0}};
    case Hop1: return {{0.3, 1}, {0.3, 0}};
    case Hop2: return {{0.2, 1}, {0.3, 0}};
  }
  throw std::runtime_error("gait not implemented for the monoped");
}
Stride BipedStride(int g) {
  const unsigned I = 0, b = 2 /*R*/, P = 1 /*L*/, B = 3;
  switch (g) {
    case Stand: return {{0.2, B}};
    case Flight: return {{0.5, I}};
    case Walk1: case Walk2: return {{0.3, b}, {0.05, B}, {0.3, P}, {0.05, B}};
    case Run1: case Run3: return {{0.15, b}, {0.4, I}, {0.15 + 0.15, P}, {0.4, I}, {0.15, b}};
    case Hop1: return {{0.15, B}, {0.5, I}, {0.15, B}};
    case Hop2: return {{0.15, b}, {0.4, I}, {0.15, b}};
    case Hop3: return {{0.2, P}, {0.2, I}, {0.2, P}};
    case Hop5: return {{0.2, P}, {0.3, I}, {0.2, b}, {0.2, B}};
  }
  throw std::runtime_error("gait not implemented for the biped");
}
Stride QuadrupedStride(int g) {
  // LF = 1, RF = 2, LH = 4, RH = 8
  const unsigned II = 0x0, IP = 0x1, Pb = 0x6, bP = 0x9, BI = 0xC, IB = 0x3, PP = 0x5, bb = 0xA,
                 Bb = 0xE, BP = 0xD, bB = 0xB, PB = 0x7, BB = 0xF;
  const Stride walk_overlap = {{0.25, bB}, {0.13, bb}, {0.25, Bb}, {0.13, Pb}, {0.25, PB}, {0.13, PP}, {0.25, BP}, {0.13, bP}};
  const Stride gallop = {{0.2, Bb}, {0.3, BI}, {0.2, BP}, {0.2, bP}, {0.2, bB}, {0.3, IB}, {0.2, PB}, {0.2, Pb}};
  switch (g) {
    case Stand: return {{0.3, BB}};
    case Flight: return {{0.3, Bb}};
    case Walk1: return {{0.3, bB}, {0.2, BB}, {0.3, Bb}, {0.2, BB}, {0.3, PB}, {0.2, BB}, {0.3, BP}, {0.2, BB}};
    case Walk2: return walk_overlap;
    case Walk2E: return WithoutTransition(walk_overlap);
    case Run1: return {{0.3, bP}, {0.2, BB}, {0.3, Pb}, {0.2, BB}};
    case Run2: return {{0.4, bP}, {0.1, II}, {0.4, Pb}, {0.1, II}};
    case Run2E: return {{0.4, bP}};
    case Run3: return {{0.3, PP}, {0.1, II}, {0.3, bb}, {0.1, II}};
    case Run3E: return {{0.3, PP}};
    case Hop1: return {{0.3, BI}, {0.1, II}, {0.3, IB}, {0.1, II}};
    case Hop1E: return {{0.3, BI}};
    case Hop2: return {{0.3, BB}, {0.4, II}, {0.3, BB}};
    case Hop3: return gallop;
    case Hop3E: return WithoutTransition(gallop);
    case Hop5: return {{0.1, Bb}, {0.2, BB}, {0.1, IP}, {0.1, Bb}, {0.2, BB}, {0.1, IP}};
  }
  throw std::runtime_error("gait not implemented for the quadruped");
}

std::vector<int> ComboGaits(int n_ee, int combo) {
  if (combo < 0 || combo > 4) throw std::runtime_error("gait combo must be C0..C4");
  static const std::vector<int> mono[5] = {{Stand, Hop1, Hop1, Hop1, Hop1, Stand}, {Stand, Hop1, Hop1, Hop1, Stand},
                                           {Stand, Hop1, Hop1, Hop1, Hop1, Stand}, {Stand, Hop2, Hop2, Hop2, Stand},
                                           {Stand, Hop2, Hop2, Hop2, Hop2, Hop2, Stand}};
  static const std::vector<int> bip[5] = {{Stand, Walk1, Walk1, Walk1, Walk1, Stand}, {Stand, Run1, Run1, Run1, Run1, Stand},
                                          {Stand, Hop1, Hop1, Hop1, Stand}, {Stand, Hop1, Hop2, Hop2, Stand},
                                          {Stand, Hop5, Hop5, Hop5, Stand}};
  static const std::vector<int> quad[5] = {{Stand, Walk2, Walk2, Walk2, Walk2E, Stand}, {Stand, Run2, Run2, Run2, Run2E, Stand},
                                           {Stand, Run3, Run3, Run3, Run3E, Stand}, {Stand, Hop1, Hop1, Hop1, Hop1E, Stand},
                                           {Stand, Hop3, Hop3, Hop3, Hop3E, Stand}};
  switch (n_ee) {
    case 1: return mono[combo];
    case 2: return bip[combo];
    case 4: return quad[combo];
  }
  throw std::runtime_error("gait generator not implemented for this leg count");
}
}  // namespace

void SpecDefaults(towr_b200_spec* s) {
  std::memset(s, 0, sizeof(*s));
  const int def[] = {TOWR_B200_C_TERRAIN, TOWR_B200_C_DYNAMIC, TOWR_B200_C_BASE_ACC, TOWR_B200_C_ENDEFFECTOR_ROM,
                     TOWR_B200_C_FORCE, TOWR_B200_C_SWING, TOWR_B200_C_BASE_HEIGHT};
  s->n_constraints = 7;
  for (int i = 0; i < 7; ++i) s->constraints[i] = def[i];
  s->duration_base_polynomial = 0.1;
  s->force_polynomials_per_stance_phase = 3;
  s->torque_polynomials_per_stance_phase = 3;
  s->ee_polynomials_per_swing_phase = 2;
  s->force_limit_in_normal_direction = 1000;
  s->torque_tx_min = -100.0; s->torque_tx_max = 100.0;
  s->torque_ty_min = -100.0; s->torque_ty_max = 100.0;
  s->torque_k_friction = 2.0 / 3.0;
  s->dt_constraint_range_of_motion = 0.08;
  s->dt_constraint_dynamic = 0.1;
  s->dt_constraint_base_motion = s->duration_base_polynomial / 4.;
  s->dt_constraint_force = 0.02;
  s->dt_constraint_torque = 0.02;
  s->bound_phase_duration[0] = 0.2;
  s->bound_phase_duration[1] = 1.0;
  s->dt_cost_energy = 0.02;               // parameters.cc:69-71
  s->energy_cost_torque_weight = 1.0;
  s->dt_cost_ang_mom = 0.02;
  s->enable_swing_ee_base_pos_tracking = 0;            // parameters.h:233-235
  s->swing_ee_base_pos_tracking_weight = 1e-2;
  s->dt_cost_swing_ee_base_pos_tracking = 0.05;
  for (int i = 0; i < 3; ++i)
    s->bounds_final_lin_pos[i] = s->bounds_final_lin_vel[i] = s->bounds_final_ang_pos[i] = s->bounds_final_ang_vel[i] = 1;
}

void SpecRobot(towr_b200_spec* s, int robot) {
  if (robot < 0 || robot > 3) throw std::runtime_error("robot model not implemented");
  const RobotDef& r = kRobots[robot];
  s->n_ee = r.n_ee;
  s->mass = r.mass;
  for (int i = 0; i < 6; ++i) s->inertia[i] = r.I[i];
  for (int e = 0; e < r.n_ee; ++e)
    for (int d = 0; d < 3; ++d) {
      s->nominal_stance[e][d] = r.nominal[e][d];
      s->max_dev[e][d] = r.dev[d];
      s->min_dev[e][d] = -r.dev[d];
    }
}

void SpecGait(towr_b200_spec* s, int combo, double total_time) {
  const int E = s->n_ee;
  Stride all;
  for (int g : ComboGaits(E, combo)) {
    Stride st = E == 1 ? MonopedStride(g) : E == 2 ? BipedStride(g) : QuadrupedStride(g);
    all.insert(all.end(), st.begin(), st.end());
  }
  // per foot: time accumulated until its contact flag flips (gait_generator.cc:77-105)
  for (int e = 0; e < E; ++e) {
    std::vector<double> d;
    double acc = 0.0;
    for (size_t p = 0; p + 1 < all.size(); ++p) {
      acc += all[p].t;
      bool now = (all[p].mask >> e) & 1u, next = (all[p + 1].mask >> e) & 1u;
      if (now != next) {
        d.push_back(acc);
        acc = 0.0;
      }
    }
    d.push_back(acc + all.back().t);
    if (d.size() > TOWR_B200_MAX_PHASES) throw std::runtime_error("too many phases");
    double total = 0.0;
    for (double v : d) total += v;
    s->n_phases[e] = static_cast<int>(d.size());
    for (size_t i = 0; i < d.size(); ++i) s->phase_durations[e][i] = (d[i] / total) * total_time;
    s->in_contact_at_start[e] = (all.front().mask >> e) & 1u;
  }
}

void SpecNominalStates(towr_b200_spec* s, double gx, double gy, double gyaw) {
  // towr_ros_app.cc:47-58: feet under their nominal position on z = 0, base at -z_nominal
  const double z_ground = 0.0;
  for (int e = 0; e < s->n_ee; ++e) {
    s->initial_ee_W[e][0] = s->nominal_stance[e][0];
    s->initial_ee_W[e][1] = s->nominal_stance[e][1];
    s->initial_ee_W[e][2] = z_ground;
  }
  for (int i = 0; i < 6; ++i) s->initial_base_lin[i] = s->initial_base_ang[i] = s->final_base_lin[i] = s->final_base_ang[i] = 0.0;
  s->initial_base_lin[2] = -s->nominal_stance[0][2] + z_ground;
  s->final_base_lin[0] = gx;
  s->final_base_lin[1] = gy;
  s->final_base_lin[2] = s->initial_base_lin[2];
  s->final_base_ang[2] = gyaw;
}

}  // namespace tb
