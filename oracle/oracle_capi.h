/* ORACLE — TEST INFRASTRUCTURE ONLY (see towr_oracle.h header).
 * Plain-C surface of the CPU oracle, loaded with ctypes by tests/, smoke() and
 * bench.py's cpu_baseline / --impl reference legs.  Not part of the product ABI
 * (that one is include/towr_b200.h). */
#ifndef ORACLE_CAPI_H_
#define ORACLE_CAPI_H_

#ifdef __cplusplus
extern "C" {
#endif

#define ORC_MAX_EE 4
#define ORC_MAX_PHASES 32
#define ORC_MAX_CONSTRAINTS 16
#define ORC_MAX_COSTS 4
#define ORC_MAX_NODE_COSTS 32
#define ORC_MAX_EE_LINEAR 4
#define ORC_MAX_EE_LINEAR_TERMS 4
#define ORC_MAX_STANCES 17
#define ORC_MAX_TERRAIN_KNOTS 16

/* The public members of towr::NlpFormulation (nlp_formulation.h:100-105) plus the
 * fields of towr::Parameters (parameters.h:135-336) the hot path reads. */
typedef struct orc_spec {
  int n_ee;
  /* RobotModel: SingleRigidBodyDynamics(mass, Ixx,Iyy,Izz,Ixy,Ixz,Iyz) + KinematicModel */
  double mass;
  double inertia[6];
  double nominal_stance[ORC_MAX_EE][3];
  double max_dev[ORC_MAX_EE][3];
  double min_dev[ORC_MAX_EE][3];
  int terrain_id;                         /* HeightMap::TerrainID, 7 = FiveStepStairs of hopper_example.cc */
  /* Parameters */
  int n_phases[ORC_MAX_EE];
  double phase_durations[ORC_MAX_EE][ORC_MAX_PHASES];
  int in_contact_at_start[ORC_MAX_EE];
  int n_constraints;
  int constraints[ORC_MAX_CONSTRAINTS];   /* Parameters::ConstraintName values, in order */
  double dt_constraint_dynamic, dt_constraint_range_of_motion, dt_constraint_base_motion;
  double dt_constraint_force, dt_constraint_torque;
  double duration_base_polynomial;
  int ee_polynomials_per_swing_phase, force_polynomials_per_stance_phase, torque_polynomials_per_stance_phase;
  double force_limit_in_normal_direction;
  double torque_tx_min, torque_tx_max, torque_ty_min, torque_ty_max, torque_k_friction;
  double bound_phase_duration[2];
  int bounds_final_lin_pos[3], bounds_final_lin_vel[3], bounds_final_ang_pos[3], bounds_final_ang_vel[3]; /* flag per dim */
  /* BaseState initial/final: p(3) then v(3); initial_ee_W_ */
  double initial_base_lin[6], initial_base_ang[6], final_base_lin[6], final_base_ang[6];
  double initial_ee_W[ORC_MAX_EE][3];
  /* Parameters::costs_ (parameters.h:163,192): (CostName, weight) pairs in order */
  int n_costs;
  int cost_ids[ORC_MAX_COSTS];
  double cost_weights[ORC_MAX_COSTS];
  /* Parameters::enable_stance_tracking, ee_stance_position_ (parameters.h:179,225) */
  int enable_stance_tracking;
  int n_stance_positions[ORC_MAX_EE];
  double ee_stance_position[ORC_MAX_EE][ORC_MAX_STANCES][2];
  /* terrain_id 8: a user HeightMap given as data (a profile along x; mode 0 steps, 1 ramps) - the restatement of "derive
   * from HeightMap and override GetHeight / GetHeightDerivWrtX" (test/hopper_example.cc:53-91) for a piecewise profile */
  int terrain_table_n, terrain_table_mode;
  double terrain_table_x[ORC_MAX_TERRAIN_KNOTS], terrain_table_h[ORC_MAX_TERRAIN_KNOTS];
  /* Parameters::dt_cost_energy_, energy_cost_torque_weight_, dt_cost_ang_mom_ (parameters.h:245-247) */
  double dt_cost_energy, energy_cost_torque_weight, dt_cost_ang_mom;
  /* Parameters::enable_swing_ee_base_pos_tracking, swing_ee_base_pos_tracking_weight_, dt_cost_swing_ee_base_pos_tracking_
   * (parameters.h:233-235) */
  int enable_swing_ee_base_pos_tracking;
  double swing_ee_base_pos_tracking_weight, dt_cost_swing_ee_base_pos_tracking;
  /* individual NodeCost(nodes_id, deriv, dim, weight) terms (node_cost.cc:36-45), as the hand-built examples add them
   * (test/squat_jump_example.cc:205-211); set: 0 base-lin, 1 base-ang, 2 ee-motion_, 3 ee-ang_, 4 ee-force_, 5 ee-torque_ */
  int n_node_costs;
  int node_cost_set[ORC_MAX_NODE_COSTS], node_cost_ee[ORC_MAX_NODE_COSTS];
  int node_cost_deriv[ORC_MAX_NODE_COSTS], node_cost_dim[ORC_MAX_NODE_COSTS];
  double node_cost_weight[ORC_MAX_NODE_COSTS];
  /* Parameters::ee_linear_constraints_ (parameters.h:315-327): target 0 Motion / 1 Angle, deriv 0 kPos / 1 kVel */
  int n_ee_linear;
  int ee_linear_target[ORC_MAX_EE_LINEAR], ee_linear_deriv[ORC_MAX_EE_LINEAR], ee_linear_n_terms[ORC_MAX_EE_LINEAR];
  int ee_linear_term_ee[ORC_MAX_EE_LINEAR][ORC_MAX_EE_LINEAR_TERMS];
  int ee_linear_term_dim[ORC_MAX_EE_LINEAR][ORC_MAX_EE_LINEAR_TERMS];
  double ee_linear_term_coeff[ORC_MAX_EE_LINEAR][ORC_MAX_EE_LINEAR_TERMS];
  double ee_linear_tolerance[ORC_MAX_EE_LINEAR], ee_linear_dt[ORC_MAX_EE_LINEAR];
} orc_spec;

/* Parameters::Parameters() defaults (parameters.cc:40-105), enable_stance_tracking=false. */
void orc_spec_defaults(orc_spec* s);
/* RobotModel(robot) constants: 0 Monoped, 1 Biped, 2 Hyq, 3 Anymal.  Sets n_ee. */
int orc_spec_robot(orc_spec* s, int robot);
/* GaitGenerator::MakeGaitGenerator(n_ee)->SetCombo(combo); phase durations scaled to T (towr_ros_app.cc:71-77). */
int orc_spec_gait(orc_spec* s, int combo, double T);
/* towr_ros_app.cc:47-58: feet at nominal xy on z=0, base above; goal (x,y,yaw), final z = initial z. */
void orc_spec_nominal_states(orc_spec* s, double goal_x, double goal_y, double goal_yaw);

typedef struct orc_problem orc_problem;
orc_problem* orc_create(const orc_spec* s);
void orc_destroy(orc_problem* p);
const char* orc_last_error(void);

int orc_n_vars(const orc_problem* p);
int orc_n_rows(const orc_problem* p);
int orc_n_var_sets(const orc_problem* p);
int orc_n_con_sets(const orc_problem* p);
const char* orc_var_set_name(const orc_problem* p, int i);
int orc_var_set_rows(const orc_problem* p, int i);
const char* orc_con_set_name(const orc_problem* p, int i);
int orc_con_set_rows(const orc_problem* p, int i);

void orc_x0(const orc_problem* p, double* x);                       /* initial guess = GetVariableValues() at set-up */
void orc_var_bounds(const orc_problem* p, double* lower, double* upper);
void orc_con_bounds(const orc_problem* p, double* lower, double* upper);
/* eval_g */
void orc_eval_g(orc_problem* p, const double* x, double* g);
/* number of stored Jacobian entries at x (x == NULL: at the current variables) */
int orc_nnz(orc_problem* p, const double* x);
/* eval_jac_g: rows/cols/vals may each be NULL; returns the number of stored entries */
int orc_eval_jac(orc_problem* p, const double* x, int* rows, int* cols, double* vals);

/* CPU baseline: n problems of ONE spec, x is n x n_vars row-major, g n x n_rows, vals n x nnz (nnz at x0).
 * OpenMP over problems, one object graph per thread.  Returns 0, or -1 on a pattern-size mismatch. */
int orc_batch_eval(const orc_spec* s, int n, const double* x, double* g, double* vals, int n_threads);

/* The same with one object graph per thread built once at pool creation (what the timed CPU baseline uses: the timed
 * region then holds evaluations only).  x: n x n_vars, g: n x n_rows, vals: n x orc_pool_nnz (each may be NULL). */
typedef struct orc_pool orc_pool;
orc_pool* orc_pool_create(const orc_spec* s, int n_threads);
void orc_pool_destroy(orc_pool* p);
int orc_pool_threads(const orc_pool* p);
int orc_pool_nnz(const orc_pool* p);
int orc_pool_eval(orc_pool* p, int n, const double* x, double* g, double* vals);

/* TEST SWITCH (no counterpart in the reference): 0 = std::pow everywhere (default, the reference's arithmetic);
 * 1 = the polynomials of a PhaseSpline (optimised durations) use correctly rounded x^2, x^3, x^4 - what the device
 * computes at run time, where glibc's pow is not available.  Global; set before evaluating. */
void orc_set_pow_mode(int mode);
int orc_get_pow_mode(void);
void orc_pow_compare(double x, int n, double* std_pow, double* exact_pow, double* quad_rounded);

/* debug hook: see oracle_capi.cc */
void orc_debug_spline(orc_problem* p, const double* x, int kind, int ee, double t, double* out13);
void orc_debug_poly_nodes(orc_problem* p, int kind, int ee, int poly, double* out12);

/* unit hooks for known-answer tests */
void orc_hermite_point(const double* p0, const double* v0, const double* p1, const double* v1, double T, double t,
                       double* pos, double* vel, double* acc);
/* out[dxdt*4 + {start_p,start_v,end_p,end_v}] for dxdt in pos,vel,acc */
void orc_hermite_basis(double T, double t, double* out12);
void orc_hermite_dpos_dT(const double* p0, const double* v0, const double* p1, const double* v1, double T, double t, double* out3);
void orc_terrain(int terrain_id, double x, double y, double* height, double* dh /*2*/, double* basis /*3x3 n,t1,t2*/,
                 double* dbasis /*3x2x3: [basis][dim][xyz]*/);
void orc_sample_times(double T, double dt, double* out, int* n);    /* time_discretization_constraint.cc:41-49 */

/* utils/save_data.cpp:9-132 (SaveTrajectoryToCSV): the sampling loop and the values of one CSV row per sample
 * (19 + 25 n_ee doubles: time, base, per foot ...; is_contact as 0/1).  rows may be NULL; returns the row count. */
/* eval_f / eval_grad_f: ifopt::Problem::EvaluateCostFunction / EvaluateCostFunctionGradient over the cost terms of
 * NlpFormulation::GetCosts (nlp_formulation.cc:605-681; NodeCost, costs/node_cost.cc:36-76).  grad is n_vars doubles. */
double orc_eval_f(orc_problem* p, const double* x);
void orc_eval_grad_f(orc_problem* p, const double* x, double* grad);
int orc_trajectory_width(const orc_problem* p);
int orc_sample_trajectory(orc_problem* p, const double* x, double dt, double* rows);

#ifdef __cplusplus
}
#endif
#endif
