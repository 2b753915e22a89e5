/* towr_b200.h — C ABI of the batched B200 evaluation backend for TOWR's NLP.
 *
 * One eval = for one trajectory-optimisation problem, given its variable vector x,
 * all constraint values g and all stored values of the sparse constraint Jacobian.
 * This library evaluates many problems of one *structure class* per launch on a B200.
 *
 * What each entry point replaces in the reference (paths relative to towr/):
 *   - the callbacks ifopt::Problem makes per Ipopt iteration:
 *     VariableSet::SetVariables      src/variables/nodes_variables.cc:68-76, phase_durations.cc:79-100
 *     ConstraintSet::GetValues       src/constraints/time_discretization_constraint.cc:65-75 (+ node-based sets)
 *     ConstraintSet::FillJacobianBlock  src/constraints/time_discretization_constraint.cc:89-96 (+ node-based sets)
 *     -> towr_b200_batch_eval / towr_b200_batch_eval_device
 *   - ConstraintSet::GetBounds / VariableSet::GetBounds / GetValues (initial guess)
 *     src/nlp_formulation.cc:121-363, nodes_variables.cc:118-122,219-263
 *     -> towr_b200_problem_x0 / _var_bounds / _con_bounds   (host, set-up only)
 *   - the first eval_jac_g(values == NULL) walk that defines the sparsity pattern
 *     -> towr_b200_structure_pattern
 *   - NlpFormulation's public members + Parameters    include/towr/nlp_formulation.h:100-105,
 *     include/towr/parameters.h:135-336   -> towr_b200_spec
 *
 * Conventions: every function returns 0 on success and a negative towr_b200_status
 * otherwise (never throws across the boundary); towr_b200_last_error() gives the text
 * for the calling thread.  Buffers are caller-owned.  There is NO CPU fallback: the
 * eval entry points fail with TOWR_B200_ERR_CUDA when no CUDA device is usable.
 */
#ifndef TOWR_B200_H_
#define TOWR_B200_H_

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TOWR_B200_MAX_EE 4
#define TOWR_B200_MAX_PHASES 32
#define TOWR_B200_MAX_CONSTRAINTS 16
#define TOWR_B200_MAX_COSTS 4
#define TOWR_B200_MAX_NODE_COSTS 32
#define TOWR_B200_MAX_EE_LINEAR 4
#define TOWR_B200_MAX_EE_LINEAR_TERMS 4
/* variable sets a NodeCost term can name (variable_names.h: base-lin, base-ang, ee-motion_, ee-ang_, ee-force_, ee-torque_) */
enum { TOWR_B200_NODES_BASE_LIN = 0, TOWR_B200_NODES_BASE_ANG = 1, TOWR_B200_NODES_EE_MOTION = 2, TOWR_B200_NODES_EE_ANG = 3,
       TOWR_B200_NODES_EE_FORCE = 4, TOWR_B200_NODES_EE_TORQUE = 5 };
#define TOWR_B200_MAX_STANCES 17   /* stance phases of one foot: (MAX_PHASES + 1) / 2 */

typedef enum towr_b200_status {
  TOWR_B200_OK = 0,
  TOWR_B200_ERR_ARG = -1,        /* bad argument / inconsistent spec (std::runtime_error in the reference) */
  TOWR_B200_ERR_UNSUPPORTED = -2,/* valid in the reference, not built here yet */
  TOWR_B200_ERR_CUDA = -3,       /* CUDA runtime error or no device */
  TOWR_B200_ERR_STRUCTURE = -4   /* problem does not belong to the batch's structure class */
} towr_b200_status;

/* towr::Parameters::ConstraintName, include/towr/parameters.h:142-153 */
enum { TOWR_B200_C_DYNAMIC = 0, TOWR_B200_C_ENDEFFECTOR_ROM, TOWR_B200_C_TOTAL_TIME, TOWR_B200_C_TERRAIN,
       TOWR_B200_C_TERRAIN_HARD, TOWR_B200_C_FORCE, TOWR_B200_C_TORQUE, TOWR_B200_C_SWING, TOWR_B200_C_BASE_ROM,
       TOWR_B200_C_BASE_ACC, TOWR_B200_C_BASE_HEIGHT };
/* towr::RobotModel::Robot, include/towr/models/robot_model.h */
enum { TOWR_B200_MONOPED = 0, TOWR_B200_BIPED, TOWR_B200_HYQ, TOWR_B200_ANYMAL };
/* towr::HeightMap::TerrainID, include/towr/terrain/height_map.h; 7 = FiveStepStairs of test/hopper_example.cc:53-86;
 * 8 = a height map given as DATA in the spec (terrain_table_*): what a user's own HeightMap subclass is in the reference
 * (hopper_example.cc:53-91 defines one), without editing an enum and recompiling kernels */
enum { TOWR_B200_FLAT = 0, TOWR_B200_BLOCK, TOWR_B200_STAIRS, TOWR_B200_GAP, TOWR_B200_SLOPE, TOWR_B200_CHIMNEY,
       TOWR_B200_CHIMNEY_LR, TOWR_B200_FIVE_STEP_STAIRS, TOWR_B200_TERRAIN_TABLE, TOWR_B200_TERRAIN_COUNT };
#define TOWR_B200_MAX_TERRAIN_KNOTS 16
enum { TOWR_B200_TABLE_STEPS = 0, TOWR_B200_TABLE_RAMPS = 1 };
/* towr::Parameters::CostName, include/towr/parameters.h:157-161 */
enum { TOWR_B200_COST_FORCES = 0, TOWR_B200_COST_EE_MOTION = 1, TOWR_B200_COST_ENERGY = 2, TOWR_B200_COST_ANG_MOM = 3 };
/* what to evaluate */
enum { TOWR_B200_EVAL_G = 1, TOWR_B200_EVAL_JAC = 2, TOWR_B200_EVAL_BOTH = 3,
       /* host-buffer interface only: like EVAL_JAC (jac_values_out is n x nnz), but only the stored entries that depend
        * on x are brought back; the other positions of jac_values_out keep what the caller put there - the constants:
        * a few of them, between two dynamic runs, are rewritten with that same constant (towr_b200_batch_dynamic_index) */
       TOWR_B200_EVAL_JAC_DYNAMIC = 4 };

/* One NlpFormulation: the six public members of towr::NlpFormulation plus the Parameters fields
 * the path reads.  Same meaning as the reference fields of the same name. */
typedef struct towr_b200_spec {
  int n_ee;
  /* model_: SingleRigidBodyDynamics(mass, Ixx,Iyy,Izz,Ixy,Ixz,Iyz) + KinematicModel */
  double mass;
  double inertia[6];
  double nominal_stance[TOWR_B200_MAX_EE][3];
  double max_dev[TOWR_B200_MAX_EE][3];
  double min_dev[TOWR_B200_MAX_EE][3];
  int terrain_id;
  /* params_ */
  int n_phases[TOWR_B200_MAX_EE];
  double phase_durations[TOWR_B200_MAX_EE][TOWR_B200_MAX_PHASES];
  int in_contact_at_start[TOWR_B200_MAX_EE];
  int n_constraints;
  int constraints[TOWR_B200_MAX_CONSTRAINTS];
  double dt_constraint_dynamic, dt_constraint_range_of_motion, dt_constraint_base_motion;
  double dt_constraint_force, dt_constraint_torque;
  double duration_base_polynomial;
  int ee_polynomials_per_swing_phase, force_polynomials_per_stance_phase, torque_polynomials_per_stance_phase;
  double force_limit_in_normal_direction;
  double torque_tx_min, torque_tx_max, torque_ty_min, torque_ty_max, torque_k_friction;
  double bound_phase_duration[2];
  int bounds_final_lin_pos[3], bounds_final_lin_vel[3], bounds_final_ang_pos[3], bounds_final_ang_vel[3];
  /* initial_base_, final_base_ (p xyz then v xyz), initial_ee_W_ */
  double initial_base_lin[6], initial_base_ang[6], final_base_lin[6], final_base_ang[6];
  double initial_ee_W[TOWR_B200_MAX_EE][3];
  /* params_.costs_ (parameters.h:163,192): (CostName, weight) pairs in order; empty by default (parameters.cc:90) */
  int n_costs;
  int cost_ids[TOWR_B200_MAX_COSTS];
  double cost_weights[TOWR_B200_MAX_COSTS];
  /* params_.enable_stance_tracking / ee_stance_position_ (parameters.h:179,225; nlp_formulation.cc:218-231): the fork's
   * default foothold bounds - x, y of the k-th stance phase of foot e fixed to ee_stance_position[e][k] instead of the
   * start / final foothold bounds.  Bounds only: g and the Jacobian do not depend on it.  towr_b200_spec_defaults leaves
   * it OFF (upstream behaviour); the reference's default is ON and then needs the positions (hopper_example.cc:114-122),
   * or it throws std::out_of_range - here TOWR_B200_ERR_ARG. */
  int enable_stance_tracking;
  int n_stance_positions[TOWR_B200_MAX_EE];
  double ee_stance_position[TOWR_B200_MAX_EE][TOWR_B200_MAX_STANCES][2];
  /* terrain_ as data (terrain_id == TOWR_B200_TERRAIN_TABLE), per problem: a profile along x, constant in y, with knots
   * terrain_table_x[0] < ... < terrain_table_x[n-1] (1 <= n <= TOWR_B200_MAX_TERRAIN_KNOTS) and heights terrain_table_h.
   *   TABLE_STEPS: h = 0 for x < x[0], h[i] for x[i] <= x < x[i+1], h[n-1] from x[n-1] on; no slope
   *                (GetHeightDerivWrtX = 0, like the reference's Stairs, height_map_examples.cc:66-80)
   *   TABLE_RAMPS: h = h[0] for x < x[0], linear between knots with GetHeightDerivWrtX = (h[i+1]-h[i])/(x[i+1]-x[i]) for
   *                x[i] <= x < x[i+1], h[n-1] and no slope from x[n-1] on (like Slope / Block, :43-63,115-145)
   * Second derivatives are 0 (HeightMap's defaults).  A structure class created from a spec with a table terrain admits
   * problems with any terrain of the same curvature; one created on an analytic terrain refuses table terrains
   * (towr_b200_structure_accepts: the kernels that read the table are a separate instantiation). */
  int terrain_table_n, terrain_table_mode;
  double terrain_table_x[TOWR_B200_MAX_TERRAIN_KNOTS], terrain_table_h[TOWR_B200_MAX_TERRAIN_KNOTS];
  /* time-discretised costs (parameters.h:245-247, parameters.cc:69-71): sampling period and torque weight of EnergyCost
   * (costs/energy_cost.cc), sampling period of AngularMomentumCost.  Class properties, like the cost list itself. */
  double dt_cost_energy, energy_cost_torque_weight, dt_cost_ang_mom;
  /* params_.enable_swing_ee_base_pos_tracking, swing_ee_base_pos_tracking_weight_, dt_cost_swing_ee_base_pos_tracking_
   * (parameters.h:233-235; off by default): one EEBasePosCost per foot (nlp_formulation.cc:612-626, costs/
   * ee_base_pos_cost.cc) that keeps the swinging foot near its initial position in the base frame.  The switch, weight and
   * period are class properties; the reference position comes from each problem's initial_base_ / initial_ee_W. */
  int enable_swing_ee_base_pos_tracking;
  double swing_ee_base_pos_tracking_weight, dt_cost_swing_ee_base_pos_tracking;
  /* Individual NodeCost terms (towr::NodeCost(nodes_id, deriv, dim, weight), costs/node_cost.cc:36-45), for callers that
   * add them one by one instead of through a cost family: the reference's hand-built formulations do
   * (test/squat_jump_example.cc:205-211, yaw_spin_example.cc:171-176, procedural_example.cc:243-249).  Term i:
   * weight * (node value)^2 over the nodes of variable set node_cost_set[i] (TOWR_B200_NODES_*; foot node_cost_ee[i] for
   * the ee sets), derivative node_cost_deriv[i] (0 position, 1 velocity), dimension node_cost_dim[i].  Added after the
   * families of cost_ids.  Class properties. */
  int n_node_costs;
  int node_cost_set[TOWR_B200_MAX_NODE_COSTS], node_cost_ee[TOWR_B200_MAX_NODE_COSTS];
  int node_cost_deriv[TOWR_B200_MAX_NODE_COSTS], node_cost_dim[TOWR_B200_MAX_NODE_COSTS];
  double node_cost_weight[TOWR_B200_MAX_NODE_COSTS];
  /* params_.ee_linear_constraints_ (parameters.h:315-327; constraints/ee_linear_constraint.cc): definition i keeps
   * |sum_j coeff_j * spline_{ee_j}(t).at(deriv)(dim_j)| <= tolerance at t = 0, dt, ...; target 0 = the feet's motion splines,
   * 1 = their angle splines; deriv 0 = kPos, 1 = kVel.  Appended after the constraints of `constraints`
   * (nlp_formulation.cc:373-375), fixed or optimised durations.  Class properties (the tolerance too: it is a row bound of
 * the class's definition). */
  int n_ee_linear;
  int ee_linear_target[TOWR_B200_MAX_EE_LINEAR], ee_linear_deriv[TOWR_B200_MAX_EE_LINEAR], ee_linear_n_terms[TOWR_B200_MAX_EE_LINEAR];
  int ee_linear_term_ee[TOWR_B200_MAX_EE_LINEAR][TOWR_B200_MAX_EE_LINEAR_TERMS];
  int ee_linear_term_dim[TOWR_B200_MAX_EE_LINEAR][TOWR_B200_MAX_EE_LINEAR_TERMS];
  double ee_linear_term_coeff[TOWR_B200_MAX_EE_LINEAR][TOWR_B200_MAX_EE_LINEAR_TERMS];
  double ee_linear_tolerance[TOWR_B200_MAX_EE_LINEAR], ee_linear_dt[TOWR_B200_MAX_EE_LINEAR];
} towr_b200_spec;

const char* towr_b200_last_error(void);
/* 1 when a CUDA device is visible to this process */
int towr_b200_device_count(void);

/* ---- formulation helpers (host; mirror the reference's set-up code) --------------------------- */
/* Parameters::Parameters() defaults, src/parameters.cc:40-105 (enable_stance_tracking off) */
int towr_b200_spec_defaults(towr_b200_spec* s);
/* RobotModel(robot): include/towr/models/examples/{monoped,biped,hyq,anymal}_model.h */
int towr_b200_spec_robot(towr_b200_spec* s, int robot);
/* GaitGenerator::MakeGaitGenerator(n_ee)->SetCombo(combo), durations scaled to T
 * (src/initialization/gait_generator.cc:54-105, towr_ros/src/towr_ros_app.cc:71-77) */
int towr_b200_spec_gait(towr_b200_spec* s, int combo, double total_time);
/* towr_ros/src/towr_ros_app.cc:47-58: feet at nominal xy on z = 0, base above them; goal pose */
int towr_b200_spec_nominal_states(towr_b200_spec* s, double goal_x, double goal_y, double goal_yaw);
/* Parameters::OptimizePhaseDurations(), src/parameters.cc:109-112 */
int towr_b200_spec_optimize_durations(towr_b200_spec* s);

/* ---- structure class: everything the sparsity pattern depends on ----------------------------- */
typedef struct towr_b200_structure towr_b200_structure;
int towr_b200_structure_create(const towr_b200_spec* s, towr_b200_structure** out);
void towr_b200_structure_destroy(towr_b200_structure* st);
int towr_b200_structure_n_vars(const towr_b200_structure* st);
int towr_b200_structure_n_rows(const towr_b200_structure* st);
int towr_b200_structure_nnz(const towr_b200_structure* st);
int towr_b200_structure_n_var_sets(const towr_b200_structure* st);
int towr_b200_structure_n_con_sets(const towr_b200_structure* st);
const char* towr_b200_structure_var_set_name(const towr_b200_structure* st, int i);
int towr_b200_structure_var_set_rows(const towr_b200_structure* st, int i);
const char* towr_b200_structure_con_set_name(const towr_b200_structure* st, int i);
int towr_b200_structure_con_set_rows(const towr_b200_structure* st, int i);
/* row-major sorted (row, col) of every stored Jacobian entry, ifopt order; arrays of nnz ints */
int towr_b200_structure_pattern(const towr_b200_structure* st, int* rows, int* cols);
/* 0 when `s` has the same structure class as `st`, TOWR_B200_ERR_STRUCTURE otherwise */
int towr_b200_structure_accepts(const towr_b200_structure* st, const towr_b200_spec* s);

/* ---- per-problem host quantities (set-up only) ------------------------------------------------ */
int towr_b200_problem_x0(const towr_b200_structure* st, const towr_b200_spec* s, double* x);
int towr_b200_problem_var_bounds(const towr_b200_structure* st, const towr_b200_spec* s, double* lower, double* upper);
int towr_b200_problem_con_bounds(const towr_b200_structure* st, const towr_b200_spec* s, double* lower, double* upper);

/* ---- a batch of problems of one structure class, resident on one GPU -------------------------- */
typedef struct towr_b200_batch towr_b200_batch;
int towr_b200_batch_create(const towr_b200_structure* st, int n_problems, int device, towr_b200_batch** out);
void towr_b200_batch_destroy(towr_b200_batch* b);
int towr_b200_batch_size(const towr_b200_batch* b);
/* per-problem constants the kernels read: mass, inertia, terrain id (from `s`; structure must match) */
int towr_b200_batch_set_problem(towr_b200_batch* b, int i, const towr_b200_spec* s);
/* same constants for all problems */
int towr_b200_batch_set_all_problems(towr_b200_batch* b, const towr_b200_spec* s);

/* Host-buffer evaluation (what an Ipopt callback sees).  problem_ids may be NULL (= 0..n-1).
 * x: n x n_vars row-major.  g_out: n x n_rows, jac_values_out: n x nnz (with TOWR_B200_EVAL_JAC_DYNAMIC only its
 * x-dependent positions are written); each may be NULL when `what` does not ask for it.  Copies H2D, launches, copies D2H,
 * synchronises; large batches go through in chunks on two streams so that one chunk's D2H overlaps the next chunk's
 * H2D and kernel. */
int towr_b200_batch_eval(towr_b200_batch* b, const int* problem_ids, int n, int what, const double* x,
                         double* g_out, double* jac_values_out);
/* Device-buffer evaluation of problems 0..n-1: pointers are device memory on the batch's GPU,
 * same shapes; `stream` is a cudaStream_t (NULL = CUDA's default stream).  Asynchronous: the launch is ordered
 * on `stream` only. */
int towr_b200_batch_eval_device(towr_b200_batch* b, int n, int what, const double* d_x, double* d_g,
                                double* d_jac_values, void* stream);
/* Entries of the Jacobian that depend on x.  The other stored values are constants of a problem: Hermite basis values,
 * +-1, the mass times a basis value (what the reference recomputes in every FillJacobianBlock, e.g.
 * src/constraints/spline_acc_constraint.cc:66-80, swing_constraint.cc:86-108, the linear rows of
 * single_rigid_body_dynamics.cc:104-122) and, when EVERY problem of the batch is on a terrain without slope (Flat,
 * Stairs, FiveStepStairs), the terrain-basis entries of the force / torque / terrain constraints as well
 * (force_constraint_discretized.cc:146-152: b_i . J_f with constant b_i).  ANYmal trot on flat ground: 49 % of the
 * entries.  A caller on the host interface fills its n x nnz value arrays with them once
 * (towr_b200_batch_constant_values) and then evaluates with TOWR_B200_EVAL_JAC_DYNAMIC: every eval rewrites only the
 * dynamic entries, in place - one strided device-to-host copy per run of consecutive dynamic entries (61 runs for ANYmal
 * trot) - so the constants never cross PCIe again.  The index depends on the terrains of the batch's problems: fill the
 * constants after the last towr_b200_batch_set_problem. */
int towr_b200_batch_n_dynamic(const towr_b200_batch* b);
/* entries per problem an EVAL_JAC_DYNAMIC eval actually copies device -> host: n_dynamic plus the constants between
 * dynamic runs that are closer than 64 entries (they travel along so that the strided copies have long rows) */
int towr_b200_batch_dynamic_transfer_entries(const towr_b200_batch* b);
/* ascending storage positions (indices into the nnz stored entries) of the dynamic entries; n_dynamic ints */
int towr_b200_batch_dynamic_index(const towr_b200_batch* b, int* index);
/* values[nnz]: the constant entries of problem i, 0.0 at the dynamic positions (host, set-up only) */
int towr_b200_batch_constant_values(const towr_b200_batch* b, int i, double* values);
/* Per-problem status flags, accumulated over every eval since the last call of this function (which reads and clears
 * them): bit 1 (value 2) = a non-finite number in the problem's x, spline nodes or durations, or in one of its stored
 * Jacobian values.  g is not scanned separately: with EVAL_G alone a finite x that produces a non-finite g (a phase
 * duration of 0 gives 1/T^2 = inf) is only flagged through the Jacobian pass (EVAL_JAC / EVAL_BOTH); bit 0 is reserved.
 * Indexed by launch slot: with problem_ids, flag i belongs to the i-th requested problem.  The read-out is ordered after
 * all evals issued through this batch, including towr_b200_batch_eval_device launches on caller streams. */
int towr_b200_batch_status(towr_b200_batch* b, int n, int* status_out);
/* kernel launches issued by this batch so far (for bench.py's gpu_launches) */
long long towr_b200_batch_launch_count(const towr_b200_batch* b);
/* page-locked (pinned) host memory, visible to every device: host buffers handed to towr_b200_batch_eval copy at the full
 * PCIe rate and asynchronously only from such memory (the callback coalescer stages its requests in it) */
int towr_b200_host_alloc(size_t bytes, void** out);
void towr_b200_host_free(void* p);
/* Host placement for multi-GPU jobs (one process or solver-thread group per GPU): the NUMA node the device hangs off
 * (/sys/bus/pci/devices/<bus id>/numa_node; -1 when the platform does not say), and a call that restricts the CALLING
 * thread - and the threads and pinned allocations it makes afterwards - to that node's CPUs (sched_setaffinity).  Returns
 * the number of CPUs bound to, 0 when there is nothing to bind to (single node, no information), < 0 on error.  Without
 * it every rank's staging buffers land on the node the process happened to start on and eight GPUs share one socket's
 * memory bandwidth (round 1: end-to-end scaling efficiency 0.22 at N = 8).  No counterpart in the reference (CPU only). */
int towr_b200_device_numa_node(int device);
int towr_b200_bind_thread_to_device_node(int device);
/* human-readable description of the kernel instantiation and launch geometry this batch uses (for logs / bench.py) */
int towr_b200_batch_kernel_info(const towr_b200_batch* b, char* buf, int n);
/* the batch's own non-blocking cudaStream_t (the one towr_b200_batch_eval works on) */
void* towr_b200_batch_stream(const towr_b200_batch* b);

/* ---- the other two Ipopt callbacks: eval_f / eval_grad_f --------------------------------------------------------------
 * Replaces ifopt::Problem::EvaluateCostFunction / EvaluateCostFunctionGradient over the cost terms of
 * NlpFormulation::GetCosts (towr/src/nlp_formulation.cc:605-681): the NodeCost families (towr/src/costs/node_cost.cc:
 * 36-76) ForcesCostID and EEMotionCostID, as test/hopper_example.cc:147-148 enables them, and the time-discretised
 * EnergyCostID (towr/src/costs/energy_cost.cc: weight dt sum_t sum_ee |f|^2 + torque_weight |tau|^2, with the gradient
 * wrt force / torque nodes and, for optimised durations, the ee-schedule) and AngMomCostID
 * (towr/src/costs/angular_momentum_cost.cc: weight dt sum_t |R I_b R^T omega|^2, gradient wrt base-ang nodes; I_b is the
 * batch problem's inertia), sampled with spec.dt_cost_energy / dt_cost_ang_mom.  spec.n_costs = 0 (the reference's
 * default, parameters.cc:90) gives f = 0 and a zero gradient.  what: bit 0 = f (n doubles), bit 1 = dense gradient
 * (n x n_vars doubles).  spec.enable_swing_ee_base_pos_tracking adds one EEBasePosCost per foot
 * (towr/src/costs/ee_base_pos_cost.cc) after the listed costs, as GetCosts does: every cost term of the reference is built. */
int towr_b200_batch_eval_cost(towr_b200_batch* b, int n, int what, const double* x, double* f_out, double* grad_out);
int towr_b200_batch_eval_cost_device(towr_b200_batch* b, int n, int what, const double* d_x, double* d_f, double* d_grad,
                                     void* stream);

/* Allocation-free eval_f / eval_grad_f.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  Same kernel and results as towr_b200_batch_eval_cost_device, but it needs no batch and performs no cudaMalloc /
 * cudaFree or any other device allocation: the kernel's work tables (column pointers and weights of the class's cost
 * terms) go into a fixed file-scope device array of the library (about 208 KB, held by every GPU that loads the library
 * until the process exits) by two cudaMemcpyAsync on `stream`.  Limits: n_vars <= TOWR_B200_STATIC_COST_MAX_VARS and at
 * most TOWR_B200_STATIC_COST_MAX_WEIGHTS (variable, node value) weights, and at most
 * TOWR_B200_STATIC_COST_MAX_SAMPLED_BYTES of EnergyCost / AngularMomentumCost tables, else TOWR_B200_ERR_ARG with a
 * message naming the limit.  Without a batch there are no per-problem constants: AngularMomentumCost uses the inertia
 * and EEBasePosCost the initial base / foot positions of the spec the structure was created from.  d_x (n x n_vars), d_f (n) and d_grad (n x n_vars) are device memory on `device`, as for the _device entry
 * point.  ALL calls share the one array, which keeps the last call's tables (each call rewrites the part it uses;
 * nothing in it has to start at zero): calls that overlap on the GPU - different streams, or threads - race; order
 * them on one stream. */
#define TOWR_B200_STATIC_COST_MAX_VARS 4096
#define TOWR_B200_STATIC_COST_MAX_WEIGHTS 16384
#define TOWR_B200_STATIC_COST_MAX_SAMPLED_BYTES 65536   /* class tables of EnergyCost / AngularMomentumCost */
int towr_b200_eval_cost_static(const towr_b200_structure* st, int device, int n, int what, const double* d_x, double* d_f,
                               double* d_grad, void* stream);

/* ---- trajectory sampling: the step on the far side of the solver -------------------------------------------------
 * Replaces towr::SaveTrajectoryToCSV's sampling loop (towr/src/utils/save_data.cpp:54-119), i.e. SplineHolder's
 * NodeSpline::GetPoint (towr/src/variables/spline.cc:80-93, towr/include/towr/variables/spline_holder.h:48-84) and
 * PhaseDurations::IsContactPhase (towr/src/variables/phase_durations.cc:113-118), for a whole batch of solution
 * vectors x.  One row per sample time t = 0, dt, 2 dt, ... <= T + 1e-9 (repeated addition, save_data.cpp:55,118):
 *   time | base pos, vel, acc | base euler, rates, accelerations | per foot: pos, vel, acc, euler, rates, accelerations,
 *   contact force, contact torque, is_contact                       (19 + 25 n_ee columns, save_data.cpp:27-46) */
int towr_b200_trajectory_width(const towr_b200_structure* st);
int towr_b200_trajectory_samples(const towr_b200_structure* st, double dt);          /* rows per problem, < 0 on error */
/* host buffers: x is n x n_vars, rows_out is n x samples x width, both row-major */
int towr_b200_batch_sample_trajectory(towr_b200_batch* b, int n, double dt, const double* x, double* rows_out);
/* device buffers already on the batch's device; stream NULL = the default stream */
int towr_b200_batch_sample_trajectory_device(towr_b200_batch* b, int n, double dt, const double* d_x, double* d_rows,
                                             void* stream);
/* SplineHolder::<spline>->GetPoint(t) / PhaseDurations::IsContactPhase(t) at ARBITRARY global times (spline_holder.h:48-84,
 * spline.cc:80-93, phase_durations.cc:113-118): the same rows as _sample_trajectory for the n_times times of `times`
 * (each in [0, total time]; ERR_ARG otherwise), rows_out n x n_times x width.  Tables for the list are built per call. */
int towr_b200_batch_sample_points(towr_b200_batch* b, int n, const double* times, int n_times, const double* x,
                                  double* rows_out);
/* one problem's rows in the reference's CSV format (header line, std::fixed, 6 decimals; save_data.cpp:24-46,68-116) */
int towr_b200_trajectory_write_csv(const towr_b200_structure* st, const double* rows, int n_rows, const char* filename);

#ifdef __cplusplus
}
#endif
#endif /* TOWR_B200_H_ */
