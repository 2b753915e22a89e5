"""ORACLE — TEST INFRASTRUCTURE ONLY.

ctypes binding of oracle/_build/liboracle.so (oracle_capi.h).  May be imported only by
tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "liboracle.so")

MAX_EE, MAX_PHASES, MAX_CONSTRAINTS = 4, 32, 16

# towr::Parameters::ConstraintName (parameters.h:142-153 order)
Dynamic, EndeffectorRom, TotalTime, Terrain, TerrainHard, Force, Torque, Swing, BaseRom, BaseAcc, BaseHeight = range(11)
Monoped, Biped, Hyq, Anymal = range(4)
FlatID, BlockID, StairsID, GapID, SlopeID, ChimneyID, ChimneyLRID, FiveStepStairsID, TableTerrainID = range(9)
TABLE_STEPS, TABLE_RAMPS = 0, 1


class Spec(C.Structure):
    _fields_ = [
        ("n_ee", C.c_int),
        ("mass", C.c_double),
        ("inertia", C.c_double * 6),
        ("nominal_stance", (C.c_double * 3) * MAX_EE),
        ("max_dev", (C.c_double * 3) * MAX_EE),
        ("min_dev", (C.c_double * 3) * MAX_EE),
        ("terrain_id", C.c_int),
        ("n_phases", C.c_int * MAX_EE),
        ("phase_durations", (C.c_double * MAX_PHASES) * MAX_EE),
        ("in_contact_at_start", C.c_int * MAX_EE),
        ("n_constraints", C.c_int),
        ("constraints", C.c_int * MAX_CONSTRAINTS),
        ("dt_constraint_dynamic", C.c_double),
        ("dt_constraint_range_of_motion", C.c_double),
        ("dt_constraint_base_motion", C.c_double),
        ("dt_constraint_force", C.c_double),
        ("dt_constraint_torque", C.c_double),
        ("duration_base_polynomial", C.c_double),
        ("ee_polynomials_per_swing_phase", C.c_int),
        ("force_polynomials_per_stance_phase", C.c_int),
        ("torque_polynomials_per_stance_phase", C.c_int),
        ("force_limit_in_normal_direction", C.c_double),
        ("torque_tx_min", C.c_double), ("torque_tx_max", C.c_double),
        ("torque_ty_min", C.c_double), ("torque_ty_max", C.c_double),
        ("torque_k_friction", C.c_double),
        ("bound_phase_duration", C.c_double * 2),
        ("bounds_final_lin_pos", C.c_int * 3), ("bounds_final_lin_vel", C.c_int * 3),
        ("bounds_final_ang_pos", C.c_int * 3), ("bounds_final_ang_vel", C.c_int * 3),
        ("initial_base_lin", C.c_double * 6), ("initial_base_ang", C.c_double * 6),
        ("final_base_lin", C.c_double * 6), ("final_base_ang", C.c_double * 6),
        ("initial_ee_W", (C.c_double * 3) * MAX_EE),
        ("n_costs", C.c_int),
        ("cost_ids", C.c_int * 4),
        ("cost_weights", C.c_double * 4),
        ("enable_stance_tracking", C.c_int),
        ("n_stance_positions", C.c_int * MAX_EE),
        ("ee_stance_position", ((C.c_double * 2) * 17) * MAX_EE),
        ("terrain_table_n", C.c_int), ("terrain_table_mode", C.c_int),
        ("terrain_table_x", C.c_double * 16), ("terrain_table_h", C.c_double * 16),
        ("dt_cost_energy", C.c_double), ("energy_cost_torque_weight", C.c_double), ("dt_cost_ang_mom", C.c_double),
        ("enable_swing_ee_base_pos_tracking", C.c_int), ("swing_ee_base_pos_tracking_weight", C.c_double),
        ("dt_cost_swing_ee_base_pos_tracking", C.c_double),
        ("n_node_costs", C.c_int), ("node_cost_set", C.c_int * 32), ("node_cost_ee", C.c_int * 32),
        ("node_cost_deriv", C.c_int * 32), ("node_cost_dim", C.c_int * 32), ("node_cost_weight", C.c_double * 32),
        ("n_ee_linear", C.c_int), ("ee_linear_target", C.c_int * 4), ("ee_linear_deriv", C.c_int * 4),
        ("ee_linear_n_terms", C.c_int * 4), ("ee_linear_term_ee", (C.c_int * 4) * 4), ("ee_linear_term_dim", (C.c_int * 4) * 4),
        ("ee_linear_term_coeff", (C.c_double * 4) * 4), ("ee_linear_tolerance", C.c_double * 4), ("ee_linear_dt", C.c_double * 4),
    ]


def build():
    """Compile the C++ restatement (make in oracle/)."""
    env = dict(os.environ)
    env.pop("CXX", None)
    subprocess.check_call(["make", "-C", _HERE], env=env, stdout=subprocess.DEVNULL)


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB_PATH):
        build()
    L = C.CDLL(_LIB_PATH)
    dp, ip, sp = C.POINTER(C.c_double), C.POINTER(C.c_int), C.POINTER(Spec)
    L.orc_spec_defaults.argtypes = [sp]
    L.orc_spec_robot.argtypes = [sp, C.c_int]
    L.orc_spec_gait.argtypes = [sp, C.c_int, C.c_double]
    L.orc_spec_nominal_states.argtypes = [sp, C.c_double, C.c_double, C.c_double]
    L.orc_create.argtypes = [sp]
    L.orc_create.restype = C.c_void_p
    L.orc_destroy.argtypes = [C.c_void_p]
    L.orc_last_error.restype = C.c_char_p
    for f in ("orc_n_vars", "orc_n_rows", "orc_n_var_sets", "orc_n_con_sets"):
        getattr(L, f).argtypes = [C.c_void_p]
    for f in ("orc_var_set_name", "orc_con_set_name"):
        getattr(L, f).argtypes = [C.c_void_p, C.c_int]
        getattr(L, f).restype = C.c_char_p
    for f in ("orc_var_set_rows", "orc_con_set_rows"):
        getattr(L, f).argtypes = [C.c_void_p, C.c_int]
    L.orc_x0.argtypes = [C.c_void_p, dp]
    L.orc_var_bounds.argtypes = [C.c_void_p, dp, dp]
    L.orc_con_bounds.argtypes = [C.c_void_p, dp, dp]
    L.orc_eval_g.argtypes = [C.c_void_p, dp, dp]
    L.orc_nnz.argtypes = [C.c_void_p, dp]
    L.orc_eval_jac.argtypes = [C.c_void_p, dp, ip, ip, dp]
    L.orc_batch_eval.argtypes = [sp, C.c_int, dp, dp, dp, C.c_int]
    L.orc_pool_create.argtypes = [sp, C.c_int]
    L.orc_pool_create.restype = C.c_void_p
    L.orc_pool_destroy.argtypes = [C.c_void_p]
    L.orc_pool_threads.argtypes = [C.c_void_p]
    L.orc_pool_nnz.argtypes = [C.c_void_p]
    L.orc_pool_eval.argtypes = [C.c_void_p, C.c_int, dp, dp, dp]
    L.orc_debug_spline.argtypes = [C.c_void_p, dp, C.c_int, C.c_int, C.c_double, dp]
    L.orc_set_pow_mode.argtypes = [C.c_int]
    L.orc_pow_compare.argtypes = [C.c_double, C.c_int, dp, dp, dp]
    L.orc_hermite_point.argtypes = [dp, dp, dp, dp, C.c_double, C.c_double, dp, dp, dp]
    L.orc_hermite_basis.argtypes = [C.c_double, C.c_double, dp]
    L.orc_hermite_dpos_dT.argtypes = [dp, dp, dp, dp, C.c_double, C.c_double, dp]
    L.orc_terrain.argtypes = [C.c_int, C.c_double, C.c_double, dp, dp, dp, dp]
    L.orc_sample_times.argtypes = [C.c_double, C.c_double, dp, ip]
    L.orc_trajectory_width.argtypes = [C.c_void_p]
    L.orc_eval_f.argtypes = [C.c_void_p, dp]
    L.orc_eval_f.restype = C.c_double
    L.orc_eval_grad_f.argtypes = [C.c_void_p, dp, dp]
    L.orc_sample_trajectory.argtypes = [C.c_void_p, dp, C.c_double, dp]
    _lib = L
    return L


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double)) if a is not None else None


def _ip(a):
    return a.ctypes.data_as(C.POINTER(C.c_int)) if a is not None else None


def make_spec(robot, combo, T, goal=(1.0, 0.0, 0.0), terrain=FlatID, optimize_durations=False, extra_constraints=()):
    """NlpFormulation set-up in the style of towr_ros_app.cc:47-117."""
    L = lib()
    s = Spec()
    L.orc_spec_defaults(C.byref(s))
    if L.orc_spec_robot(C.byref(s), robot) != 0:
        raise RuntimeError(L.orc_last_error().decode())
    if L.orc_spec_gait(C.byref(s), combo, T) != 0:
        raise RuntimeError(L.orc_last_error().decode())
    L.orc_spec_nominal_states(C.byref(s), *goal)
    s.terrain_id = terrain
    for c in extra_constraints:
        s.constraints[s.n_constraints] = c
        s.n_constraints += 1
    if optimize_durations:
        s.constraints[s.n_constraints] = TotalTime
        s.n_constraints += 1
    return s


class Problem:
    """One NLP instance of the oracle (ifopt::Problem restated)."""

    def __init__(self, spec):
        self.spec = spec
        self._L = lib()
        self._p = self._L.orc_create(C.byref(spec))
        if not self._p:
            raise RuntimeError(self._L.orc_last_error().decode())
        self.n_vars = self._L.orc_n_vars(self._p)
        self.n_rows = self._L.orc_n_rows(self._p)

    def __del__(self):
        if getattr(self, "_p", None):
            self._L.orc_destroy(self._p)
            self._p = None

    def var_sets(self):
        return [(self._L.orc_var_set_name(self._p, i).decode(), self._L.orc_var_set_rows(self._p, i))
                for i in range(self._L.orc_n_var_sets(self._p))]

    def con_sets(self):
        return [(self._L.orc_con_set_name(self._p, i).decode(), self._L.orc_con_set_rows(self._p, i))
                for i in range(self._L.orc_n_con_sets(self._p))]

    def x0(self):
        x = np.empty(self.n_vars)
        self._L.orc_x0(self._p, _dp(x))
        return x

    def var_bounds(self):
        lo, up = np.empty(self.n_vars), np.empty(self.n_vars)
        self._L.orc_var_bounds(self._p, _dp(lo), _dp(up))
        return lo, up

    def con_bounds(self):
        lo, up = np.empty(self.n_rows), np.empty(self.n_rows)
        self._L.orc_con_bounds(self._p, _dp(lo), _dp(up))
        return lo, up

    def sample_trajectory(self, x, dt):
        """Rows of SaveTrajectoryToCSV (utils/save_data.cpp) for the solution vector x: (n_samples, 19 + 25 n_ee)."""
        x = np.ascontiguousarray(x, dtype=np.float64)
        w = self._L.orc_trajectory_width(self._p)
        n = self._L.orc_sample_trajectory(self._p, _dp(x), float(dt), None)
        rows = np.empty((n, w))
        self._L.orc_sample_trajectory(self._p, _dp(x), float(dt), _dp(rows))
        return rows

    def eval_f(self, x):
        x = np.ascontiguousarray(x, dtype=np.float64)
        return float(self._L.orc_eval_f(self._p, _dp(x)))

    def eval_grad_f(self, x):
        x = np.ascontiguousarray(x, dtype=np.float64)
        grad = np.empty(self.n_vars)
        self._L.orc_eval_grad_f(self._p, _dp(x), _dp(grad))
        return grad

    def eval_g(self, x):
        x = np.ascontiguousarray(x, dtype=np.float64)
        g = np.empty(self.n_rows)
        self._L.orc_eval_g(self._p, _dp(x), _dp(g))
        return g

    def eval_jac(self, x):
        x = np.ascontiguousarray(x, dtype=np.float64)
        n = self._L.orc_nnz(self._p, _dp(x))
        r, c, v = np.empty(n, np.int32), np.empty(n, np.int32), np.empty(n)
        self._L.orc_eval_jac(self._p, _dp(x), _ip(r), _ip(c), _dp(v))
        return r, c, v


def batch_eval(spec, x, nnz, want_g=True, want_jac=True, n_threads=0):
    x = np.ascontiguousarray(x, dtype=np.float64)
    n, nv = x.shape
    p = Problem(spec)
    assert nv == p.n_vars
    g = np.empty((n, p.n_rows)) if want_g else None
    v = np.empty((n, nnz)) if want_jac else None
    rc = lib().orc_batch_eval(C.byref(spec), n, _dp(x), _dp(g), _dp(v), n_threads)
    if rc != 0:
        raise RuntimeError("oracle pattern size changed across the batch")
    return g, v


class Pool:
    """One oracle object graph per host thread, built once (orc_pool): the timed CPU baseline evaluates through it."""

    def __init__(self, spec, n_threads):
        self._L = lib()
        self._p = self._L.orc_pool_create(C.byref(spec), int(n_threads))
        if not self._p:
            raise RuntimeError(self._L.orc_last_error().decode())
        self.n_threads = self._L.orc_pool_threads(self._p)
        self.nnz = self._L.orc_pool_nnz(self._p)
        probe = Problem(spec)
        self.n_vars, self.n_rows = probe.n_vars, probe.n_rows
        self._x0 = probe.x0()

    def x0(self):
        return self._x0.copy()

    def __del__(self):
        if getattr(self, "_p", None):
            self._L.orc_pool_destroy(self._p)
            self._p = None

    def eval(self, x, g=None, vals=None):
        x = np.ascontiguousarray(x, dtype=np.float64)
        n = x.shape[0]
        assert x.shape[1] == self.n_vars
        g = np.empty((n, self.n_rows)) if g is None else g
        vals = np.empty((n, self.nnz)) if vals is None else vals
        if self._L.orc_pool_eval(self._p, n, _dp(x), _dp(g), _dp(vals)) != 0:
            raise RuntimeError("oracle pattern size changed across the batch")
        return g, vals


class pow_mode:
    """with pow_mode(1): the oracle's PhaseSpline polynomials use correctly rounded integer powers (orc_set_pow_mode)."""

    def __init__(self, mode):
        self.mode = mode

    def __enter__(self):
        self.old = lib().orc_get_pow_mode()
        lib().orc_set_pow_mode(self.mode)

    def __exit__(self, *a):
        lib().orc_set_pow_mode(self.old)
