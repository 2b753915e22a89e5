"""towr2025_b200 — batched B200 evaluation backend for TOWR's NLP (constraint values + sparse-Jacobian values).

Python here is plumbing only: a ctypes view of the C ABI in include/towr_b200.h (libtowr_b200.so, built from
towr2025_b200/csrc by `make` / build()).  The product is the CUDA library; there is no CPU evaluation path —
every eval call fails loudly when the library or a GPU is missing.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
LIB_PATH = os.environ.get("TOWR_B200_LIB") or os.path.join(_HERE, "libtowr_b200.so")   # override: experiments only
CSRC = os.path.join(_HERE, "csrc")

MAX_EE, MAX_PHASES, MAX_CONSTRAINTS = 4, 32, 16

# towr::Parameters::ConstraintName
Dynamic, EndeffectorRom, TotalTime, Terrain, TerrainHard, Force, Torque, Swing, BaseRom, BaseAcc, BaseHeight = range(11)
Monoped, Biped, Hyq, Anymal = range(4)
FlatID, BlockID, StairsID, GapID, SlopeID, ChimneyID, ChimneyLRID, FiveStepStairsID, TableTerrainID = range(9)
TABLE_STEPS, TABLE_RAMPS = 0, 1
EVAL_G, EVAL_JAC, EVAL_BOTH, EVAL_JAC_DYNAMIC = 1, 2, 3, 4
C0, C1, C2, C3, C4 = range(5)


class Spec(C.Structure):
    """towr_b200_spec (include/towr_b200.h)."""
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

    def copy(self):
        s = Spec()
        C.memmove(C.byref(s), C.byref(self), C.sizeof(Spec))
        return s


class TowrB200Error(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("towr_b200 error %d: %s" % (code, msg))
        self.code = code


def build(verbose=False):
    """Compile libtowr_b200.so in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    env = dict(os.environ)
    out = None if verbose else subprocess.DEVNULL
    subprocess.check_call(["make", "-C", CSRC, "-j4"], env=env, stdout=out)
    return LIB_PATH


_lib = None

_EXPORTS = """towr_b200_last_error towr_b200_device_count towr_b200_spec_defaults towr_b200_spec_robot towr_b200_spec_gait
towr_b200_spec_nominal_states towr_b200_spec_optimize_durations towr_b200_structure_create towr_b200_structure_destroy
towr_b200_structure_n_vars towr_b200_structure_n_rows towr_b200_structure_nnz towr_b200_structure_n_var_sets
towr_b200_structure_n_con_sets towr_b200_structure_var_set_name towr_b200_structure_var_set_rows
towr_b200_structure_con_set_name towr_b200_structure_con_set_rows towr_b200_structure_pattern towr_b200_structure_accepts
towr_b200_problem_x0 towr_b200_problem_var_bounds towr_b200_problem_con_bounds towr_b200_batch_create
towr_b200_batch_destroy towr_b200_batch_size towr_b200_batch_set_problem towr_b200_batch_set_all_problems
towr_b200_batch_eval towr_b200_batch_eval_device towr_b200_batch_status towr_b200_batch_launch_count
towr_b200_batch_stream towr_b200_trajectory_width towr_b200_trajectory_samples towr_b200_batch_sample_trajectory
towr_b200_batch_sample_trajectory_device towr_b200_trajectory_write_csv towr_b200_batch_eval_cost
towr_b200_batch_eval_cost_device towr_b200_eval_cost_static towr_b200_batch_n_dynamic
towr_b200_batch_dynamic_index towr_b200_batch_constant_values towr_b200_batch_dynamic_transfer_entries towr_b200_batch_kernel_info towr_b200_host_alloc towr_b200_host_free
towr_b200_device_numa_node towr_b200_bind_thread_to_device_node towr_b200_batch_sample_points""".split()


def lib():
    """The loaded C-ABI library.  Raises if it has not been built: there is no fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise TowrB200Error(-3, "libtowr_b200.so is not built (run towr2025_b200.build() or make -C towr2025_b200/csrc)")
    L = C.CDLL(LIB_PATH)
    dp, ip, sp, vp = C.POINTER(C.c_double), C.POINTER(C.c_int), C.POINTER(Spec), C.c_void_p
    L.towr_b200_last_error.restype = C.c_char_p
    L.towr_b200_spec_defaults.argtypes = [sp]
    L.towr_b200_spec_robot.argtypes = [sp, C.c_int]
    L.towr_b200_spec_gait.argtypes = [sp, C.c_int, C.c_double]
    L.towr_b200_spec_nominal_states.argtypes = [sp, C.c_double, C.c_double, C.c_double]
    L.towr_b200_spec_optimize_durations.argtypes = [sp]
    L.towr_b200_structure_create.argtypes = [sp, C.POINTER(vp)]
    L.towr_b200_structure_destroy.argtypes = [vp]
    L.towr_b200_structure_destroy.restype = None
    for f in ("n_vars", "n_rows", "nnz", "n_var_sets", "n_con_sets"):
        getattr(L, "towr_b200_structure_" + f).argtypes = [vp]
    for f in ("var_set_name", "con_set_name"):
        getattr(L, "towr_b200_structure_" + f).argtypes = [vp, C.c_int]
        getattr(L, "towr_b200_structure_" + f).restype = C.c_char_p
    for f in ("var_set_rows", "con_set_rows"):
        getattr(L, "towr_b200_structure_" + f).argtypes = [vp, C.c_int]
    L.towr_b200_structure_pattern.argtypes = [vp, ip, ip]
    L.towr_b200_batch_n_dynamic.argtypes = [vp]
    L.towr_b200_batch_dynamic_transfer_entries.argtypes = [vp]
    L.towr_b200_batch_dynamic_index.argtypes = [vp, ip]
    L.towr_b200_batch_constant_values.argtypes = [vp, C.c_int, dp]
    L.towr_b200_structure_accepts.argtypes = [vp, sp]
    L.towr_b200_problem_x0.argtypes = [vp, sp, dp]
    L.towr_b200_problem_var_bounds.argtypes = [vp, sp, dp, dp]
    L.towr_b200_problem_con_bounds.argtypes = [vp, sp, dp, dp]
    L.towr_b200_batch_create.argtypes = [vp, C.c_int, C.c_int, C.POINTER(vp)]
    L.towr_b200_batch_destroy.argtypes = [vp]
    L.towr_b200_batch_destroy.restype = None
    L.towr_b200_batch_size.argtypes = [vp]
    L.towr_b200_batch_set_problem.argtypes = [vp, C.c_int, sp]
    L.towr_b200_batch_set_all_problems.argtypes = [vp, sp]
    L.towr_b200_batch_eval.argtypes = [vp, ip, C.c_int, C.c_int, dp, dp, dp]
    L.towr_b200_batch_eval_device.argtypes = [vp, C.c_int, C.c_int, vp, vp, vp, vp]
    L.towr_b200_batch_status.argtypes = [vp, C.c_int, ip]
    L.towr_b200_batch_eval_cost.argtypes = [vp, C.c_int, C.c_int, dp, dp, dp]
    L.towr_b200_batch_eval_cost_device.argtypes = [vp, C.c_int, C.c_int, vp, vp, vp, vp]
    L.towr_b200_eval_cost_static.argtypes = [vp, C.c_int, C.c_int, C.c_int, vp, vp, vp, vp]
    L.towr_b200_trajectory_width.argtypes = [vp]
    L.towr_b200_trajectory_samples.argtypes = [vp, C.c_double]
    L.towr_b200_batch_sample_trajectory.argtypes = [vp, C.c_int, C.c_double, dp, dp]
    L.towr_b200_batch_sample_trajectory_device.argtypes = [vp, C.c_int, C.c_double, vp, vp, vp]
    if hasattr(L, "towr_b200_batch_sample_points"):      # absent only in older builds loaded through TOWR_B200_LIB (A/B runs)
        L.towr_b200_batch_sample_points.argtypes = [vp, C.c_int, dp, C.c_int, dp, dp]
    L.towr_b200_trajectory_write_csv.argtypes = [vp, dp, C.c_int, C.c_char_p]
    L.towr_b200_batch_kernel_info.argtypes = [vp, C.c_char_p, C.c_int]
    L.towr_b200_batch_launch_count.argtypes = [vp]
    L.towr_b200_batch_launch_count.restype = C.c_longlong
    L.towr_b200_batch_stream.argtypes = [vp]
    L.towr_b200_batch_stream.restype = C.c_void_p
    _lib = L
    return L


def exported_symbols():
    return list(_EXPORTS)


def device_numa_node(device=0):
    """NUMA node of the GPU (towr_b200_device_numa_node); -1 when the platform does not say."""
    L = lib()
    L.towr_b200_device_numa_node.argtypes = [C.c_int]
    return int(L.towr_b200_device_numa_node(int(device)))


def bind_thread_to_device_node(device=0):
    """Restrict the calling thread (and what it starts or pins afterwards) to the CPUs of the GPU's NUMA node
    (towr_b200_bind_thread_to_device_node).  Returns the number of CPUs bound to, 0 when there is nothing to bind to."""
    L = lib()
    L.towr_b200_bind_thread_to_device_node.argtypes = [C.c_int]
    rc = int(L.towr_b200_bind_thread_to_device_node(int(device)))
    if rc < 0:
        raise TowrB200Error(rc, L.towr_b200_last_error().decode())
    return rc


def _check(rc):
    if rc != 0:
        raise TowrB200Error(rc, lib().towr_b200_last_error().decode())


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double)) if a is not None else None


def _ip(a):
    return a.ctypes.data_as(C.POINTER(C.c_int)) if a is not None else None


def make_spec(robot, combo, total_time, goal=(1.0, 0.0, 0.0), terrain=FlatID, optimize_durations=False,
              extra_constraints=()):
    """An NlpFormulation set up the way towr_ros_app.cc:47-117 does it."""
    L = lib()
    s = Spec()
    _check(L.towr_b200_spec_defaults(C.byref(s)))
    _check(L.towr_b200_spec_robot(C.byref(s), robot))
    _check(L.towr_b200_spec_gait(C.byref(s), combo, total_time))
    _check(L.towr_b200_spec_nominal_states(C.byref(s), *goal))
    s.terrain_id = terrain
    for c in extra_constraints:
        s.constraints[s.n_constraints] = c
        s.n_constraints += 1
    if optimize_durations:
        _check(L.towr_b200_spec_optimize_durations(C.byref(s)))
    return s


class Structure:
    """A structure class: sizes, names and the device-independent sparsity pattern."""

    def __init__(self, spec):
        self._L = lib()
        h = C.c_void_p()
        _check(self._L.towr_b200_structure_create(C.byref(spec), C.byref(h)))
        self._h = h
        self.n_vars = self._L.towr_b200_structure_n_vars(h)
        self.n_rows = self._L.towr_b200_structure_n_rows(h)
        self.nnz = self._L.towr_b200_structure_nnz(h)

    def __del__(self):
        if getattr(self, "_h", None):
            self._L.towr_b200_structure_destroy(self._h)
            self._h = None

    def var_sets(self):
        L, h = self._L, self._h
        return [(L.towr_b200_structure_var_set_name(h, i).decode(), L.towr_b200_structure_var_set_rows(h, i))
                for i in range(L.towr_b200_structure_n_var_sets(h))]

    def con_sets(self):
        L, h = self._L, self._h
        return [(L.towr_b200_structure_con_set_name(h, i).decode(), L.towr_b200_structure_con_set_rows(h, i))
                for i in range(L.towr_b200_structure_n_con_sets(h))]

    def trajectory_width(self):
        """Columns of one trajectory row (towr_b200_trajectory_width): 19 + 25 n_ee."""
        return self._L.towr_b200_trajectory_width(self._h)

    def trajectory_samples(self, dt):
        """Rows per problem for the sampling period dt (towr_b200_trajectory_samples)."""
        n = self._L.towr_b200_trajectory_samples(self._h, float(dt))
        if n < 0:
            raise TowrB200Error(-1, self._L.towr_b200_last_error().decode())
        return n

    def write_csv(self, rows, filename):
        """One problem's trajectory rows in the reference's CSV format (towr_b200_trajectory_write_csv)."""
        rows = np.ascontiguousarray(rows, dtype=np.float64)
        _check(self._L.towr_b200_trajectory_write_csv(self._h, _dp(rows), rows.shape[0], filename.encode()))

    def pattern(self):
        r, c = np.empty(self.nnz, np.int32), np.empty(self.nnz, np.int32)
        _check(self._L.towr_b200_structure_pattern(self._h, _ip(r), _ip(c)))
        return r, c

    def accepts(self, spec):
        return self._L.towr_b200_structure_accepts(self._h, C.byref(spec)) == 0

    def x0(self, spec):
        x = np.empty(self.n_vars)
        _check(self._L.towr_b200_problem_x0(self._h, C.byref(spec), _dp(x)))
        return x

    def var_bounds(self, spec):
        lo, up = np.empty(self.n_vars), np.empty(self.n_vars)
        _check(self._L.towr_b200_problem_var_bounds(self._h, C.byref(spec), _dp(lo), _dp(up)))
        return lo, up

    def con_bounds(self, spec):
        lo, up = np.empty(self.n_rows), np.empty(self.n_rows)
        _check(self._L.towr_b200_problem_con_bounds(self._h, C.byref(spec), _dp(lo), _dp(up)))
        return lo, up

    def eval_cost_static(self, n, what, d_x, d_f, d_grad, device=0, stream=None):
        """Allocation-free eval_f / eval_grad_f on device pointers (towr_b200_eval_cost_static); asynchronous on `stream`."""
        _check(self._L.towr_b200_eval_cost_static(self._h, device, n, what, C.c_void_p(d_x), C.c_void_p(d_f or 0),
                                                   C.c_void_p(d_grad or 0), C.c_void_p(stream or 0)))

    @property
    def bytes_per_eval(self):
        """Algorithmic HBM bytes of one eval: read x, write g, write every stored Jacobian value."""
        return 8 * (self.n_vars + self.n_rows + self.nnz)


class Batch:
    """n problems of one structure class resident on one GPU."""

    def __init__(self, structure, n_problems, device=0):
        self._L = lib()
        self.structure = structure
        self.n = n_problems
        h = C.c_void_p()
        _check(self._L.towr_b200_batch_create(structure._h, n_problems, device, C.byref(h)))
        self._h = h

    def __del__(self):
        if getattr(self, "_h", None):
            self._L.towr_b200_batch_destroy(self._h)
            self._h = None

    def set_problem(self, i, spec):
        _check(self._L.towr_b200_batch_set_problem(self._h, i, C.byref(spec)))

    def set_all(self, spec):
        _check(self._L.towr_b200_batch_set_all_problems(self._h, C.byref(spec)))

    def eval(self, x, what=EVAL_BOTH, ids=None, g_out=None, jac_out=None):
        """Host-buffer evaluation through the C ABI (copies in, launches, copies out)."""
        x = np.ascontiguousarray(x, dtype=np.float64)
        n = x.shape[0]
        st = self.structure
        assert x.shape[1] == st.n_vars
        g = (g_out if g_out is not None else np.empty((n, st.n_rows))) if what & EVAL_G else None
        v = None
        if what & EVAL_JAC:
            v = jac_out if jac_out is not None else np.empty((n, st.nnz))
        elif what & EVAL_JAC_DYNAMIC:           # only the x-dependent positions of the n x nnz array are written
            if jac_out is None:                 # convenience: start from the constants of problems 0..n-1
                jac_out = np.stack([self.constant_values(i if ids is None else int(ids[i])) for i in range(n)])
            v = jac_out
        idp = None
        if ids is not None:
            ids = np.ascontiguousarray(ids, dtype=np.int32)
            idp = _ip(ids)
        _check(self._L.towr_b200_batch_eval(self._h, idp, n, what, _dp(x), _dp(g), _dp(v)))
        return g, v

    def eval_cost(self, x, what=3):
        """Host-buffer eval_f / eval_grad_f (towr_b200_batch_eval_cost): (f[n], grad[n, n_vars])."""
        x = np.ascontiguousarray(x, dtype=np.float64)
        n = x.shape[0]
        f = np.empty(n) if what & 1 else None
        grad = np.empty((n, self.structure.n_vars)) if what & 2 else None
        _check(self._L.towr_b200_batch_eval_cost(self._h, n, what, _dp(x), _dp(f), _dp(grad)))
        return f, grad

    def sample_trajectory(self, x, dt):
        """Host-buffer trajectory sampling (towr_b200_batch_sample_trajectory): (n, samples, width) rows."""
        x = np.ascontiguousarray(x, dtype=np.float64)
        n = x.shape[0]
        rows = np.empty((n, self.structure.trajectory_samples(dt), self.structure.trajectory_width()))
        _check(self._L.towr_b200_batch_sample_trajectory(self._h, n, float(dt), _dp(x), _dp(rows)))
        return rows

    def sample_points(self, x, times):
        """SplineHolder GetPoint / IsContactPhase rows at arbitrary global times (towr_b200_batch_sample_points):
        (n, len(times), width)."""
        x = np.ascontiguousarray(x, dtype=np.float64)
        times = np.ascontiguousarray(times, dtype=np.float64)
        rows = np.empty((x.shape[0], times.shape[0], self.structure.trajectory_width()))
        _check(self._L.towr_b200_batch_sample_points(self._h, x.shape[0], _dp(times), times.shape[0], _dp(x), _dp(rows)))
        return rows

    def sample_trajectory_device(self, n, dt, d_x, d_rows, stream=None):
        """Device-pointer trajectory sampling, asynchronous on `stream`."""
        _check(self._L.towr_b200_batch_sample_trajectory_device(self._h, n, float(dt), C.c_void_p(d_x), C.c_void_p(d_rows),
                                                                C.c_void_p(stream or 0)))

    def eval_device(self, n, what, d_x, d_g, d_vals, stream=None):
        """Device-pointer evaluation (ints from tensor.data_ptr()); asynchronous on `stream`
        (a cudaStream_t as int; None/0 = CUDA's default stream)."""
        _check(self._L.towr_b200_batch_eval_device(self._h, n, what, C.c_void_p(d_x), C.c_void_p(d_g or 0),
                                                    C.c_void_p(d_vals or 0), C.c_void_p(stream or 0)))

    def eval_cost_device(self, n, what, d_x, d_f, d_grad, stream=None):
        """Device-pointer eval_f / eval_grad_f (towr_b200_batch_eval_cost_device); asynchronous on `stream`."""
        _check(self._L.towr_b200_batch_eval_cost_device(self._h, n, what, C.c_void_p(d_x), C.c_void_p(d_f or 0),
                                                         C.c_void_p(d_grad or 0), C.c_void_p(stream or 0)))

    @property
    def n_dynamic(self):
        """Stored Jacobian entries that depend on x for this batch (towr_b200_batch_n_dynamic)."""
        return self._L.towr_b200_batch_n_dynamic(self._h)

    @property
    def dynamic_transfer_entries(self):
        """Entries per problem an EVAL_JAC_DYNAMIC eval copies back (towr_b200_batch_dynamic_transfer_entries)."""
        return self._L.towr_b200_batch_dynamic_transfer_entries(self._h)

    def dynamic_index(self):
        idx = np.empty(self.n_dynamic, np.int32)
        _check(self._L.towr_b200_batch_dynamic_index(self._h, _ip(idx)))
        return idx

    def constant_values(self, i):
        """nnz values: the x-independent entries of problem i, 0.0 at the dynamic positions."""
        v = np.empty(self.structure.nnz)
        _check(self._L.towr_b200_batch_constant_values(self._h, i, _dp(v)))
        return v

    def status(self, n=None):
        n = self.n if n is None else n
        s = np.empty(n, np.int32)
        _check(self._L.towr_b200_batch_status(self._h, n, _ip(s)))
        return s

    @property
    def kernel_info(self):
        buf = C.create_string_buffer(512)
        _check(self._L.towr_b200_batch_kernel_info(self._h, buf, 512))
        return buf.value.decode()

    @property
    def launches(self):
        return int(self._L.towr_b200_batch_launch_count(self._h))
