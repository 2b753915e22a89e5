"""ORACLE, SECOND LEG — TEST INFRASTRUCTURE ONLY.

An independent restatement of the reference's hot path in numpy / scipy.sparse, written from the reference's Eigen
expressions (file:line cited per function), NOT from the C++ oracle in this directory.  It exists so that
tests/test_oracle_cross.py can check the C++ oracle (values, Jacobian values, stored-entry pattern, x0, bounds) against a
second derivation: the reference's own unit tests are empty stubs, so two restatements that agree are the
strongest pin available here.

May be imported only by tests/.  Slow by design (seconds per evaluation); nothing in the product or bench.py uses it.

Eigen semantics kept on purpose (they decide which Jacobian entries are STORED):
  * a sparse product / sum keeps every structurally reachable entry, also when its value is 0.0;
  * `coeffRef(r, c)` stores the entry whatever the value;
  * `dense.sparseView()` drops exact zeros, `dense.sparseView(1.0, -1.0)` keeps all;
  * `jac.row(i) = expr` / `jac.middleRows(i, n) = expr` replace the rows' stored entries by those of expr.
scipy.sparse prunes numerical zeros in products, so every matrix here is a pair (v, p): v carries the values, p is the
same expression evaluated on 0/1 pattern matrices (non-negative, so nothing ever cancels).
"""
import math

import numpy as np
import scipy.sparse as sp

X, Y, Z = 0, 1, 2
kPos, kVel, kAcc = 0, 1, 2
AX, AY, AZ, LX, LY, LZ = range(6)
INF = 1.0e20                      # ifopt::inf (ifopt/bounds.h)
NO_BOUND = (-INF, INF)
BOUND_ZERO = (0.0, 0.0)
BOUND_GREATER_ZERO = (0.0, INF)
BOUND_SMALLER_ZERO = (-INF, 0.0)


# ------------------------------------------------------------------------------------------------ sparse pair
def _csr(m):
    return m if isinstance(m, sp.csr_matrix) else sp.csr_matrix(m)


def _ones(m):
    m = _csr(m)
    m.data[:] = 1.0
    return m


class Sp:
    """Eigen::SparseMatrix<double, RowMajor> as used by the reference: values v and stored-entry pattern p."""
    __slots__ = ("v", "p")

    def __init__(self, v, p):
        self.v = _csr(v)
        self.p = _ones(p)

    @property
    def shape(self):
        return self.v.shape

    @staticmethod
    def zeros(r, c):
        return Sp(sp.csr_matrix((r, c)), sp.csr_matrix((r, c)))

    @staticmethod
    def from_dense(m, keep_all):
        """keep_all: `.sparseView(1.0, -1.0)`; otherwise `.sparseView()`, which drops exact zeros."""
        m = np.atleast_2d(np.asarray(m, dtype=float))
        pat = np.ones_like(m) if keep_all else (m != 0.0).astype(float)
        return Sp(sp.csr_matrix(m), sp.csr_matrix(pat))

    @staticmethod
    def from_entries(shape, entries):
        """entries: dict (row, col) -> value built by coeffRef accumulation; every key is stored."""
        if not entries:
            return Sp.zeros(*shape)
        rc = np.array(list(entries.keys()), dtype=int)
        vals = np.array(list(entries.values()), dtype=float)
        v = sp.coo_matrix((vals, (rc[:, 0], rc[:, 1])), shape=shape)
        p = sp.coo_matrix((np.ones(len(vals)), (rc[:, 0], rc[:, 1])), shape=shape)
        return Sp(v, p)

    def __matmul__(self, o):
        return Sp(self.v @ o.v, self.p @ o.p)

    def __add__(self, o):
        return Sp(self.v + o.v, self.p + o.p)

    def __sub__(self, o):
        return Sp(self.v - o.v, self.p + o.p)

    def __neg__(self):
        return Sp(-self.v, self.p.copy())

    def __rmul__(self, s):
        return Sp(self.v * float(s), self.p.copy())

    __mul__ = __rmul__

    def T(self):
        return Sp(self.v.T, self.p.T)

    def row(self, i):
        return Sp(self.v[i, :], self.p[i, :])

    @staticmethod
    def vstack(rows):
        return Sp(sp.vstack([r.v for r in rows], format="csr"), sp.vstack([r.p for r in rows], format="csr"))

    def dense(self):
        return np.asarray(self.v.todense())

    def stored(self):
        """(rows, cols, vals) of the stored entries in row-major order."""
        p = self.p.tocoo()
        order = np.lexsort((p.col, p.row))
        r, c = p.row[order], p.col[order]
        if not len(r):
            return r, c, np.zeros(0)
        # v + 0*p has p's pattern or more; read it back at p's positions through a dense row-major key lookup
        v = self.v.tocoo()
        table = dict(zip(zip(v.row.tolist(), v.col.tolist()), v.data.tolist()))
        vals = np.array([table.get(k, 0.0) for k in zip(r.tolist(), c.tolist())])
        return r, c, vals


def cross(v):
    """models/single_rigid_body_dynamics.cc:46-54 — six coeffRef entries, the diagonal is not stored."""
    return Sp.from_entries((3, 3), {(0, 1): -v[2], (0, 2): v[1], (1, 0): v[2], (1, 2): -v[0], (2, 0): -v[1], (2, 1): v[0]})


# ------------------------------------------------------------------------------------------------ polynomial
class CubicHermite:
    """helpers/polynomial.cc:74-257.  std::pow(t, n) is restated as math.pow (the same libm call)."""

    def __init__(self, dim):
        self.dim = dim
        self.T = 0.0
        self.n0 = np.zeros((2, dim))
        self.n1 = np.zeros((2, dim))
        self.coeff = np.zeros((4, dim))

    def update_coeff(self):                                           # polynomial.cc:97-104
        T = self.T
        p0, v0, p1, v1 = self.n0[kPos], self.n0[kVel], self.n1[kPos], self.n1[kVel]
        with np.errstate(divide="ignore", invalid="ignore"):
            self.coeff[0] = p0
            self.coeff[1] = v0
            self.coeff[2] = -(3 * (p0 - p1) + T * (2 * v0 + v1)) / math.pow(T, 2)
            self.coeff[3] = (2 * (p0 - p1) + T * (v0 + v1)) / math.pow(T, 3)

    def point(self, t):                                               # polynomial.cc:47-72
        out = np.zeros((3, self.dim))
        for d in (kPos, kVel, kAcc):
            for c in range(4):
                if d == kPos:
                    w = math.pow(t, c)
                elif d == kVel:
                    w = c * math.pow(t, c - 1) if c >= 1 else 0.0
                else:
                    w = c * (c - 1) * math.pow(t, c - 2) if c >= 2 else 0.0
                out[d] += w * self.coeff[c]
        return out

    def d_wrt_start(self, dxdt, node_deriv, t):                       # polynomial.cc:106-196
        T = self.T
        t2, t3, T2, T3 = math.pow(t, 2), math.pow(t, 3), math.pow(T, 2), math.pow(T, 3)
        if dxdt == kPos:
            return (2 * t3) / T3 - (3 * t2) / T2 + 1 if node_deriv == kPos else t - (2 * t2) / T + t3 / T2
        if dxdt == kVel:
            return (6 * t2) / T3 - (6 * t) / T2 if node_deriv == kPos else (3 * t2) / T2 - (4 * t) / T + 1
        return (12 * t) / T3 - 6 / T2 if node_deriv == kPos else (6 * t) / T2 - 4 / T

    def d_wrt_end(self, dxdt, node_deriv, t):                         # polynomial.cc:122-236
        T = self.T
        t2, t3, T2, T3 = math.pow(t, 2), math.pow(t, 3), math.pow(T, 2), math.pow(T, 3)
        if dxdt == kPos:
            return (3 * t2) / T2 - (2 * t3) / T3 if node_deriv == kPos else t3 / T2 - t2 / T
        if dxdt == kVel:
            return (6 * t) / T2 - (6 * t2) / T3 if node_deriv == kPos else (3 * t2) / T2 - (2 * t) / T
        return 6 / T2 - (12 * t) / T3 if node_deriv == kPos else (6 * t) / T2 - 2 / T

    def dpos_dT(self, t):                                             # polynomial.cc:238-257
        x0, v0, x1, v1 = self.n0[kPos], self.n0[kVel], self.n1[kPos], self.n1[kVel]
        T = self.T
        t2, t3 = math.pow(t, 2), math.pow(t, 3)
        T2, T3, T4 = math.pow(T, 2), math.pow(T, 3), math.pow(T, 4)
        return ((t3 * (v0 + v1)) / T3 - (t2 * (2 * v0 + v1)) / T2 - (3 * t3 * (2 * x0 - 2 * x1 + T * v0 + T * v1)) / T4
                + (2 * t2 * (3 * x0 - 3 * x1 + 2 * T * v0 + T * v1)) / T3)


# ------------------------------------------------------------------------------------------------ variables
class NodesVariables:
    """variables/nodes_variables.cc:44-263.  nodes[id, deriv, dim]; index_info[idx] = [(id, deriv, dim), ...]."""

    def __init__(self, name, n_nodes, dim):
        self.name = name
        self.dim = dim
        self.nodes = np.zeros((n_nodes, 2, dim))
        self.index_info = []
        self.bounds = []
        self.observers = []

    def rows(self):
        return len(self.index_info)

    def set_number_of_variables(self):
        self.bounds = [NO_BOUND] * self.rows()

    def opt_index(self, node, deriv, dim):                            # nodes_variables.cc:44-53 (first match)
        for idx, infos in enumerate(self.index_info):
            for nvi in infos:
                if nvi == (node, deriv, dim):
                    return idx
        return -1

    def values(self):                                                 # :55-64 (the last info of an index wins)
        x = np.zeros(self.rows())
        for idx, infos in enumerate(self.index_info):
            for (n, d, k) in infos:
                x[idx] = self.nodes[n, d, k]
        return x

    def set_variables(self, x):                                       # :66-74
        for idx, infos in enumerate(self.index_info):
            for (n, d, k) in infos:
                self.nodes[n, d, k] = x[idx]
        for o in self.observers:
            o.update_nodes()

    def polynomial_count(self):
        return self.nodes.shape[0] - 1

    def set_by_linear_interpolation(self, initial, final, t_total):  # :123-146
        initial, final = np.asarray(initial, float), np.asarray(final, float)
        dp = final - initial
        average_velocity = dp / t_total
        num_nodes = self.nodes.shape[0]
        for infos in self.index_info:
            for (n, d, k) in infos:
                if d == kPos:
                    pos = initial + n / float(num_nodes - 1) * dp
                    self.nodes[n, kPos, k] = pos[k]
                if d == kVel:
                    self.nodes[n, kVel, k] = average_velocity[k]

    def set_by_linear_interpolation_relative_to_base(self, ee0_W, eeT_W, b0, bT, rpy0, rpyT, t_total):   # :148-203
        ee0_W, eeT_W, b0, bT, rpy0, rpyT = (np.asarray(a, float) for a in (ee0_W, eeT_W, b0, bT, rpy0, rpyT))
        num_nodes = self.nodes.shape[0]
        if num_nodes < 2:
            return
        R0, RT = rotation_base_to_world(rpy0), rotation_base_to_world(rpyT)
        r0_B = R0.T @ (ee0_W - b0)
        rT_B = RT.T @ (eeT_W - bT)
        dp_B = rT_B - r0_B
        avg_vel_B = dp_B / t_total
        base_avg_vel_W = (bT - b0) / t_total
        for infos in self.index_info:
            for (n, d, k) in infos:
                alpha = n / float(num_nodes - 1)
                base_pos_W = (1.0 - alpha) * b0 + alpha * bT
                base_rpy_W = (1.0 - alpha) * rpy0 + alpha * rpyT
                R = rotation_base_to_world(base_rpy_W)
                r_B = r0_B + alpha * dp_B
                if d == kPos:
                    self.nodes[n, kPos, k] = (base_pos_W + R @ r_B)[k]
                if d == kVel:
                    self.nodes[n, kVel, k] = (base_avg_vel_W + R @ avg_vel_B)[k]
        self.set_variables(self.values())

    def add_bound(self, node, deriv, dim, lower, upper):              # :220-227
        for idx, infos in enumerate(self.index_info):
            for nvi in infos:
                if nvi == (node, deriv, dim):
                    self.bounds[idx] = (lower, upper)

    def add_bounds(self, node, deriv, dims, val, upper=None):         # :205-218, :229-236
        for k in dims:
            self.add_bound(node, deriv, k, val[k], val[k] if upper is None else upper[k])

    def add_start_bound(self, deriv, dims, val):                      # :238-242
        self.add_bounds(0, deriv, dims, val)

    def add_final_bound(self, deriv, dims, val):                      # :244-249
        self.add_bounds(self.nodes.shape[0] - 1, deriv, dims, val)


class NodesVariablesAll(NodesVariables):
    """variables/nodes_variables_all.cc:35-61."""

    def __init__(self, n_nodes, dim, name):
        super().__init__(name, n_nodes, dim)
        per_node = 2 * dim
        for idx in range(n_nodes * per_node):
            internal = idx % per_node
            self.index_info.append([(idx // per_node, kPos if internal < dim else kVel, internal % dim)])
        self.set_number_of_variables()


class NodesVariablesPhaseBased(NodesVariables):
    """variables/nodes_variables_phase_based.cc:39-213.  poly_info[i] = (phase, poly_in_phase, n_polys_in_phase, is_constant)."""

    def __init__(self, phase_count, first_phase_constant, name, n_polys_in_changing_phase, dim=3):
        self.poly_info = []
        constant = first_phase_constant
        for i in range(phase_count):                                  # BuildPolyInfos :39-59
            if constant:
                self.poly_info.append((i, 0, 1, True))
            else:
                for j in range(n_polys_in_changing_phase):
                    self.poly_info.append((i, j, n_polys_in_changing_phase, False))
            constant = not constant
        super().__init__(name, len(self.poly_info) + 1, dim)

    def convert_phase_to_poly_durations(self, phase_durations):       # :78-88
        return [phase_durations[info[0]] / info[2] for info in self.poly_info]

    def d_poly_duration_d_phase_duration(self, poly):                 # :90-96
        return 1.0 / self.poly_info[poly][2]

    def prev_polys_in_phase(self, poly):                              # :98-102
        return self.poly_info[poly][1]

    def adjacent_poly_ids(self, node):                                # :166-181
        last = self.nodes.shape[0] - 1
        if node == 0:
            return [0]
        if node == last:
            return [last - 1]
        return [node - 1, node]

    def is_constant_node(self, node):                                 # :104-113
        return any(self.poly_info[p][3] for p in self.adjacent_poly_ids(node))

    def non_constant_nodes(self):                                     # :121-130
        return [n for n in range(self.nodes.shape[0]) if not self.is_constant_node(n)]

    def phase_of(self, node):                                         # :132-138
        assert not self.is_constant_node(node)
        return self.poly_info[self.adjacent_poly_ids(node)[0]][0]

    def poly_at_start_of_phase(self, phase):                          # :140-148
        for i, info in enumerate(self.poly_info):
            if info[0] == phase:
                return i
        raise RuntimeError("phase not found")

    def node_at_start_of_phase(self, phase):                          # :157-162 (node id = poly id + Side::Start)
        return self.poly_at_start_of_phase(phase)

    def value_at_start_of_phase(self, phase):                         # :150-155
        return self.nodes[self.node_at_start_of_phase(phase), kPos].copy()


class NodesVariablesEEMotion(NodesVariablesPhaseBased):
    """:199-240 — swing nodes: x, y pos+vel and z pos (z vel fixed 0); a stance phase's two nodes share 3 positions."""

    def __init__(self, phase_count, in_contact_at_start, name, n_polys):
        super().__init__(phase_count, in_contact_at_start, name, n_polys)
        n = 0
        n_nodes = self.nodes.shape[0]
        while n < n_nodes:
            if not self.is_constant_node(n):
                for k in range(3):
                    self.index_info.append([(n, kPos, k)])
                    if k == Z:
                        self.nodes[n, kVel, Z] = 0.0
                    else:
                        self.index_info.append([(n, kVel, k)])
            else:
                self.nodes[n, kVel] = 0.0
                self.nodes[n + 1, kVel] = 0.0
                for k in range(3):
                    self.index_info.append([(n, kPos, k), (n + 1, kPos, k)])
                n += 1
            n += 1
        self.set_number_of_variables()


class NodesVariablesEEAng(NodesVariablesPhaseBased):
    """:352-396 — as EEMotion, but a swing node optimises all three velocities."""

    def __init__(self, phase_count, in_contact_at_start, name, n_polys):
        super().__init__(phase_count, in_contact_at_start, name, n_polys)
        n = 0
        n_nodes = self.nodes.shape[0]
        while n < n_nodes:
            if not self.is_constant_node(n):
                for k in range(3):
                    self.index_info.append([(n, kPos, k)])
                    self.index_info.append([(n, kVel, k)])
            else:
                self.nodes[n, kVel] = 0.0
                self.nodes[n + 1, kVel] = 0.0
                for k in range(3):
                    self.index_info.append([(n, kPos, k), (n + 1, kPos, k)])
                n += 1
            n += 1
        self.set_number_of_variables()


class NodesVariablesEELoad(NodesVariablesPhaseBased):
    """NodesVariablesEEForce :242-285 and NodesVariablesEETorque :287-350: the swing phase is the constant one (all zero,
    nothing optimised); a stance node optimises pos+vel in three dims."""

    def __init__(self, phase_count, in_contact_at_start, name, n_polys):
        super().__init__(phase_count, not in_contact_at_start, name, n_polys)
        n = 0
        n_nodes = self.nodes.shape[0]
        while n < n_nodes:
            if not self.is_constant_node(n):
                for k in range(3):
                    self.index_info.append([(n, kPos, k)])
                    self.index_info.append([(n, kVel, k)])
            else:
                self.nodes[n] = 0.0
                self.nodes[n + 1] = 0.0
                n += 1
            n += 1
        self.set_number_of_variables()


class PhaseDurations:
    """variables/phase_durations.cc:37-154."""

    def __init__(self, ee, timings, first_in_contact, min_duration, max_duration):
        self.name = "ee-schedule%d" % ee                              # variable_names.h id::EESchedule
        self.durations = [float(t) for t in timings]
        self.t_total = 0.0
        for t in self.durations:                                      # std::accumulate, left to right
            self.t_total += t
        self.bound = (min_duration, max_duration)
        self.initial_contact_state = bool(first_in_contact)
        self.observers = []

    def rows(self):
        return len(self.durations) - 1

    def values(self):
        return np.array(self.durations[:-1])

    def set_variables(self, x):                                       # :79-88
        for i in range(self.rows()):
            self.durations[i] = float(x[i])
        self.durations[-1] = self.t_total - float(np.sum(x))
        for o in self.observers:
            o.update_polynomial_durations()

    @property
    def bounds(self):
        return [self.bound] * self.rows()

    def is_contact_phase(self, t):                                    # :119-124
        phase = segment_id(t, self.durations)
        return self.initial_contact_state if phase % 2 == 0 else not self.initial_contact_state

    def jacobian_of_pos(self, current_phase, dx_dT, xd):              # :126-154; sparseView(1.0,-1.0) keeps every entry
        n_dim = len(xd)
        jac = np.zeros((n_dim, self.rows()))
        in_last_phase = current_phase == len(self.durations) - 1
        if not in_last_phase:
            jac[:, current_phase] = dx_dT
        for phase in range(current_phase):
            jac[:, phase] = -1 * xd
            if in_last_phase:
                jac[:, phase] -= dx_dT
        return Sp.from_dense(jac, keep_all=True)


# ------------------------------------------------------------------------------------------------ splines
def segment_id(t_global, durations):                                  # helpers/spline.cc:48-63
    eps = 1e-10
    assert t_global >= 0.0
    t = 0.0
    for i, d in enumerate(durations):
        t += d
        if t >= t_global - eps:
            return i
    raise AssertionError("t beyond the spline")


def local_time(t_global, durations):                                  # spline.cc:65-75
    i = segment_id(t_global, durations)
    t_local = t_global
    for k in range(i):
        t_local -= durations[k]
    return i, t_local


class NodeSpline:
    """helpers/spline.cc:35-118 + helpers/node_spline.cc:35-112."""

    def __init__(self, nodes, poly_durations):
        self.vars = nodes
        self.polys = [CubicHermite(nodes.dim) for _ in poly_durations]
        for p, d in zip(self.polys, poly_durations):
            p.T = float(d)
        for p in self.polys:
            p.update_coeff()
        self._memo = {}                                               # (poly, t_local, dxdt) -> Sp, dropped on every update
        self.version = 0
        nodes.observers.append(self)
        self.update_nodes()
        self.structure = {}                                           # jac_wrt_nodes_structure_: empty (n_dim x n_vars)

    def poly_durations(self):
        return [p.T for p in self.polys]

    def total_time(self):
        t = 0.0
        for p in self.polys:
            t += p.T
        return t

    def update_nodes(self):                                           # node_spline.cc:45-54
        for i, p in enumerate(self.polys):
            p.n0 = self.vars.nodes[i].copy()
            p.n1 = self.vars.nodes[i + 1].copy()
        for p in self.polys:
            p.update_coeff()
        self._memo.clear()
        self.version += 1

    def point(self, t_global):                                        # spline.cc:77-84
        i, t_local = local_time(t_global, self.poly_durations())
        return self.polys[i].point(t_local)

    def point_local(self, poly, t_local):                             # spline.cc:86-90
        return self.polys[poly].point(t_local)

    def fill_jacobian_wrt_nodes(self, poly, t_local, dxdt, entries, fill_with_zeros):    # node_spline.cc:82-112
        for idx, infos in enumerate(self.vars.index_info):
            for (n, d, k) in infos:
                for side in (0, 1):                                   # Side::Start, Side::End; node id = poly + side
                    if poly + side == n:
                        if side == 0:
                            val = self.polys[poly].d_wrt_start(dxdt, d, t_local)
                        else:
                            val = self.polys[poly].d_wrt_end(dxdt, d, t_local)
                        if fill_with_zeros:
                            val = 0.0
                        entries[(k, idx)] = entries.get((k, idx), 0.0) + val

    def jacobian_wrt_nodes_local(self, poly, t_local, dxdt):          # node_spline.cc:72-80
        key = (poly, t_local, dxdt)
        if key not in self._memo:                                     # the same (t, dxdt) is asked for by several constraints
            entries = dict(self.structure)
            self.fill_jacobian_wrt_nodes(poly, t_local, dxdt, entries, False)
            self._memo[key] = Sp.from_entries((self.vars.dim, self.vars.rows()), entries)
        return self._memo[key]

    def jacobian_wrt_nodes(self, t_global, dxdt):                     # node_spline.cc:63-70
        i, t_local = local_time(t_global, self.poly_durations())
        return self.jacobian_wrt_nodes_local(i, t_local, dxdt)


class PhaseSpline(NodeSpline):
    """helpers/phase_spline.cc:35-93."""

    def __init__(self, nodes, phase_durations):
        super().__init__(nodes, nodes.convert_phase_to_poly_durations(phase_durations.durations))
        self.phase_durations = phase_durations
        phase_durations.observers.append(self)
        self.update_polynomial_durations()
        for i in range(nodes.polynomial_count()):                     # :47-48: every reachable entry stored as 0
            self.fill_jacobian_wrt_nodes(i, 0.0, kPos, self.structure, True)

    def update_polynomial_durations(self):                            # :54-65
        durations = self.vars.convert_phase_to_poly_durations(self.phase_durations.durations)
        for p, d in zip(self.polys, durations):
            p.T = d
        for p in self.polys:
            p.update_coeff()
        self._memo.clear()
        self.version += 1

    def d_pos_d_phase_duration(self, t_global):                       # :79-93
        poly, t_local = local_time(t_global, self.poly_durations())
        vel = self.point(t_global)[kVel]
        dxdT = self.polys[poly].dpos_dT(t_local)
        inner = self.vars.d_poly_duration_d_phase_duration(poly)
        prev = self.vars.prev_polys_in_phase(poly)
        return inner * (dxdT - prev * vel)

    def jacobian_of_pos_wrt_durations(self, t_global):                # :67-77
        dx_dT = self.d_pos_d_phase_duration(t_global)
        xd = self.point(t_global)[kVel]
        phase = segment_id(t_global, self.phase_durations.durations)
        return self.phase_durations.jacobian_of_pos(phase, dx_dT, xd)


# ------------------------------------------------------------------------------------------------ Euler angles
def rotation_base_to_world(xyz):                                      # helpers/euler_converter.cc:207-221
    x, y, z = xyz
    s, c = math.sin, math.cos
    return np.array([
        [c(y) * c(z), c(z) * s(x) * s(y) - c(x) * s(z), s(x) * s(z) + c(x) * c(z) * s(y)],
        [c(y) * s(z), c(x) * c(z) + s(x) * s(y) * s(z), c(x) * s(y) * s(z) - c(z) * s(x)],
        [-s(y), c(y) * s(x), c(x) * c(y)]])


def euler_M(xyz):                                                     # euler_converter.cc:133-146 (coeffRef: 5 stored)
    z, y = xyz[Z], xyz[Y]
    return Sp.from_entries((3, 3), {(0, Y): -math.sin(z), (0, X): math.cos(y) * math.cos(z),
                                    (1, Y): math.cos(z), (1, X): math.cos(y) * math.sin(z),
                                    (2, Z): 1.0, (2, X): -math.sin(y)})


def euler_Mdot(xyz, xyz_d):                                           # euler_converter.cc:148-166 (5 stored)
    z, zd, y, yd = xyz[Z], xyz_d[Z], xyz[Y], xyz_d[Y]
    s, c = math.sin, math.cos
    return Sp.from_entries((3, 3), {(0, Y): -c(z) * zd, (0, X): -c(z) * s(y) * yd - c(y) * s(z) * zd,
                                    (1, Y): -s(z) * zd, (1, X): c(y) * c(z) * zd - s(y) * s(z) * yd,
                                    (2, X): -c(y) * yd})


class EulerConverter:
    """helpers/euler_converter.cc:37-304 over the base-ang NodeSpline."""

    def __init__(self, euler_spline):
        self.euler = euler_spline
        self.n = euler_spline.vars.rows()
        self._memo, self._memo_version = {}, -1

    def rotation(self, t):                                            # :200-205, dense Matrix3d for the callers that want one
        return rotation_base_to_world(self.euler.point(t)[kPos])

    def rotation_sp(self, t):                                         # MatrixSXd: M.sparseView(1.0, -1.0), all 9 stored
        return Sp.from_dense(self.rotation(t), keep_all=True)

    def angular_velocity(self, t):                                    # :59-71
        ori = self.euler.point(t)
        return euler_M(ori[kPos]).dense() @ ori[kVel]

    def angular_acceleration(self, t):                                # :73-84
        ori = self.euler.point(t)
        return euler_Mdot(ori[kPos], ori[kVel]).dense() @ ori[kVel] + euler_M(ori[kPos]).dense() @ ori[kAcc]

    def _jac(self, t, deriv, dim):                                    # :299-304
        return self.euler.jacobian_wrt_nodes(t, deriv).row(dim)

    def _empty_row(self):
        return Sp.zeros(1, self.n)

    def d_M_wrt_nodes(self, t, ang_dim):                              # :168-198
        ori = self.euler.point(t)
        z, y = ori[kPos][Z], ori[kPos][Y]
        s, c = math.sin, math.cos
        jac_z, jac_y = self._jac(t, kPos, Z), self._jac(t, kPos, Y)
        rows = [self._empty_row(), self._empty_row(), self._empty_row()]
        if ang_dim == X:
            rows[Y] = -c(z) * jac_z
            rows[X] = -c(z) * s(y) * jac_y - c(y) * s(z) * jac_z
        elif ang_dim == Y:
            rows[Y] = -s(z) * jac_z
            rows[X] = c(y) * c(z) * jac_z - s(y) * s(z) * jac_y
        else:
            rows[X] = -c(y) * jac_y
        return Sp.vstack(rows)

    def d_Mdot_wrt_nodes(self, t, ang_dim):                           # :270-297
        ori = self.euler.point(t)
        z, zd, y, yd = ori[kPos][Z], ori[kVel][Z], ori[kPos][Y], ori[kVel][Y]
        s, c = math.sin, math.cos
        jac_z, jac_y = self._jac(t, kPos, Z), self._jac(t, kPos, Y)
        jac_zd, jac_yd = self._jac(t, kVel, Z), self._jac(t, kVel, Y)
        rows = [self._empty_row(), self._empty_row(), self._empty_row()]
        if ang_dim == X:
            rows[Y] = s(z) * zd * jac_z - c(z) * jac_zd
            rows[X] = (s(y) * s(z) * yd * jac_z - c(y) * s(z) * jac_zd - c(y) * c(z) * yd * jac_y - c(y) * c(z) * zd * jac_z
                       - c(z) * s(y) * jac_yd + s(y) * s(z) * jac_y * zd)
        elif ang_dim == Y:
            rows[Y] = -s(z) * jac_zd - c(z) * zd * jac_z
            rows[X] = (c(y) * c(z) * jac_zd - s(y) * s(z) * jac_yd - c(y) * s(z) * yd * jac_y - c(z) * s(y) * yd * jac_z
                       - c(z) * s(y) * jac_y * zd - c(y) * s(z) * zd * jac_z)
        else:
            rows[X] = s(y) * yd * jac_y - c(y) * jac_yd
        return Sp.vstack(rows)

    def d_ang_vel_wrt_nodes(self, t):                                 # :86-102
        ori = self.euler.point(t)
        vel = Sp.from_dense(ori[kVel], keep_all=True)                 # 1 x 3 row, sparseView(1.0, -1.0)
        dVel_du = self.euler.jacobian_wrt_nodes(t, kVel)
        M = euler_M(ori[kPos])
        return Sp.vstack([vel @ self.d_M_wrt_nodes(t, dim) + M.row(dim) @ dVel_du for dim in (X, Y, Z)])

    def d_ang_acc_wrt_nodes(self, t):                                 # :104-131
        ori = self.euler.point(t)
        vel = Sp.from_dense(ori[kVel], keep_all=True)
        acc = Sp.from_dense(ori[kAcc], keep_all=True)
        dVel_du = self.euler.jacobian_wrt_nodes(t, kVel)
        dAcc_du = self.euler.jacobian_wrt_nodes(t, kAcc)
        M, Mdot = euler_M(ori[kPos]), euler_Mdot(ori[kPos], ori[kVel])
        rows = []
        for dim in (X, Y, Z):
            rows.append(vel @ self.d_Mdot_wrt_nodes(t, dim) + Mdot.row(dim) @ dVel_du
                        + acc @ self.d_M_wrt_nodes(t, dim) + M.row(dim) @ dAcc_du)
        return Sp.vstack(rows)

    def d_rotation_wrt_nodes(self, t):                                # :241-268: 3x3 table of 1 x n rows
        if self._memo_version != self.euler.version:
            self._memo, self._memo_version = {}, self.euler.version
        if t not in self._memo:
            self._memo[t] = self._d_rotation_wrt_nodes(t)
        return self._memo[t]

    def _d_rotation_wrt_nodes(self, t):
        ori = self.euler.point(t)
        x, y, z = ori[kPos]
        s, c = math.sin, math.cos
        jx, jy, jz = self._jac(t, kPos, X), self._jac(t, kPos, Y), self._jac(t, kPos, Z)
        R = [[None] * 3 for _ in range(3)]
        R[X][X] = -c(z) * s(y) * jy - c(y) * s(z) * jz
        R[X][Y] = (s(x) * s(z) * jx - c(x) * c(z) * jz - s(x) * s(y) * s(z) * jz + c(x) * c(z) * s(y) * jx
                   + c(y) * c(z) * s(x) * jy)
        R[X][Z] = (c(x) * s(z) * jx + c(z) * s(x) * jz - c(z) * s(x) * s(y) * jx - c(x) * s(y) * s(z) * jz
                   + c(x) * c(y) * c(z) * jy)
        R[Y][X] = c(y) * c(z) * jz - s(y) * s(z) * jy
        R[Y][Y] = (c(x) * s(y) * s(z) * jx - c(x) * s(z) * jz - c(z) * s(x) * jx + c(y) * s(x) * s(z) * jy
                   + c(z) * s(x) * s(y) * jz)
        R[Y][Z] = (s(x) * s(z) * jz - c(x) * c(z) * jx - s(x) * s(y) * s(z) * jx + c(x) * c(y) * s(z) * jy
                   + c(x) * c(z) * s(y) * jz)
        R[Z][X] = -c(y) * jy
        R[Z][Y] = c(x) * c(y) * jx - s(x) * s(y) * jy
        R[Z][Z] = -c(y) * s(x) * jx - c(x) * s(y) * jy
        return R

    def d_rot_vec_mult(self, t, v, inverse):                          # :223-239
        Rd = self.d_rotation_wrt_nodes(t)
        rows = []
        for row in (X, Y, Z):
            acc = self._empty_row()
            for col in (X, Y, Z):
                jac_row = Rd[col][row] if inverse else Rd[row][col]
                acc = acc + v[col] * jac_row
            rows.append(acc)
        return Sp.vstack(rows)


# ------------------------------------------------------------------------------------------------ terrain
class HeightMap:
    """terrain/height_map.cc:62-148; friction_coeff_ = 0.5 (height_map.h:136)."""
    mu = 0.5

    def height(self, x, y):
        return 0.0

    def dx(self, x, y):
        return 0.0

    def dy(self, x, y):
        return 0.0

    def dxx(self, x, y):
        return 0.0

    def dxy(self, x, y):
        return 0.0

    def dyx(self, x, y):
        return 0.0

    def dyy(self, x, y):
        return 0.0

    def d_height(self, dim, x, y):                                    # :50-59
        return self.dx(x, y) if dim == X else self.dy(x, y)

    def dd_height(self, dim1, dim2, x, y):                            # :134-148
        if dim1 == X:
            return self.dxx(x, y) if dim2 == X else self.dxy(x, y)
        return self.dyx(x, y) if dim2 == X else self.dyy(x, y)

    def basis(self, which, x, y, deriv=None):                         # :68-133; which: 0 normal, 1 tangent1, 2 tangent2
        requested = deriv is None
        if which == 0:
            n = np.zeros(3)
            for dim in (X, Y):
                n[dim] = -self.d_height(dim, x, y) if requested else -self.dd_height(dim, deriv, x, y)
            n[Z] = 1.0 if requested else 0.0
            return n
        if which == 1:
            return np.array([1.0 if requested else 0.0, 0.0,
                             self.d_height(X, x, y) if requested else self.dd_height(X, deriv, x, y)])
        return np.array([0.0, 1.0 if requested else 0.0,
                         self.d_height(Y, x, y) if requested else self.dd_height(Y, deriv, x, y)])

    def normalized_basis(self, which, x, y):                          # :62-66
        v = self.basis(which, x, y)
        return v / math.sqrt(v @ v)

    def d_normalized_basis(self, which, dim, x, y):                   # :82-91, :127-132
        dv = self.basis(which, x, y, dim)
        v = self.basis(which, x, y)
        sq = v @ v
        norm = math.sqrt(sq)
        unit = np.zeros(3)
        unit[dim] = 1.0
        d_norm = 1 / sq * (norm * unit - v[dim] * (v / norm))
        return d_norm * dv


class FlatGround(HeightMap):
    pass


class Block(HeightMap):                                               # height_map_examples.cc:43-63, .h:58-71
    start, length, h, eps = 0.7, 3.5, 0.5, 0.03
    slope = h / eps

    def height(self, x, y):
        out = 0.0
        if self.start <= x <= self.start + self.eps:
            out = self.slope * (x - self.start)
        if self.start + self.eps <= x <= self.start + self.length:
            out = self.h
        return out

    def dx(self, x, y):
        return self.slope if self.start <= x <= self.start + self.eps else 0.0


class Stairs(HeightMap):                                              # .cc:66-80, .h:76-87
    def height(self, x, y):
        h = 0.0
        if x >= 1.0:
            h = 0.2
        if x >= 1.0 + 0.4:
            h = 0.4
        if x >= 1.0 + 0.4 + 1.0:
            h = 0.0
        return h


class Gap(HeightMap):                                                 # .cc:83-112, .h:92-117
    gap_start, w, h = 1.0, 0.5, 1.5
    dxw = w / 2.0
    xc = gap_start + dxw
    gap_end = gap_start + w
    a = (4 * h) / (w * w)
    b = -(8 * h * xc) / (w * w)
    c = -(h * (w - 2 * xc) * (w + 2 * xc)) / (w * w)

    def _inside(self, x):
        return self.gap_start <= x <= self.gap_end

    def height(self, x, y):
        return self.a * x * x + self.b * x + self.c if self._inside(x) else 0.0

    def dx(self, x, y):
        return 2 * self.a * x + self.b if self._inside(x) else 0.0

    def dxx(self, x, y):
        return 2 * self.a if self._inside(x) else 0.0


class Slope(HeightMap):                                               # .cc:115-145, .h:122-137
    start, up, down, hc = 1.0, 1.0, 1.0, 0.7
    x_down = start + up
    x_flat = x_down + down
    slope = hc / up

    def height(self, x, y):
        z = 0.0
        if x >= self.start:
            z = self.slope * (x - self.start)
        if x >= self.x_down:
            z = self.hc - self.slope * (x - self.x_down)
        if x >= self.x_flat:
            z = 0.0
        return z

    def dx(self, x, y):
        d = 0.0
        if x >= self.start:
            d = self.slope
        if x >= self.x_down:
            d = -self.slope
        if x >= self.x_flat:
            d = 0.0
        return d


class Chimney(HeightMap):                                             # .cc:148-168, .h:142-155
    x_start, length, y_start, slope = 1.0, 1.5, 0.5, 3.0
    x_end = x_start + length

    def height(self, x, y):
        return self.slope * (y - self.y_start) if self.x_start <= x <= self.x_end else 0.0

    def dy(self, x, y):
        return self.slope if self.x_start <= x <= self.x_end else 0.0


class ChimneyLR(HeightMap):                                           # .cc:171-198, .h:160-175
    x_start, length, y_start, slope = 0.5, 1.0, 0.5, 2.0
    x_end1 = x_start + length
    x_end2 = x_start + 2 * length

    def height(self, x, y):
        z = 0.0
        if self.x_start <= x <= self.x_end1:
            z = self.slope * (y - self.y_start)
        if self.x_end1 <= x <= self.x_end2:
            z = -self.slope * (y + self.y_start)
        return z

    def dy(self, x, y):
        d = 0.0
        if self.x_start <= x <= self.x_end1:
            d = self.slope
        if self.x_end1 <= x <= self.x_end2:
            d = -self.slope
        return d


class FiveStepStairs(HeightMap):                                      # test/hopper_example.cc:53-91
    def height(self, x, y):
        if x < 0.5:
            return 0.0
        step = int((x - 0.5) / 0.3)
        if step >= 5:
            return 5 * 0.15
        return (step + 1) * 0.15


class TableTerrain(HeightMap):
    """A user height map given as data (orc_spec::terrain_table_*, terrain id 8): profile along x.  mode 0: steps (0 before
    the first knot, h[i] on [x[i], x[i+1]), h[-1] from the last knot on); mode 1: ramps (h[0] before the first knot, linear
    between knots with the segment's slope as d h / d x, h[-1] and no slope from the last knot on)."""

    def __init__(self, mode, xs, hs):
        self.steps, self.xs, self.hs = mode == 0, np.asarray(xs, float), np.asarray(hs, float)

    def _i(self, x):
        return int(np.searchsorted(self.xs, x, side="right")) - 1

    def height(self, x, y):
        i = self._i(x)
        if i < 0:
            return 0.0 if self.steps else float(self.hs[0])
        if self.steps or i + 1 == len(self.xs):
            return float(self.hs[i])
        return float(self.hs[i] + self.dx(x, y) * (x - self.xs[i]))

    def dx(self, x, y):
        i = self._i(x)
        if self.steps or i < 0 or i + 1 == len(self.xs):
            return 0.0
        return float((self.hs[i + 1] - self.hs[i]) / (self.xs[i + 1] - self.xs[i]))


TERRAINS = [FlatGround, Block, Stairs, Gap, Slope, Chimney, ChimneyLR, FiveStepStairs]


# ------------------------------------------------------------------------------------------------ dynamic model
class SingleRigidBodyDynamics:
    """models/single_rigid_body_dynamics.cc:35-204, models/dynamic_model.cc:35-66 (g = 9.80665)."""

    def __init__(self, mass, inertia6, n_ee):
        Ixx, Iyy, Izz, Ixy, Ixz, Iyz = inertia6
        I = np.array([[Ixx, -Ixy, -Ixz], [-Ixy, Iyy, -Iyz], [-Ixz, -Iyz, Izz]])
        self.I_b = Sp.from_dense(I, keep_all=False)                   # inertia_b.sparseView(): exact zeros dropped
        self.m, self.g, self.n_ee = mass, 9.80665, n_ee

    def set_current(self, com, com_acc, w_R_b, omega, omega_dot, force, pos, torque):
        self.com, self.com_acc, self.R, self.omega, self.omega_dot = com, com_acc, w_R_b, omega, omega_dot
        self.force, self.pos, self.torque = force, pos, torque

    def _R(self):
        return Sp.from_dense(self.R, keep_all=False)                  # w_R_b_.sparseView()

    def _I_w(self):
        return self._R() @ self.I_b @ Sp.from_dense(self.R.T, keep_all=False)

    def dynamic_violation(self):                                      # :76-101
        f_sum, tau_sum = np.zeros(3), np.zeros(3)
        for ee in range(self.n_ee):
            f = self.force[ee]
            tau_sum = tau_sum + (np.cross(f, self.com - self.pos[ee]) + self.torque[ee])
            f_sum = f_sum + f
        I_w = self._I_w().dense()
        acc = np.zeros(6)
        acc[AX:AX + 3] = I_w @ self.omega_dot + cross(self.omega).dense() @ (I_w @ self.omega) - tau_sum
        acc[LX:LX + 3] = self.m * self.com_acc - f_sum - np.array([0.0, 0.0, -self.m * self.g])
        return acc

    @staticmethod
    def _stack6(ang, lin, n):
        return Sp.vstack([ang if ang is not None else Sp.zeros(3, n), lin if lin is not None else Sp.zeros(3, n)])

    def jac_wrt_base_lin(self, jac_pos, jac_acc):                     # :103-121
        n = jac_pos.shape[1]
        jac_tau_sum = Sp.zeros(3, n)
        for f in self.force:
            jac_tau_sum = jac_tau_sum + cross(f) @ jac_pos
        return self._stack6(-jac_tau_sum, self.m * jac_acc, n)

    def jac_wrt_base_ang(self, conv, t):                              # :123-155
        R, I_b, I_w = self._R(), self.I_b, self._I_w()
        Rt_dense = self.R.T
        I_b_dense = I_b.dense()
        v11 = I_b_dense @ Rt_dense @ self.omega_dot
        jac11 = conv.d_rot_vec_mult(t, v11, False)
        jac12 = R @ I_b @ conv.d_rot_vec_mult(t, self.omega_dot, True)
        jac_ang_acc = conv.d_ang_acc_wrt_nodes(t)
        jac13 = I_w @ jac_ang_acc
        jac1 = jac11 + jac12 + jac13
        v21 = I_b_dense @ Rt_dense @ self.omega
        jac21 = conv.d_rot_vec_mult(t, v21, False)
        jac22 = R @ I_b @ conv.d_rot_vec_mult(t, self.omega, True)
        jac_ang_vel = conv.d_ang_vel_wrt_nodes(t)
        jac23 = I_w @ jac_ang_vel
        jac2 = cross(self.omega) @ (jac21 + jac22 + jac23) - cross(I_w.dense() @ self.omega) @ jac_ang_vel
        n = jac_ang_vel.shape[1]
        return self._stack6(jac1 + jac2, None, n)

    def jac_wrt_force(self, jac_force, ee):                           # :157-170
        r = self.com - self.pos[ee]
        jac_tau = -(cross(r) @ jac_force)
        return self._stack6(-jac_tau, -jac_force, jac_force.shape[1])

    def jac_wrt_torque(self, jac_torque, ee):                         # :172-181
        return self._stack6(-jac_torque, None, jac_torque.shape[1])

    def jac_wrt_ee_pos(self, jac_ee_pos, ee):                         # :183-194
        jac_tau = cross(self.force[ee]) @ (-jac_ee_pos)
        return self._stack6(-jac_tau, None, jac_tau.shape[1])


# ------------------------------------------------------------------------------------------------ constraints
class ConstraintSet:
    """ifopt::ConstraintSet (ifopt 2.0.1, not vendored; towr/CMakeLists.txt:5): values, bounds, and one Jacobian block per
    variable set, filled into an EMPTY rows x n block (composite.cc: `Jacobian jac(GetRows(), n); FillJacobianBlock`)."""
    name = "?"
    n_rows = 0

    def link(self, variables):
        pass

    def values(self):
        raise NotImplementedError

    def bounds(self):
        raise NotImplementedError

    def block(self, var_set):
        """Sp (n_rows x var_set.rows()) or None when the set does not depend on var_set."""
        raise NotImplementedError


def sample_times(T, dt):                                              # time_discretization_constraint.cc:36-49
    t = 0.0
    out = [t]
    for _ in range(int(math.floor(T / dt))):
        t += dt
        out.append(t)
    out.append(T)
    return out


class TimeDiscretizationConstraint(ConstraintSet):
    """constraints/time_discretization_constraint.cc:36-98.  rows_per_instance rows per sample; `instance_block` returns
    the rows of one sample, which `jac.middleRows(row, n) = ...` puts in place (absent rows stay empty)."""
    rows_per_instance = 0

    def __init__(self, T, dt, name):
        self.name = name
        self.dts = sample_times(T, dt)

    def finish(self):
        self.n_rows = len(self.dts) * self.rows_per_instance

    def values(self):
        g = np.zeros(self.n_rows)
        for k, t in enumerate(self.dts):
            g[k * self.rows_per_instance:(k + 1) * self.rows_per_instance] = self.instance_values(t, k)
        return g

    def bounds(self):
        out = []
        for k, t in enumerate(self.dts):
            out.extend(self.instance_bounds(t, k))
        return out

    def block(self, var_set):
        if not self.depends_on(var_set.name):
            return None
        n = var_set.rows()
        rows = []
        for k, t in enumerate(self.dts):
            b = self.instance_block(t, k, var_set.name, n)
            rows.append(b if b is not None else Sp.zeros(self.rows_per_instance, n))
        return Sp.vstack(rows)


class DynamicConstraint(TimeDiscretizationConstraint):
    """constraints/dynamic_constraint.cc:38-148."""
    rows_per_instance = 6

    def __init__(self, model, T, dt, s):
        super().__init__(T, dt, "dynamic")
        self.model, self.s = model, s
        self.finish()

    def update_model(self, t):                                        # :124-146
        s = self.s
        com = s.base_linear.point(t)
        conv = s.angular_converter
        n_ee = self.model.n_ee
        self.model.set_current(com[kPos], com[kAcc], conv.rotation(t), conv.angular_velocity(t), conv.angular_acceleration(t),
                               [s.ee_force[e].point(t)[kPos] for e in range(n_ee)],
                               [s.ee_motion[e].point(t)[kPos] for e in range(n_ee)],
                               [s.ee_torque[e].point(t)[kPos] for e in range(n_ee)])

    def instance_values(self, t, k):                                  # :65-70
        self.update_model(t)
        return self.model.dynamic_violation()

    def instance_bounds(self, t, k):
        return [BOUND_ZERO] * 6

    def depends_on(self, name):
        return not name.startswith("ee-ang_")                         # :79-122: for ee-ang sets jac_model stays empty

    def instance_block(self, t, k, name, n):                          # :79-122
        self.update_model(t)
        s, m = self.s, self.model
        jac_model = Sp.zeros(6, n)
        if name == "base-lin":
            jac_model = m.jac_wrt_base_lin(s.base_linear.jacobian_wrt_nodes(t, kPos), s.base_linear.jacobian_wrt_nodes(t, kAcc))
        if name == "base-ang":
            jac_model = m.jac_wrt_base_ang(s.angular_converter, t)
        for ee in range(m.n_ee):
            if name == "ee-force_%d" % ee:
                jac_model = m.jac_wrt_force(s.ee_force[ee].jacobian_wrt_nodes(t, kPos), ee)
            if name == "ee-torque_%d" % ee:
                jac_model = m.jac_wrt_torque(s.ee_torque[ee].jacobian_wrt_nodes(t, kPos), ee)
            if name == "ee-motion_%d" % ee:
                jac_model = m.jac_wrt_ee_pos(s.ee_motion[ee].jacobian_wrt_nodes(t, kPos), ee)
            if name == "ee-schedule%d" % ee:
                jac_model = jac_model + m.jac_wrt_force(s.ee_force[ee].jacobian_of_pos_wrt_durations(t), ee)
                jac_model = jac_model + m.jac_wrt_ee_pos(s.ee_motion[ee].jacobian_of_pos_wrt_durations(t), ee)
        return jac_model


class RangeOfMotionConstraint(TimeDiscretizationConstraint):
    """constraints/range_of_motion_constraint.cc:37-138 (no swing relaxation: rom_swing_relax_dims_ is empty by default)."""
    rows_per_instance = 3

    def __init__(self, kin, T, dt, ee, s):
        super().__init__(T, dt, "rangeofmotion-%d" % ee)
        self.s, self.ee = s, ee
        self.nominal, self.max_dev, self.min_dev = kin["nominal"][ee], kin["max_dev"][ee], kin["min_dev"][ee]
        self.finish()

    def instance_values(self, t, k):                                  # :71-82
        s = self.s
        base_W = s.base_linear.point(t)[kPos]
        ee_W = s.ee_motion[self.ee].point(t)[kPos]
        b_R_w = s.angular_converter.rotation(t).T
        return b_R_w @ (ee_W - base_W)

    def instance_bounds(self, t, k):                                  # :84-103
        return [(self.nominal[d] + self.min_dev[d], self.nominal[d] + self.max_dev[d]) for d in range(3)]

    def depends_on(self, name):
        return name in ("base-lin", "base-ang", "ee-motion_%d" % self.ee, "ee-schedule%d" % self.ee)

    def instance_block(self, t, k, name, n):                          # :105-136
        s = self.s
        b_R_w = s.angular_converter.rotation_sp(t).T()
        if name == "base-lin":
            return (-1 * b_R_w) @ s.base_linear.jacobian_wrt_nodes(t, kPos)
        if name == "base-ang":
            r_W = s.ee_motion[self.ee].point(t)[kPos] - s.base_linear.point(t)[kPos]
            return s.angular_converter.d_rot_vec_mult(t, r_W, True)
        if name == "ee-motion_%d" % self.ee:
            return b_R_w @ s.ee_motion[self.ee].jacobian_wrt_nodes(t, kPos)
        return b_R_w @ s.ee_motion[self.ee].jacobian_of_pos_wrt_durations(t)


class BaseMotionConstraint(TimeDiscretizationConstraint):
    """constraints/base_motion_constraint.cc:37-93.  Rows per sample in Dim6D order AX, AY, AZ, LX, LY, LZ."""
    rows_per_instance = 6

    def __init__(self, T, dt, ax, ay, lz, s):
        super().__init__(T, dt, "baseMotion")
        self.s = s
        self.node_bounds = [ax, ay, NO_BOUND, NO_BOUND, NO_BOUND, lz]
        self.finish()

    def instance_values(self, t, k):
        return np.concatenate([self.s.base_angular.point(t)[kPos], self.s.base_linear.point(t)[kPos]])

    def instance_bounds(self, t, k):
        return list(self.node_bounds)

    def depends_on(self, name):
        return name in ("base-lin", "base-ang")

    def instance_block(self, t, k, name, n):
        if name == "base-ang":
            return Sp.vstack([self.s.base_angular.jacobian_wrt_nodes(t, kPos), Sp.zeros(3, n)])
        return Sp.vstack([Sp.zeros(3, n), self.s.base_linear.jacobian_wrt_nodes(t, kPos)])


class EELinearConstraint(TimeDiscretizationConstraint):
    """constraints/ee_linear_constraint.cc:5-48: |sum_j coeff_j * spline_{ee_j}(t).at(deriv)(dim_j)| <= tolerance."""
    rows_per_instance = 1

    def __init__(self, terms, splines, var_names, deriv, tolerance, T, dt):
        super().__init__(T, dt, "ee-linear")
        self.terms, self.splines, self.var_names, self.deriv, self.tolerance = terms, splines, var_names, deriv, tolerance
        self.finish()

    def instance_values(self, t, k):                                  # :20-29
        val = 0.0
        for (ee, dim, coeff) in self.terms:
            val += coeff * self.splines[ee].point(t)[self.deriv][dim]
        return np.array([val])

    def instance_bounds(self, t, k):                                  # :31-34
        return [(-self.tolerance, self.tolerance)]

    def depends_on(self, name):
        return any(name == self.var_names[ee] for (ee, _, _) in self.terms)

    def instance_block(self, t, k, name, n):                          # :36-48: jac.row(k) += coeff * jac_nodes.row(dim)
        row = Sp.zeros(1, n)
        for (ee, dim, coeff) in self.terms:
            if name == self.var_names[ee]:
                row = row + coeff * self.splines[ee].jacobian_wrt_nodes(t, self.deriv).row(dim)
        return row


def _accumulate_linear_form(J, b, dst_row, entries):                  # force_constraint_discretized.cc:36-49
    r, c, v = J.stored()
    for i in range(len(r)):
        entries[(dst_row, c[i])] = entries.get((dst_row, c[i]), 0.0) + b[r[i]] * v[i]


def _accumulate_scaled_row(J, src_row, scale, dst_row, entries):      # force_constraint_discretized.cc:51-67
    if scale == 0.0:
        return
    r, c, v = J.stored()
    for i in range(len(r)):
        if r[i] == src_row:
            entries[(dst_row, c[i])] = entries.get((dst_row, c[i]), 0.0) + scale * v[i]


class ForceConstraintDiscretized(TimeDiscretizationConstraint):
    """constraints/force_constraint_discretized.cc:71-221.  The block is filled by coeffRef accumulation into `jac` itself,
    so an entry is stored only where some accumulation touched it (scale == 0.0 skips, :58)."""
    rows_per_instance = 5

    def __init__(self, terrain, T, dt, fn_max, ee, s):
        super().__init__(T, dt, "force-disc-ee-force_%d" % ee)
        self.terrain, self.fn_max, self.mu, self.ee, self.s = terrain, fn_max, terrain.mu, ee, s
        self.finish()

    def _basis(self, t):
        p = self.s.ee_motion[self.ee].point(t)[kPos]
        f = self.s.ee_force[self.ee].point(t)[kPos]
        n = self.terrain.normalized_basis(0, p[X], p[Y])
        t1 = self.terrain.normalized_basis(1, p[X], p[Y])
        t2 = self.terrain.normalized_basis(2, p[X], p[Y])
        return p, f, [n, t1 - self.mu * n, t1 + self.mu * n, t2 - self.mu * n, t2 + self.mu * n]

    def instance_values(self, t, k):                                  # :98-117
        p, f, b = self._basis(t)
        return np.array([f @ bi for bi in b])

    def instance_bounds(self, t, k):                                  # :119-128
        return [(0.0, self.fn_max), BOUND_SMALLER_ZERO, BOUND_GREATER_ZERO, BOUND_SMALLER_ZERO, BOUND_GREATER_ZERO]

    def depends_on(self, name):
        return name in ("ee-force_%d" % self.ee, "ee-motion_%d" % self.ee, "ee-schedule%d" % self.ee)

    def _scales(self, p, f, dim):
        dn = self.terrain.d_normalized_basis(0, dim, p[X], p[Y])
        dt1 = self.terrain.d_normalized_basis(1, dim, p[X], p[Y])
        dt2 = self.terrain.d_normalized_basis(2, dim, p[X], p[Y])
        mu = self.mu
        return [f @ dn, f @ (dt1 - mu * dn), f @ (dt1 + mu * dn), f @ (dt2 - mu * dn), f @ (dt2 + mu * dn)]

    def instance_block(self, t, k, name, n):                          # :130-219
        p, f, b = self._basis(t)
        s, ee = self.s, self.ee
        entries = {}
        if name == "ee-force_%d" % ee:
            Jf = s.ee_force[ee].jacobian_wrt_nodes(t, kPos)
            for r in range(5):
                _accumulate_linear_form(Jf, b[r], r, entries)
        if name == "ee-motion_%d" % ee:
            Jp = s.ee_motion[ee].jacobian_wrt_nodes(t, kPos)
            for dim in (X, Y):
                sc = self._scales(p, f, dim)
                for r in range(5):
                    _accumulate_scaled_row(Jp, dim, sc[r], r, entries)
        if name == "ee-schedule%d" % ee:
            Jf_dT = s.ee_force[ee].jacobian_of_pos_wrt_durations(t)
            Jp_dT = s.ee_motion[ee].jacobian_of_pos_wrt_durations(t)
            for r in range(5):
                _accumulate_linear_form(Jf_dT, b[r], r, entries)
            for dim in (X, Y):
                sc = self._scales(p, f, dim)
                for r in range(5):
                    _accumulate_scaled_row(Jp_dT, dim, sc[r], r, entries)
        return Sp.from_entries((5, n), entries)


class TorqueConstraintDiscretized(TimeDiscretizationConstraint):
    """constraints/torque_constraint_discretized.cc:66-235."""
    rows_per_instance = 4

    def __init__(self, terrain, T, dt, tx_min, tx_max, ty_min, ty_max, k_friction, ee, s):
        super().__init__(T, dt, "torque-disc-ee-torque_%d" % ee)
        self.terrain, self.mu, self.ee, self.s = terrain, terrain.mu, ee, s
        self.lim = (tx_min, tx_max, ty_min, ty_max)
        self.kf = k_friction
        self.finish()

    def _state(self, t):
        s, ee = self.s, self.ee
        p = s.ee_motion[ee].point(t)[kPos]
        f = s.ee_force[ee].point(t)[kPos]
        tau = s.ee_torque[ee].point(t)[kPos]
        n = self.terrain.normalized_basis(0, p[X], p[Y])
        t1 = self.terrain.normalized_basis(1, p[X], p[Y])
        t2 = self.terrain.normalized_basis(2, p[X], p[Y])
        return p, f, tau, n, t1, t2

    def instance_values(self, t, k):                                  # :100-123
        p, f, tau, n, t1, t2 = self._state(t)
        tau_n = tau @ n
        tz_lim = self.kf * self.mu * (f @ n)
        return np.array([tau @ t1, tau @ t2, tau_n - tz_lim, -tau_n - tz_lim])

    def instance_bounds(self, t, k):                                  # :125-133
        return [(self.lim[0], self.lim[1]), (self.lim[2], self.lim[3]), BOUND_SMALLER_ZERO, BOUND_SMALLER_ZERO]

    def depends_on(self, name):
        e = self.ee
        return name in ("ee-torque_%d" % e, "ee-force_%d" % e, "ee-motion_%d" % e, "ee-schedule%d" % e)

    def _motion_scales(self, p, f, tau, dim):
        dn = self.terrain.d_normalized_basis(0, dim, p[X], p[Y])
        dt1 = self.terrain.d_normalized_basis(1, dim, p[X], p[Y])
        dt2 = self.terrain.d_normalized_basis(2, dim, p[X], p[Y])
        s_tau_n = tau @ dn
        s_lim = self.kf * self.mu * (f @ dn)
        return [tau @ dt1, tau @ dt2, s_tau_n - s_lim, -s_tau_n - s_lim]

    def instance_block(self, t, k, name, n_cols):                     # :135-233
        p, f, tau, n, t1, t2 = self._state(t)
        s, ee = self.s, self.ee
        entries = {}
        b = -self.kf * self.mu * n
        if name == "ee-torque_%d" % ee:
            J = s.ee_torque[ee].jacobian_wrt_nodes(t, kPos)
            for r, vec in enumerate((t1, t2, n, -n)):
                _accumulate_linear_form(J, vec, r, entries)
        if name == "ee-force_%d" % ee:
            J = s.ee_force[ee].jacobian_wrt_nodes(t, kPos)
            _accumulate_linear_form(J, b, 2, entries)
            _accumulate_linear_form(J, b, 3, entries)
        if name == "ee-motion_%d" % ee:
            Jp = s.ee_motion[ee].jacobian_wrt_nodes(t, kPos)
            for dim in (X, Y):
                sc = self._motion_scales(p, f, tau, dim)
                for r in range(4):
                    _accumulate_scaled_row(Jp, dim, sc[r], r, entries)
        if name == "ee-schedule%d" % ee:
            Jtau_dT = s.ee_torque[ee].jacobian_of_pos_wrt_durations(t)
            Jf_dT = s.ee_force[ee].jacobian_of_pos_wrt_durations(t)
            Jp_dT = s.ee_motion[ee].jacobian_of_pos_wrt_durations(t)
            for r, vec in enumerate((t1, t2, n, -n)):
                _accumulate_linear_form(Jtau_dT, vec, r, entries)
            _accumulate_linear_form(Jf_dT, b, 2, entries)
            _accumulate_linear_form(Jf_dT, b, 3, entries)
            for dim in (X, Y):
                sc = self._motion_scales(p, f, tau, dim)
                for r in range(4):
                    _accumulate_scaled_row(Jp_dT, dim, sc[r], r, entries)
        return Sp.from_entries((4, n_cols), entries)


class TerrainConstraintHard(TimeDiscretizationConstraint):
    """constraints/terrain_constraint_hard.cc:34-132."""
    rows_per_instance = 1
    k_coeff = 0.02

    def __init__(self, terrain, T, dt, ee, s):
        super().__init__(T, dt, "terrainhard-%d" % ee)
        self.terrain, self.ee, self.s = terrain, ee, s
        self.finish()

    def _state(self, t):
        pt = self.s.ee_motion[self.ee].point(t)
        p, v = pt[kPos], pt[kVel]
        t1 = self.terrain.normalized_basis(1, p[X], p[Y])
        t2 = self.terrain.normalized_basis(2, p[X], p[Y])
        vt1, vt2 = v @ t1, v @ t2
        return p, v, t1, t2, vt1, vt2, math.sqrt(vt1 * vt1 + vt2 * vt2)

    def instance_values(self, t, k):                                  # :50-69
        p, v, t1, t2, vt1, vt2, mag = self._state(t)
        delta_z = p[Z] - self.terrain.height(p[X], p[Y])
        return np.array([delta_z - min(self.k_coeff * mag, self.k_coeff)])

    def instance_bounds(self, t, k):
        return [(0.0, 1e20)]

    def depends_on(self, name):
        return name == "ee-motion_%d" % self.ee

    def instance_block(self, t, k, name, n):                          # :79-130
        p, v, t1, t2, vt1, vt2, mag = self._state(t)
        spline = self.s.ee_motion[self.ee]
        jac_pos = spline.jacobian_wrt_nodes(t, kPos)
        jac_vel = spline.jacobian_wrt_nodes(t, kVel)
        row = jac_pos.row(Z)
        for dim in (X, Y):
            row = row - self.terrain.d_height(dim, p[X], p[Y]) * jac_pos.row(dim)
        if mag > 1e-6:
            if self.k_coeff * mag < 0.05 - 1e-6:
                tangent_deriv = (vt1 * t1 + vt2 * t2) / mag
                for dim in (X, Y, Z):
                    row = row - self.k_coeff * tangent_deriv[dim] * jac_vel.row(dim)
        return row


class TerrainConstraint(ConstraintSet):
    """constraints/terrain_constraint.cc:33-106."""

    def __init__(self, terrain, ee_motion_name, min_height, max_height):
        self.name = "terrain-" + ee_motion_name
        self.var_name, self.terrain, self.min_h, self.max_h = ee_motion_name, terrain, min_height, max_height

    def link(self, variables):
        self.ee_motion = variables[self.var_name]
        self.node_ids = list(range(1, self.ee_motion.nodes.shape[0]))
        self.n_rows = len(self.node_ids)

    def values(self):
        nodes = self.ee_motion.nodes
        return np.array([nodes[n, kPos, Z] - self.terrain.height(nodes[n, kPos, X], nodes[n, kPos, Y]) for n in self.node_ids])

    def bounds(self):
        return [BOUND_ZERO if self.ee_motion.is_constant_node(n) else (self.min_h, self.max_h) for n in self.node_ids]

    def block(self, var_set):
        if var_set.name != self.var_name:
            return None
        entries = {}
        nodes = self.ee_motion.nodes
        for row, n in enumerate(self.node_ids):
            entries[(row, self.ee_motion.opt_index(n, kPos, Z))] = 1.0
            p = nodes[n, kPos]
            for dim in (X, Y):
                entries[(row, self.ee_motion.opt_index(n, kPos, dim))] = -self.terrain.d_height(dim, p[X], p[Y])
        return Sp.from_entries((self.n_rows, var_set.rows()), entries)


class BaseHeightConstraint(ConstraintSet):
    """constraints/base_height_constraint.cc:6-78."""

    def __init__(self, terrain, safety_distance, base_lin_name):
        self.name = "base-height-" + base_lin_name
        self.var_name, self.terrain, self.safety = base_lin_name, terrain, safety_distance

    def link(self, variables):
        self.base = variables[self.var_name]
        self.node_ids = list(range(1, self.base.nodes.shape[0]))
        self.n_rows = len(self.node_ids)

    def values(self):
        nodes = self.base.nodes
        return np.array([nodes[n, kPos, Z] - self.terrain.height(nodes[n, kPos, X], nodes[n, kPos, Y]) - self.safety
                         for n in self.node_ids])

    def bounds(self):
        return [(0.0, 1e20)] * self.n_rows

    def block(self, var_set):
        if var_set.name != self.var_name:
            return None
        entries = {}
        nodes = self.base.nodes
        for row, n in enumerate(self.node_ids):
            entries[(row, self.base.opt_index(n, kPos, Z))] = 1.0
            p = nodes[n, kPos]
            for dim in (X, Y):
                entries[(row, self.base.opt_index(n, kPos, dim))] = -self.terrain.d_height(dim, p[X], p[Y])
        return Sp.from_entries((self.n_rows, var_set.rows()), entries)


class ForceConstraint(ConstraintSet):
    """constraints/force_constraint.cc:35-154 (node-based; used when dt_constraint_force_ <= 0)."""

    def __init__(self, terrain, force_limit, ee):
        self.name = "force-ee-force_%d" % ee
        self.terrain, self.fn_max, self.mu, self.ee = terrain, force_limit, terrain.mu, ee

    def link(self, variables):
        self.ee_force = variables["ee-force_%d" % self.ee]
        self.ee_motion = variables["ee-motion_%d" % self.ee]
        self.node_ids = self.ee_force.non_constant_nodes()
        self.n_rows = len(self.node_ids) * 5

    def _basis(self, f_node):
        phase = self.ee_force.phase_of(f_node)
        p = self.ee_motion.value_at_start_of_phase(phase)
        n = self.terrain.normalized_basis(0, p[X], p[Y])
        t1 = self.terrain.normalized_basis(1, p[X], p[Y])
        t2 = self.terrain.normalized_basis(2, p[X], p[Y])
        return phase, p, n, t1, t2

    def values(self):
        g = []
        mu = self.mu
        for f_node in self.node_ids:
            phase, p, n, t1, t2 = self._basis(f_node)
            f = self.ee_force.nodes[f_node, kPos]
            g += [f @ n, f @ (t1 - mu * n), f @ (t1 + mu * n), f @ (t2 - mu * n), f @ (t2 + mu * n)]
        return np.array(g)

    def bounds(self):
        out = []
        for _ in self.node_ids:
            out += [(0.0, self.fn_max), BOUND_SMALLER_ZERO, BOUND_GREATER_ZERO, BOUND_SMALLER_ZERO, BOUND_GREATER_ZERO]
        return out

    def block(self, var_set):
        mu = self.mu
        if var_set.name == self.ee_force.name:
            entries = {}
            for i, f_node in enumerate(self.node_ids):
                phase, p, n, t1, t2 = self._basis(f_node)
                for dim in (X, Y, Z):
                    idx = self.ee_force.opt_index(f_node, kPos, dim)
                    for r, val in enumerate((n[dim], t1[dim] - mu * n[dim], t1[dim] + mu * n[dim],
                                             t2[dim] - mu * n[dim], t2[dim] + mu * n[dim])):
                        entries[(5 * i + r, idx)] = val
            return Sp.from_entries((self.n_rows, var_set.rows()), entries)
        if var_set.name == self.ee_motion.name:
            entries = {}
            for i, f_node in enumerate(self.node_ids):
                phase, p, n, t1, t2 = self._basis(f_node)
                ee_node = self.ee_motion.node_at_start_of_phase(phase)
                f = self.ee_force.nodes[f_node, kPos]
                for dim in (X, Y):
                    dn = self.terrain.d_normalized_basis(0, dim, p[X], p[Y])
                    dt1 = self.terrain.d_normalized_basis(1, dim, p[X], p[Y])
                    dt2 = self.terrain.d_normalized_basis(2, dim, p[X], p[Y])
                    idx = self.ee_motion.opt_index(ee_node, kPos, dim)
                    for r, val in enumerate((f @ dn, f @ (dt1 - mu * dn), f @ (dt1 + mu * dn), f @ (dt2 - mu * dn),
                                             f @ (dt2 + mu * dn))):
                        entries[(5 * i + r, idx)] = val
            return Sp.from_entries((self.n_rows, var_set.rows()), entries)
        return None


class TorqueConstraint(ConstraintSet):
    """constraints/torque_constraint.cc:34-159 (node-based; used when dt_constraint_torque_ <= 0).  Three rows per stance
    torque node.  The friction-moment bound is evaluated with the force at the start of the node's phase as it is when
    the bounds are asked for (:94-112); the ee-motion block uses the torque at the START of the phase (:141); the
    ee-force block is empty (:155-157)."""

    def __init__(self, terrain, tx_min, tx_max, ty_min, ty_max, k_friction, ee):
        self.name = "torque-ee-torque_%d" % ee
        self.terrain, self.mu, self.ee = terrain, terrain.mu, ee
        self.lim, self.kf = (tx_min, tx_max, ty_min, ty_max), k_friction

    def link(self, variables):
        self.ee_torque = variables["ee-torque_%d" % self.ee]
        self.ee_motion = variables["ee-motion_%d" % self.ee]
        self.ee_force = variables["ee-force_%d" % self.ee]
        self.node_ids = self.ee_torque.non_constant_nodes()
        self.n_rows = len(self.node_ids) * 3

    def _frame(self, t_node):
        phase = self.ee_torque.phase_of(t_node)
        p = self.ee_motion.value_at_start_of_phase(phase)
        return phase, p, [self.terrain.normalized_basis(b, p[X], p[Y]) for b in (1, 2, 0)]    # t1, t2, n

    def values(self):
        g = []
        for t_node in self.node_ids:
            phase, p, frame = self._frame(t_node)
            tau = self.ee_torque.nodes[t_node, kPos]
            g += [tau @ b for b in frame]
        return np.array(g)

    def bounds(self):
        out = []
        for t_node in self.node_ids:
            phase, p, frame = self._frame(t_node)
            f = self.ee_force.value_at_start_of_phase(phase)
            tz_limit = self.kf * self.mu * (f @ frame[2])
            out += [(self.lim[0], self.lim[1]), (self.lim[2], self.lim[3]), (-tz_limit, tz_limit)]
        return out

    def block(self, var_set):
        if var_set.name == self.ee_torque.name:
            entries = {}
            for i, t_node in enumerate(self.node_ids):
                phase, p, frame = self._frame(t_node)
                for dim in (X, Y, Z):
                    idx = self.ee_torque.opt_index(t_node, kPos, dim)
                    for r in range(3):
                        entries[(3 * i + r, idx)] = frame[r][dim]
            return Sp.from_entries((self.n_rows, var_set.rows()), entries)
        if var_set.name == self.ee_motion.name:
            entries = {}
            for i, t_node in enumerate(self.node_ids):
                phase, p, frame = self._frame(t_node)
                ee_node = self.ee_motion.node_at_start_of_phase(phase)
                tau = self.ee_torque.value_at_start_of_phase(phase)
                for dim in (X, Y):
                    idx = self.ee_motion.opt_index(ee_node, kPos, dim)
                    for r, b in enumerate((1, 2, 0)):
                        entries[(3 * i + r, idx)] = tau @ self.terrain.d_normalized_basis(b, dim, p[X], p[Y])
            return Sp.from_entries((self.n_rows, var_set.rows()), entries)
        return None


class SwingConstraint(ConstraintSet):
    """constraints/swing_constraint.cc:35-106; t_swing_avg_ = 0.3 (swing_constraint.h:68)."""
    t_swing_avg = 0.3

    def __init__(self, ee_motion_name):
        self.name = "swing-" + ee_motion_name
        self.var_name = ee_motion_name

    def link(self, variables):
        self.ee_motion = variables[self.var_name]
        self.node_ids = self.ee_motion.non_constant_nodes()
        self.n_rows = len(self.node_ids) * 2 * 2

    def values(self):
        g = []
        nodes = self.ee_motion.nodes
        for n in self.node_ids:
            prev, nxt = nodes[n - 1, kPos, :2], nodes[n + 1, kPos, :2]
            distance_xy = nxt - prev
            xy_center = prev + 0.5 * distance_xy
            des_vel_center = distance_xy / self.t_swing_avg
            for dim in (X, Y):
                g.append(nodes[n, kPos, dim] - xy_center[dim])
                g.append(nodes[n, kVel, dim] - des_vel_center[dim])
        return np.array(g)

    def bounds(self):
        return [BOUND_ZERO] * self.n_rows

    def block(self, var_set):
        if var_set.name != self.var_name:
            return None
        entries = {}
        oi = self.ee_motion.opt_index
        row = 0
        for n in self.node_ids:
            for dim in (X, Y):
                entries[(row, oi(n, kPos, dim))] = 1.0
                entries[(row, oi(n + 1, kPos, dim))] = -0.5
                entries[(row, oi(n - 1, kPos, dim))] = -0.5
                row += 1
                entries[(row, oi(n, kVel, dim))] = 1.0
                entries[(row, oi(n + 1, kPos, dim))] = -1.0 / self.t_swing_avg
                entries[(row, oi(n - 1, kPos, dim))] = +1.0 / self.t_swing_avg
                row += 1
        return Sp.from_entries((self.n_rows, var_set.rows()), entries)


class SplineAccConstraint(ConstraintSet):
    """constraints/spline_acc_constraint.cc:34-90."""

    def __init__(self, spline, var_name):
        self.name = "splineacc-" + var_name
        self.spline, self.var_name = spline, var_name
        self.n_dim = spline.vars.dim
        self.n_junctions = len(spline.polys) - 1
        self.T = spline.poly_durations()
        self.n_rows = self.n_dim * self.n_junctions

    def values(self):
        g = np.zeros(self.n_rows)
        for j in range(self.n_junctions):
            acc_prev = self.spline.point_local(j, self.T[j])[kAcc]
            acc_next = self.spline.point_local(j + 1, 0.0)[kAcc]
            g[j * self.n_dim:(j + 1) * self.n_dim] = acc_prev - acc_next
        return g

    def bounds(self):
        return [BOUND_ZERO] * self.n_rows

    def block(self, var_set):
        if var_set.name != self.var_name:
            return None
        rows = []
        for j in range(self.n_junctions):
            acc_prev = self.spline.jacobian_wrt_nodes_local(j, self.T[j], kAcc)
            acc_next = self.spline.jacobian_wrt_nodes_local(j + 1, 0.0, kAcc)
            rows.append(acc_prev - acc_next)
        return Sp.vstack(rows)


class TotalDurationConstraint(ConstraintSet):
    """constraints/total_duration_constraint.cc:36-77."""

    def __init__(self, T_total, ee):
        self.name = "totalduration-%d" % ee
        self.T_total, self.ee, self.n_rows = T_total, ee, 1

    def link(self, variables):
        self.phase_durations = variables["ee-schedule%d" % self.ee]

    def values(self):
        return np.array([np.sum(self.phase_durations.values())])

    def bounds(self):
        return [(0.1, self.T_total - 0.2)]

    def block(self, var_set):
        if var_set.name != self.phase_durations.name:
            return None
        return Sp.from_entries((1, var_set.rows()), {(0, c): 1.0 for c in range(var_set.rows())})


# ------------------------------------------------------------------------------------------------ costs
class NodeCost:
    """costs/node_cost.cc:36-76.  (`double cost;` is uninitialised in the reference, :56; the sum starts from 0 here.)"""

    def __init__(self, var_name, deriv, dim, weight):
        self.var_name, self.deriv, self.dim, self.weight = var_name, deriv, dim, weight

    def link(self, variables):
        self.vars = variables[self.var_name]

    def cost(self):
        c = 0.0
        for n in range(self.vars.nodes.shape[0]):
            c += self.weight * math.pow(self.vars.nodes[n, self.deriv, self.dim], 2)
        return c

    def gradient(self, var_set):
        g = np.zeros(var_set.rows())
        if var_set.name == self.var_name:
            for i, infos in enumerate(self.vars.index_info):
                for (n, d, k) in infos:
                    if d == self.deriv and k == self.dim:
                        g[i] += self.weight * 2.0 * self.vars.nodes[n, self.deriv, self.dim]
        return g


def cost_sample_times(T, dt):
    """EnergyCost::GetSampleTimes (energy_cost.cc:52-66) == AngularMomentumCost::GetSampleTimes (:51-65)."""
    if dt <= 0.0:
        return [0.0, T]
    ts, t = [], 0.0
    while t <= T + 1e-9:
        ts.append(t)
        t += dt
    return ts


class EnergyCost:
    """costs/energy_cost.cc:38-163: sum over samples of weight * dt * sum_ee (|f|^2 + torque_weight |tau|^2)."""

    def __init__(self, splines, weight, torque_weight, dt):
        self.sh, self.weight, self.torque_weight, self.dt = splines, weight, torque_weight, dt

    def link(self, variables):
        pass

    def times(self):
        return cost_sample_times(self.sh.base_linear.total_time(), self.dt)

    def cost(self):                                                   # :68-90
        if self.weight <= 0.0:
            return 0.0
        c = 0.0
        for t in self.times():
            inst = 0.0
            for ee in range(len(self.sh.ee_force)):
                f = self.sh.ee_force[ee].point(t)[kPos]
                tau = self.sh.ee_torque[ee].point(t)[kPos]
                inst += float(f @ f) + self.torque_weight * float(tau @ tau)
            c += self.weight * inst * (self.dt if self.dt > 0.0 else 1.0)
        return c

    def gradient(self, var_set):                                      # :92-161
        g = np.zeros(var_set.rows())
        if self.weight <= 0.0:
            return g
        wdt = self.weight * (self.dt if self.dt > 0.0 else 1.0)
        for t in self.times():
            for ee in range(len(self.sh.ee_force)):
                f = self.sh.ee_force[ee].point(t)[kPos]
                tau = self.sh.ee_torque[ee].point(t)[kPos]
                if var_set.name == "ee-force_%d" % ee:
                    g += ((2.0 * wdt) * f) @ self.sh.ee_force[ee].jacobian_wrt_nodes(t, kPos).dense()
                if var_set.name == "ee-torque_%d" % ee and self.torque_weight != 0.0:
                    g += ((2.0 * wdt * self.torque_weight) * tau) @ self.sh.ee_torque[ee].jacobian_wrt_nodes(t, kPos).dense()
                if var_set.name == "ee-schedule%d" % ee:
                    g += ((2.0 * wdt) * f) @ self.sh.ee_force[ee].jacobian_of_pos_wrt_durations(t).dense()
                    if self.torque_weight != 0.0:
                        g += ((2.0 * wdt * self.torque_weight) * tau) @ self.sh.ee_torque[ee].jacobian_of_pos_wrt_durations(t).dense()
        return g


class EEBasePosCost:
    """costs/ee_base_pos_cost.cc:38-160: during swing, weight * |b_R_w (p_ee - p_base) - p_ref_B|^2 per sample (no dt factor)."""

    def __init__(self, splines, ee, p_ref_B, weight, dt):
        self.sh, self.ee, self.p_ref_B, self.weight, self.dt = splines, ee, np.asarray(p_ref_B, dtype=float), weight, dt

    def link(self, variables):
        pass

    def _swing_samples(self):
        for t in cost_sample_times(self.sh.base_linear.total_time(), self.dt):
            if self.sh.phase_durations[self.ee].is_contact_phase(t):      # :78-80
                continue
            r_W = self.sh.ee_motion[self.ee].point(t)[kPos] - self.sh.base_linear.point(t)[kPos]
            b_R_w = self.sh.angular_converter.rotation(t).T
            yield t, r_W, b_R_w, b_R_w @ r_W - self.p_ref_B

    def cost(self):                                                   # :67-96
        if self.weight <= 0.0:
            return 0.0
        c = 0.0
        for _, _, _, e in self._swing_samples():
            c += self.weight * float(e @ e)
        return c

    def gradient(self, var_set):                                      # :98-160 (no ee-schedule block, :154-157)
        g = np.zeros(var_set.rows())
        if self.weight <= 0.0:
            return g
        for t, r_W, b_R_w, e in self._swing_samples():
            m = (2.0 * self.weight) * e
            if var_set.name == "ee-motion_%d" % self.ee:
                g += (m @ b_R_w) @ self.sh.ee_motion[self.ee].jacobian_wrt_nodes(t, kPos).dense()
            if var_set.name == "base-lin":
                g += (-m @ b_R_w) @ self.sh.base_linear.jacobian_wrt_nodes(t, kPos).dense()
            if var_set.name == "base-ang":
                g += m @ self.sh.angular_converter.d_rot_vec_mult(t, r_W, True).dense()
        return g


class AngularMomentumCost:
    """costs/angular_momentum_cost.cc:39-208, SingleRigidBodyDynamics branch: sum of weight * dt * |R I_b R^T omega|^2."""

    def __init__(self, splines, model, weight, dt):
        self.sh, self.I_b, self.weight, self.dt = splines, model.I_b.dense(), weight, dt

    def link(self, variables):
        pass

    def times(self):
        return cost_sample_times(self.sh.base_linear.total_time(), self.dt)

    def cost(self):                                                   # :67-101
        if self.weight <= 0.0:
            return 0.0
        conv = self.sh.angular_converter
        wdt = self.weight * (self.dt if self.dt > 0.0 else 1.0)
        c = 0.0
        for t in self.times():
            R = conv.rotation(t)
            v = (R @ self.I_b @ R.T) @ conv.angular_velocity(t)
            c += wdt * float(v @ v)
        return c

    def gradient(self, var_set):                                      # :103-206
        g = np.zeros(var_set.rows())
        if self.weight <= 0.0 or var_set.name != "base-ang":
            return g
        conv = self.sh.angular_converter
        wdt = self.weight * (self.dt if self.dt > 0.0 else 1.0)
        for t in self.times():
            R = conv.rotation(t)
            omega = conv.angular_velocity(t)
            v = (R @ self.I_b @ R.T) @ omega
            m = (2.0 * wdt) * v
            u = R.T @ omega
            v2 = self.I_b @ u
            J_R_v2 = conv.d_rot_vec_mult(t, v2, False).dense()
            term_a = conv.d_rot_vec_mult(t, omega, True).dense()
            term_b = R.T @ conv.d_ang_vel_wrt_nodes(t).dense()
            J2 = R @ (self.I_b @ (term_a + term_b))
            g += m @ J_R_v2
            g += m @ J2
        return g


# ------------------------------------------------------------------------------------------------ formulation
class SplineHolder:
    """helpers/spline_holder.cc:35-69."""

    def __init__(self, base_lin, base_ang, base_poly_durations, ee_motion, ee_ang, ee_force, ee_torque, phase_durations,
                 durations_change):
        self.base_linear = NodeSpline(base_lin, base_poly_durations)
        self.base_angular = NodeSpline(base_ang, base_poly_durations)
        self.phase_durations = phase_durations
        self.ee_motion, self.ee_ang, self.ee_force, self.ee_torque = [], [], [], []
        for ee in range(len(ee_motion)):
            for nodes, out in ((ee_motion[ee], self.ee_motion), (ee_ang[ee], self.ee_ang), (ee_force[ee], self.ee_force),
                               (ee_torque[ee], self.ee_torque)):
                if durations_change:
                    out.append(PhaseSpline(nodes, phase_durations[ee]))
                else:
                    out.append(NodeSpline(nodes, nodes.convert_phase_to_poly_durations(phase_durations[ee].durations)))
        self.angular_converter = EulerConverter(self.base_angular)   # nlp_formulation.cc:113-116 (Euler is the default)


# towr::Parameters::ConstraintName (parameters.h:142-153)
C_DYNAMIC, C_EE_ROM, C_TOTAL_TIME, C_TERRAIN, C_TERRAIN_HARD, C_FORCE, C_TORQUE, C_SWING, C_BASE_ROM, C_BASE_ACC, C_BASE_HEIGHT = range(11)
COST_FORCES, COST_EE_MOTION, COST_ENERGY, COST_ANG_MOM = 0, 1, 2, 3


class Formulation:
    """towr::NlpFormulation (nlp_formulation.cc:76-681) + towr::Parameters (parameters.cc:40-175) + ifopt::Problem.
    `spec` is any object with the fields of oracle_capi.h's orc_spec (plain input data: robot constants, gait timings,
    states, flags) — the ctypes struct of oracle_py works, nothing of the C++ oracle is called."""

    def __init__(self, spec):
        s = spec
        self.n_ee = n_ee = int(s.n_ee)
        arr = lambda a, *shape: np.array(np.ctypeslib.as_array(a), dtype=float).reshape(*shape) if not isinstance(a, np.ndarray) else a
        if int(s.terrain_id) == 8:
            n = int(s.terrain_table_n)
            self.terrain = TableTerrain(int(s.terrain_table_mode), list(s.terrain_table_x)[:n], list(s.terrain_table_h)[:n])
        else:
            self.terrain = TERRAINS[int(s.terrain_id)]()
        self.model = SingleRigidBodyDynamics(float(s.mass), [float(v) for v in s.inertia], n_ee)
        self.kin = {"nominal": [np.array(list(s.nominal_stance[e])) for e in range(n_ee)],
                    "max_dev": [np.array(list(s.max_dev[e])) for e in range(n_ee)],
                    "min_dev": [np.array(list(s.min_dev[e])) for e in range(n_ee)]}
        self.ee_phase_durations = [[float(s.phase_durations[e][i]) for i in range(int(s.n_phases[e]))] for e in range(n_ee)]
        self.in_contact_at_start = [bool(s.in_contact_at_start[e]) for e in range(n_ee)]
        self.constraint_ids = [int(s.constraints[i]) for i in range(int(s.n_constraints))]
        self.spec = s
        self.initial_base_lin = np.array(list(s.initial_base_lin)).reshape(2, 3)
        self.initial_base_ang = np.array(list(s.initial_base_ang)).reshape(2, 3)
        self.final_base_lin = np.array(list(s.final_base_lin)).reshape(2, 3)
        self.final_base_ang = np.array(list(s.final_base_ang)).reshape(2, 3)
        self.initial_ee_W = [np.array(list(s.initial_ee_W[e])) for e in range(n_ee)]
        self._build()

    # parameters.cc:114-175
    def total_time(self):
        t = 0.0
        for d in self.ee_phase_durations[0]:
            t += d
        return t

    def base_poly_durations(self):
        out = []
        dt = float(self.spec.duration_base_polynomial)
        t_left = self.total_time()
        while t_left > 1e-10:
            out.append(dt if t_left > dt else t_left)
            t_left -= dt
        return out

    def optimize_timings(self):
        return C_TOTAL_TIME in self.constraint_ids

    def _dims(self, flags):
        return [d for d in range(3) if int(flags[d])]

    def _build(self):
        s, n_ee, T = self.spec, self.n_ee, self.total_time()
        zero3 = np.zeros(3)
        # MakeBaseVariables, nlp_formulation.cc:121-186
        base_durations = self.base_poly_durations()
        n_nodes = len(base_durations) + 1
        lin = NodesVariablesAll(n_nodes, 3, "base-lin")
        fx, fy = self.final_base_lin[kPos][X], self.final_base_lin[kPos][Y]
        final_pos = np.array([fx, fy, self.terrain.height(fx, fy) - self.kin["nominal"][0][Z]])
        lin.set_by_linear_interpolation(self.initial_base_lin[kPos], final_pos, T)
        lin.add_start_bound(kPos, [X, Y, Z], self.initial_base_lin[kPos])
        lin.add_start_bound(kVel, [X, Y, Z], self.initial_base_lin[kVel])
        lin.add_start_bound(kAcc, [X, Y, Z], zero3)                   # no variable has deriv kAcc: a no-op, as there
        lin.add_final_bound(kPos, self._dims(s.bounds_final_lin_pos), self.final_base_lin[kPos])
        lin.add_final_bound(kVel, self._dims(s.bounds_final_lin_vel), self.final_base_lin[kVel])
        ang = NodesVariablesAll(n_nodes, 3, "base-ang")
        ang.set_by_linear_interpolation(self.initial_base_ang[kPos], self.final_base_ang[kPos], T)
        ang.add_start_bound(kPos, [X, Y, Z], self.initial_base_ang[kPos])
        ang.add_start_bound(kVel, [X, Y, Z], self.initial_base_ang[kVel])
        ang.add_final_bound(kPos, self._dims(s.bounds_final_ang_pos), self.final_base_ang[kPos])
        ang.add_final_bound(kVel, self._dims(s.bounds_final_ang_vel), self.final_base_ang[kVel])
        # MakeEndeffectorVariables :188-242
        ee_motion = []
        for ee in range(n_ee):
            nodes = NodesVariablesEEMotion(len(self.ee_phase_durations[ee]), self.in_contact_at_start[ee], "ee-motion_%d" % ee,
                                           int(s.ee_polynomials_per_swing_phase))
            yaw = self.final_base_ang[kPos][Z]
            w_R_b = rotation_base_to_world(np.array([0.0, 0.0, yaw]))
            final_ee = self.final_base_lin[kPos] + w_R_b @ self.kin["nominal"][ee]
            x, y = final_ee[X], final_ee[Y]
            final_ee = np.array([x, y, self.terrain.height(x, y)])
            nodes.set_by_linear_interpolation_relative_to_base(self.initial_ee_W[ee], final_ee, self.initial_base_lin[kPos],
                                                               self.final_base_lin[kPos], self.initial_base_ang[kPos],
                                                               self.final_base_ang[kPos], T)
            if int(s.enable_stance_tracking):
                stance_id = 0
                phase = 0
                phase_count = len(self.ee_phase_durations[ee])
                while phase < phase_count:                            # :218-231: `phase++` twice per stance found
                    node = nodes.node_at_start_of_phase(phase)
                    if nodes.is_constant_node(node):
                        if stance_id >= int(s.n_stance_positions[ee]):
                            raise IndexError("ee_stance_position_.at(ee).at(stance_id)")   # std::out_of_range there
                        pos = [float(s.ee_stance_position[ee][stance_id][0]), float(s.ee_stance_position[ee][stance_id][1])]
                        nodes.add_bounds(node, kPos, [X, Y], pos)
                        stance_id += 1
                        phase += 1
                    phase += 1
            else:
                nodes.add_start_bound(kPos, [X, Y, Z], self.initial_ee_W[ee])
                nodes.add_final_bound(kPos, [X, Y, Z], final_ee)
            ee_motion.append(nodes)
        # MakeEndeffectorAngVariables :244-287 (enable_stance_rpy_tracking off)
        ee_ang = []
        for ee in range(n_ee):
            nodes = NodesVariablesEEAng(len(self.ee_phase_durations[ee]), self.in_contact_at_start[ee], "ee-ang_%d" % ee,
                                        int(s.ee_polynomials_per_swing_phase))
            nodes.set_by_linear_interpolation(self.initial_base_ang[kPos], self.final_base_ang[kPos], T)
            nodes.add_start_bound(kPos, [0, 1, 2], self.initial_base_ang[kPos])
            nodes.add_final_bound(kPos, [0, 1, 2], self.final_base_ang[kPos])
            ee_ang.append(nodes)
        # MakeForceVariables :289-318, MakeTorqueVariables :320-346
        ee_force, ee_torque = [], []
        for ee in range(n_ee):
            nodes = NodesVariablesEELoad(len(self.ee_phase_durations[ee]), self.in_contact_at_start[ee], "ee-force_%d" % ee,
                                         int(s.force_polynomials_per_stance_phase))
            f_stance = np.array([0.0, 0.0, self.model.m * self.model.g / n_ee])
            nodes.set_by_linear_interpolation(f_stance, f_stance, T)
            nodes.add_start_bound(kPos, [X, Y, Z], f_stance)
            nodes.add_final_bound(kPos, [X, Y, Z], f_stance)
            nodes.add_start_bound(kVel, [X, Y, Z], zero3)
            nodes.add_final_bound(kVel, [X, Y, Z], zero3)
            ee_force.append(nodes)
        for ee in range(n_ee):
            nodes = NodesVariablesEELoad(len(self.ee_phase_durations[ee]), self.in_contact_at_start[ee], "ee-torque_%d" % ee,
                                         int(s.torque_polynomials_per_stance_phase))
            nodes.set_by_linear_interpolation(zero3, zero3, T)
            nodes.add_start_bound(kPos, [X, Y, Z], zero3)
            nodes.add_final_bound(kPos, [X, Y, Z], zero3)
            nodes.add_start_bound(kVel, [X, Y, Z], zero3)
            nodes.add_final_bound(kVel, [X, Y, Z], zero3)
            ee_torque.append(nodes)
        # MakeContactScheduleVariables :348-363
        schedule = [PhaseDurations(ee, self.ee_phase_durations[ee], self.in_contact_at_start[ee],
                                   float(s.bound_phase_duration[0]), float(s.bound_phase_duration[1])) for ee in range(n_ee)]
        # GetVariableSets :76-119
        self.var_sets = [lin, ang] + ee_motion + ee_ang + ee_force + ee_torque + (schedule if self.optimize_timings() else [])
        self.splines = SplineHolder(lin, ang, base_durations, ee_motion, ee_ang, ee_force, ee_torque, schedule,
                                    self.optimize_timings())
        by_name = {v.name: v for v in self.var_sets}
        # GetConstraints :365-398 and the Make* functions :400-602
        sh = self.splines
        self.con_sets = []
        for cid in self.constraint_ids:
            if cid == C_DYNAMIC:
                self.con_sets.append(DynamicConstraint(self.model, T, float(s.dt_constraint_dynamic), sh))
            elif cid == C_EE_ROM:
                for ee in range(n_ee):
                    self.con_sets.append(RangeOfMotionConstraint(self.kin, T, float(s.dt_constraint_range_of_motion), ee, sh))
            elif cid == C_BASE_ROM:
                self.con_sets.append(BaseMotionConstraint(T, float(s.dt_constraint_base_motion), NO_BOUND, NO_BOUND, NO_BOUND, sh))
            elif cid == C_TOTAL_TIME:
                for ee in range(n_ee):
                    self.con_sets.append(TotalDurationConstraint(T, ee))
            elif cid == C_TERRAIN:
                for ee in range(n_ee):
                    self.con_sets.append(TerrainConstraint(self.terrain, "ee-motion_%d" % ee, 0.02, float("inf")))
            elif cid == C_TERRAIN_HARD:
                for ee in range(n_ee):
                    self.con_sets.append(TerrainConstraintHard(self.terrain, T, float(s.dt_constraint_range_of_motion), ee, sh))
            elif cid == C_FORCE:
                for ee in range(n_ee):
                    if float(s.dt_constraint_force) > 0.0:
                        self.con_sets.append(ForceConstraintDiscretized(self.terrain, T, float(s.dt_constraint_force),
                                                                        float(s.force_limit_in_normal_direction), ee, sh))
                    else:
                        self.con_sets.append(ForceConstraint(self.terrain, float(s.force_limit_in_normal_direction), ee))
            elif cid == C_TORQUE:
                for ee in range(n_ee):
                    if float(s.dt_constraint_torque) <= 0.0:
                        self.con_sets.append(TorqueConstraint(self.terrain, float(s.torque_tx_min), float(s.torque_tx_max),
                                                              float(s.torque_ty_min), float(s.torque_ty_max),
                                                              float(s.torque_k_friction), ee))
                        continue
                    self.con_sets.append(TorqueConstraintDiscretized(self.terrain, T, float(s.dt_constraint_torque),
                                                                     float(s.torque_tx_min), float(s.torque_tx_max),
                                                                     float(s.torque_ty_min), float(s.torque_ty_max),
                                                                     float(s.torque_k_friction), ee, sh))
            elif cid == C_SWING:
                for ee in range(n_ee):
                    self.con_sets.append(SwingConstraint("ee-motion_%d" % ee))
            elif cid == C_BASE_ACC:
                self.con_sets.append(SplineAccConstraint(sh.base_linear, "base-lin"))
                self.con_sets.append(SplineAccConstraint(sh.base_angular, "base-ang"))
            elif cid == C_BASE_HEIGHT:
                self.con_sets.append(BaseHeightConstraint(self.terrain, 0.4, "base-lin"))
            else:
                raise RuntimeError("constraint not defined!")
        for i in range(int(s.n_ee_linear)):                           # :373-375, MakeEELinearConstraint :682-704
            motion = int(s.ee_linear_target[i]) == 0
            terms = [(int(s.ee_linear_term_ee[i][j]), int(s.ee_linear_term_dim[i][j]), float(s.ee_linear_term_coeff[i][j]))
                     for j in range(int(s.ee_linear_n_terms[i]))]
            self.con_sets.append(EELinearConstraint(
                terms, sh.ee_motion if motion else sh.ee_ang,
                [("ee-motion_%d" if motion else "ee-ang_%d") % ee for ee in range(n_ee)],
                kVel if int(s.ee_linear_deriv[i]) else kPos, float(s.ee_linear_tolerance[i]), T, float(s.ee_linear_dt[i])))
        for c in self.con_sets:
            c.link(by_name)
        # GetCosts :604-681 (NodeCost terms only)
        self.costs = []
        for i in range(int(s.n_costs)):
            cid, w = int(s.cost_ids[i]), float(s.cost_weights[i])
            if cid == COST_FORCES:
                for ee in range(n_ee):
                    for dim in (X, Y, Z):
                        self.costs += [NodeCost("ee-force_%d" % ee, kPos, dim, w), NodeCost("ee-torque_%d" % ee, kPos, dim, w),
                                       NodeCost("ee-force_%d" % ee, kVel, dim, 0.1 * w),
                                       NodeCost("ee-torque_%d" % ee, kVel, dim, 0.1 * w)]
            elif cid == COST_EE_MOTION:
                for ee in range(n_ee):
                    self.costs += [NodeCost("ee-motion_%d" % ee, kVel, X, w), NodeCost("ee-motion_%d" % ee, kVel, Y, w),
                                   NodeCost("ee-motion_%d" % ee, kVel, Z, 0.5 * w)]
            elif cid == COST_ENERGY:                                      # :636-639
                self.costs.append(EnergyCost(sh, w, float(s.energy_cost_torque_weight), float(s.dt_cost_energy)))
            elif cid == COST_ANG_MOM:                                     # :640-643
                self.costs.append(AngularMomentumCost(sh, self.model, w, float(s.dt_cost_ang_mom)))
            else:
                raise RuntimeError("cost not defined!")
        # individual NodeCost terms, as the hand-built examples add them (test/squat_jump_example.cc:205-211)
        set_names = ("base-lin", "base-ang", "ee-motion_%d", "ee-ang_%d", "ee-force_%d", "ee-torque_%d")
        for i in range(int(s.n_node_costs)):
            k = int(s.node_cost_set[i])
            name = set_names[k] if k < 2 else set_names[k] % int(s.node_cost_ee[i])
            self.costs.append(NodeCost(name, int(s.node_cost_deriv[i]), int(s.node_cost_dim[i]), float(s.node_cost_weight[i])))
        # :612-626: one EEBasePosCost per foot, reference r_B0 = b_R_w0 (initial_ee_W - initial_base.lin.p)
        if int(s.enable_swing_ee_base_pos_tracking) and float(s.swing_ee_base_pos_tracking_weight) > 0.0:
            b_R_w0 = rotation_base_to_world(self.initial_base_ang[kPos]).T
            for ee in range(n_ee):
                r_B0 = b_R_w0 @ (self.initial_ee_W[ee] - self.initial_base_lin[kPos])
                self.costs.append(EEBasePosCost(sh, ee, r_B0, float(s.swing_ee_base_pos_tracking_weight),
                                                float(s.dt_cost_swing_ee_base_pos_tracking)))
        for c in self.costs:
            c.link(by_name)

    # ---- ifopt::Problem
    def n_vars(self):
        return sum(v.rows() for v in self.var_sets)

    def n_rows(self):
        return sum(c.n_rows for c in self.con_sets)

    def x0(self):
        return np.concatenate([v.values() for v in self.var_sets])

    def var_bounds(self):
        b = [bd for v in self.var_sets for bd in v.bounds]
        return np.array([x[0] for x in b]), np.array([x[1] for x in b])

    def con_bounds(self):
        b = [bd for c in self.con_sets for bd in c.bounds()]
        return np.array([x[0] for x in b]), np.array([x[1] for x in b])

    def set_variables(self, x):
        col = 0
        for v in self.var_sets:
            v.set_variables(np.asarray(x[col:col + v.rows()], dtype=float))
            col += v.rows()

    def eval_g(self, x):
        self.set_variables(x)
        return np.concatenate([np.asarray(c.values(), dtype=float) for c in self.con_sets])

    def eval_jac(self, x):
        """(rows, cols, vals) of the stored entries in row-major order — what Ipopt's eval_jac_g receives."""
        self.set_variables(x)
        rows, cols, vals = [], [], []
        row0 = 0
        for c in self.con_sets:
            col0 = 0
            for v in self.var_sets:
                blk = c.block(v)
                if blk is not None:
                    assert blk.shape == (c.n_rows, v.rows()), (c.name, v.name, blk.shape)
                    r, cc, vv = blk.stored()
                    rows.append(r + row0)
                    cols.append(cc + col0)
                    vals.append(vv)
                col0 += v.rows()
            row0 += c.n_rows
        rows, cols, vals = np.concatenate(rows), np.concatenate(cols), np.concatenate(vals)
        order = np.lexsort((cols, rows))
        return rows[order], cols[order], vals[order]

    def eval_f(self, x):
        self.set_variables(x)
        f = 0.0
        for c in self.costs:
            f += c.cost()
        return f

    def eval_grad_f(self, x):
        self.set_variables(x)
        parts = []
        for v in self.var_sets:
            g = np.zeros(v.rows())
            for c in self.costs:
                g += c.gradient(v)
            parts.append(g)
        return np.concatenate(parts)
