"""The oracle against what can pin it without the reference's own golden vectors (there are none):
Hermite identities, closed forms, finite-difference Jacobians (the programmatic `derivative_test first-order` of
towr_ros_app.cc:105-108), physical known answers, structure sizes, and the committed oracle fixtures."""
import ctypes as C
import glob
import os

import numpy as np
import pytest

from util import CASES, mk, perturbed, tol_ratio

DP = C.POINTER(C.c_double)


def _p(a):
    return a.ctypes.data_as(DP)


def test_hermite_interpolation_identities(oracle):
    L = oracle.lib()
    rng = np.random.default_rng(0)
    for _ in range(20):
        p0, v0, p1, v1 = (rng.uniform(-2, 2, 3) for _ in range(4))
        T = rng.uniform(0.05, 1.5)
        pos, vel, acc = np.empty(3), np.empty(3), np.empty(3)
        L.orc_hermite_point(_p(p0), _p(v0), _p(p1), _p(v1), T, 0.0, _p(pos), _p(vel), _p(acc))
        assert np.array_equal(pos, p0) and np.array_equal(vel, v0)
        L.orc_hermite_point(_p(p0), _p(v0), _p(p1), _p(v1), T, T, _p(pos), _p(vel), _p(acc))
        assert np.allclose(pos, p1, rtol=0, atol=1e-12) and np.allclose(vel, v1, rtol=0, atol=1e-11)


def test_hermite_basis_matches_point_linearity(oracle):
    """pos(t) = sum_k basis_k(t) * node_k (cubic_hermite_polynomial.m:12-76 identity) and FD of dpos/dT."""
    L = oracle.lib()
    rng = np.random.default_rng(1)
    for _ in range(20):
        p0, v0, p1, v1 = (rng.uniform(-2, 2, 3) for _ in range(4))
        T = rng.uniform(0.1, 1.5)
        t = rng.uniform(0, T)
        b = np.empty(12)
        L.orc_hermite_basis(T, t, _p(b))
        pos, vel, acc = np.empty(3), np.empty(3), np.empty(3)
        L.orc_hermite_point(_p(p0), _p(v0), _p(p1), _p(v1), T, t, _p(pos), _p(vel), _p(acc))
        for k, out in enumerate((pos, vel, acc)):
            lin = b[4 * k] * p0 + b[4 * k + 1] * v0 + b[4 * k + 2] * p1 + b[4 * k + 3] * v1
            assert np.allclose(lin, out, rtol=1e-10, atol=1e-9)
        d = np.empty(3)
        L.orc_hermite_dpos_dT(_p(p0), _p(v0), _p(p1), _p(v1), T, t, _p(d))
        h = 1e-6
        pp, pm = np.empty(3), np.empty(3)
        L.orc_hermite_point(_p(p0), _p(v0), _p(p1), _p(v1), T + h, t, _p(pp), _p(vel), _p(acc))
        L.orc_hermite_point(_p(p0), _p(v0), _p(p1), _p(v1), T - h, t, _p(pm), _p(vel), _p(acc))
        assert np.allclose((pp - pm) / (2 * h), d, rtol=1e-5, atol=1e-6)


def test_sample_times_accumulate(oracle):
    """time_discretization_constraint.cc:41-49: repeated t += dt, final T appended (a duplicate when T/dt is integral)."""
    L = oracle.lib()
    out = np.empty(512)
    n = C.c_int()
    L.orc_sample_times(2.0, 0.1, _p(out), C.byref(n))
    assert n.value == 22
    t, ref = 0.0, [0.0]
    for _ in range(20):
        t += 0.1
        ref.append(t)
    ref.append(2.0)
    assert list(out[:22]) == ref
    L.orc_sample_times(2.0, 0.08, _p(out), C.byref(n))
    assert n.value == 27
    L.orc_sample_times(2.0, 0.02, _p(out), C.byref(n))
    assert n.value == 102


def test_terrain_known_answers(oracle):
    L = oracle.lib()
    h, dh, basis, db = C.c_double(), np.empty(2), np.empty(9), np.empty(18)
    L.orc_terrain(oracle.FlatID, 0.3, 0.1, C.byref(h), _p(dh), _p(basis), _p(db))
    assert h.value == 0.0 and not dh.any()
    assert np.array_equal(np.abs(basis), np.array([0, 0, 1, 1, 0, 0, 0, 1, 0], dtype=float)) and not db.any()
    L.orc_terrain(oracle.SlopeID, 1.5, 0.0, C.byref(h), _p(dh), _p(basis), _p(db))
    assert h.value == pytest.approx(0.35) and dh[0] == pytest.approx(0.7)
    n = basis[:3]
    assert np.linalg.norm(n) == pytest.approx(1.0) and n[0] == pytest.approx(-0.7 / np.sqrt(1.49))
    L.orc_terrain(oracle.GapID, 1.25, 0.0, C.byref(h), _p(dh), _p(basis), _p(db))
    assert h.value == pytest.approx(-1.5) and abs(dh[0]) < 1e-12 and db.any()      # parabola bottom, curvature present
    L.orc_terrain(oracle.BlockID, 0.7, 0.0, C.byref(h), _p(dh), _p(basis), _p(db))   # closed interval at the edge
    assert h.value == 0.0 and dh[0] == pytest.approx(0.5 / 0.03)


def test_static_stance_known_answer(oracle):
    """Initial guess of nlp_formulation.cc:303-307: f_z = m g / E per foot, zero velocity, identity R.
    With goal = start every dynamic row is 0 up to rounding, and RoM values equal the nominal stance."""
    O = oracle
    s = O.make_spec(O.Anymal, 1, 2.0, goal=(0.0, 0.0, 0.0))
    p = O.Problem(s)
    g = p.eval_g(p.x0())
    off = 0
    for name, n in p.con_sets():
        blk = g[off:off + n]
        if name == "dynamic":
            # t = 0: four feet on the ground with f_z = m g / 4 each -> exact balance; later the force splines ramp
            # to zero around the flight phases, so LZ rises up to m g (no force at all) and never above
            rows = blk.reshape(-1, 6)
            assert np.abs(rows[0]).max() < 1e-9
            assert np.abs(rows[:, [0, 1, 2, 3, 4]]).max() < 1e-9
            assert rows[:, 5].min() > -1e-9 and rows[:, 5].max() < 29.5 * 9.80665 + 1e-9
        if name == "rangeofmotion-0":
            assert np.allclose(blk.reshape(-1, 3), [0.34, 0.19, -0.42], atol=1e-12)
        off += n


@pytest.mark.parametrize("case", CASES[:6], ids=[c[0] for c in CASES[:6]])
def test_oracle_jacobian_matches_finite_differences(oracle, case):
    O = oracle
    p = O.Problem(mk(O, case))
    x = perturbed(p.x0(), 4, 0.05)
    r, c, v = p.eval_jac(x)
    rng = np.random.default_rng(9)
    for _ in range(3):
        d = rng.uniform(-1, 1, p.n_vars)
        h = 1e-6
        fd = (p.eval_g(x + h * d) - p.eval_g(x - h * d)) / (2 * h)
        jd = np.zeros(p.n_rows)
        np.add.at(jd, r, v * d[c])
        # piecewise terrains (Block/Stairs) are non-smooth at their edges: allow a handful of rows to disagree
        bad = np.abs(jd - fd) > 1e-5 * max(1.0, np.abs(fd).max())
        assert bad.sum() <= (3 if case[5] in (1, 2) else 0)


def test_oracle_reproduces_committed_fixtures(oracle):
    O = oracle
    files = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))
    assert len(files) >= 5
    by_name = {c[0]: c for c in CASES}
    for f in files:
        z = np.load(f)
        p = O.Problem(mk(O, by_name[os.path.basename(f)[:-4]]))
        r, c, v = p.eval_jac(z["x"])
        assert len(v) == int(z["nnz"])
        assert np.array_equal(r[z["jac_idx"]], z["jac_rows"]) and np.array_equal(c[z["jac_idx"]], z["jac_cols"])
        assert tol_ratio(p.eval_g(z["x"]), z["g"]) <= 1.0
        assert tol_ratio(v[z["jac_idx"]], z["jac_vals"]) <= 1.0
        assert abs(v.sum() - float(z["sum_vals"])) <= 1e-9 * float(z["sum_abs"])


def test_exact_powers_are_correctly_rounded(oracle):
    """oracle/splines.cc ExactPow (the device's Cube / Pow4 restated, used by orc_set_pow_mode(1)) equals the __float128
    product rounded to double; glibc's pow is within one ulp of it and differs for a small fraction of arguments."""
    L = oracle.lib()
    rng = np.random.default_rng(5)
    std, ex, qd = C.c_double(), C.c_double(), C.c_double()
    n, diff_std = 20000, 0
    for n_pow in (2, 3, 4):
        for x in rng.uniform(1e-3, 3.0, size=n):
            L.orc_pow_compare(float(x), n_pow, C.byref(std), C.byref(ex), C.byref(qd))
            assert ex.value == qd.value, (x, n_pow)
            assert abs(std.value - qd.value) <= np.spacing(qd.value)
            diff_std += std.value != qd.value
    assert diff_std < 0.01 * 3 * n
