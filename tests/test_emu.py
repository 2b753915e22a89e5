"""Kernel arithmetic (the device headers compiled for the host, tests/emu) against the oracle, no GPU.
Checks the table interpretation (items, entries, B constants) and the per-sample math at the parity
tolerance of BASELINE.json: 1e-12 relative / 1e-14 absolute."""
import ctypes as C

import numpy as np
import pytest

from util import CASES, GAP_CASES, check_against_oracle, is_rt, mk, perturbed, tol_ratio

DP = C.POINTER(C.c_double)


def emu_eval(E, spec, st, x):
    g = np.full(st.n_rows, np.nan)
    v = np.full(st.nnz, np.nan)
    rc = E.emu_eval(C.byref(spec), x.ctypes.data_as(DP), g.ctypes.data_as(DP), v.ctypes.data_as(DP))
    assert rc == 0, E.emu_last_error()
    return g, v


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_emulated_kernel_matches_oracle(product, oracle, emu, case):
    T, O = product, oracle
    s = mk(T, case)
    st = T.Structure(s)
    x0 = st.x0(s)
    rt = is_rt(case)
    for x in (x0, perturbed(x0, 11, 0.02 if rt else 0.1), perturbed(x0, 12, 0.05 if rt else 0.3)):
        g, v = emu_eval(emu, s, st, x)
        check_against_oracle(O, mk(O, case), st, x, g, v, rt)


@pytest.mark.parametrize("case", GAP_CASES, ids=[c[0] for c in GAP_CASES])
def test_emulated_kernel_on_gap_superset_pattern(product, oracle, emu, case):
    """Gap: the declared pattern is a superset of what the oracle stores at any x; equal values there, exact zeros elsewhere."""
    T, O = product, oracle
    s = mk(T, case)
    st = T.Structure(s)
    x0 = st.x0(s)
    n_extra = []
    for x in (x0, perturbed(x0, 3, 0.02), perturbed(x0, 4, 0.05)):
        g, v = emu_eval(emu, s, st, x)
        n_extra.append(check_against_oracle(O, mk(O, case), st, x, g, v, is_rt(case), superset=True))
    assert len(set(n_extra)) > 1          # the oracle's pattern really depends on x here


def test_emulated_kernel_segment_capacity(product, oracle, emu, monkeypatch):
    """Small G capacity -> many segments; results must not depend on the segmentation."""
    T, O = product, oracle
    s = T.make_spec(T.Anymal, T.C1, 2.0, goal=(1.1, 0.1, 0.2))
    st = T.Structure(s)
    x = perturbed(st.x0(s), 5, 0.1)
    ref = emu_eval(emu, s, st, x)
    monkeypatch.setenv("TOWR_B200_GCAP", "600")
    alt = emu_eval(emu, s, st, x)
    assert np.array_equal(ref[0], alt[0]) and np.array_equal(ref[1], alt[1])


def fine_dt_spec(M, robot, combo, T_total, dt, optimise=False):
    """BASELINE config 5's sampling convention (tools/mixed_sweep.py): dynamic dt, range of motion 0.8 dt, force 0.2 dt."""
    s = M.make_spec(robot, combo, T_total, optimize_durations=optimise)
    s.dt_constraint_dynamic = dt
    s.dt_constraint_range_of_motion = 0.8 * dt
    s.dt_constraint_force = 0.2 * dt
    return s


def test_emulated_kernel_program_parts(product, oracle, emu, monkeypatch):
    """Classes whose tables exceed one SM are split into parts (pack.cc: runs of segments with their own tables, one launch
    each).  A forced split of a small class must not change a bit; a class that needs the split (monoped, dt 0.01: 87 071
    stored entries, more than 8191 B constants) must match the oracle."""
    T, O = product, oracle
    s = T.make_spec(T.Anymal, T.C1, 2.0, goal=(1.1, 0.1, 0.2), optimize_durations=True)
    st = T.Structure(s)
    x = perturbed(st.x0(s), 5, 0.02)
    ref = emu_eval(emu, s, st, x)
    monkeypatch.setenv("TOWR_B200_MAX_BLOB", "24000")
    alt = emu_eval(emu, s, st, x)
    monkeypatch.delenv("TOWR_B200_MAX_BLOB")
    assert np.array_equal(ref[0], alt[0]) and np.array_equal(ref[1], alt[1])
    s = fine_dt_spec(T, T.Monoped, T.C0, 2.4, 0.01)
    st = T.Structure(s)
    assert st.nnz == 87071
    x = perturbed(st.x0(s), 6, 0.05)
    g, v = emu_eval(emu, s, st, x)
    check_against_oracle(O, fine_dt_spec(O, O.Monoped, 0, 2.4, 0.01), st, x, g, v, False)


def test_emulated_kernel_reproduces_committed_fixtures(product, emu):
    """tests/golden/*.npz (oracle outputs frozen by tests/golden/make_golden.py)."""
    import glob
    import os
    from util import CASES
    T = product
    by_name = {c[0]: c for c in CASES}
    files = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))
    assert files
    for f in files:
        z = np.load(f)
        s = mk(T, by_name[os.path.basename(f)[:-4]])
        st = T.Structure(s)
        g, v = emu_eval(emu, s, st, np.ascontiguousarray(z["x"]))
        rows, cols = st.pattern()
        assert st.nnz == int(z["nnz"])
        assert np.array_equal(rows[z["jac_idx"]], z["jac_rows"]) and np.array_equal(cols[z["jac_idx"]], z["jac_cols"])
        assert tol_ratio(g, z["g"]) <= 1.0 and tol_ratio(v[z["jac_idx"]], z["jac_vals"]) <= 1.0

