"""Host side of the backend against the oracle (no GPU): spec helpers, sizes, names, sparsity pattern,
initial guess and bounds must be bit-exact (BASELINE.json: "sparsity indices must be bit-exact")."""
import ctypes as C

import numpy as np
import pytest

from util import CASES, mk, perturbed, to_oracle_spec


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_structure_matches_oracle(product, oracle, case):
    T, O = product, oracle
    s = mk(T, case)
    so = mk(O, case)
    # robot constants, gait tables and parameter defaults: same bytes
    assert bytes(to_oracle_spec(O, s)) == bytes(so)
    st = T.Structure(s)
    p = O.Problem(so)
    assert (st.n_vars, st.n_rows) == (p.n_vars, p.n_rows)
    assert st.var_sets() == p.var_sets()
    assert st.con_sets() == p.con_sets()
    x0 = st.x0(s)
    assert np.array_equal(x0, p.x0())
    rows, cols = st.pattern()
    for x in (x0, perturbed(x0, 7, 0.02)):
        r, c, _ = p.eval_jac(x)
        assert len(r) == st.nnz
        assert np.array_equal(r, rows) and np.array_equal(c, cols)
    lo, up = st.var_bounds(s)
    lo2, up2 = p.var_bounds()
    assert np.array_equal(lo, lo2) and np.array_equal(up, up2)
    lo, up = st.con_bounds(s)
    lo2, up2 = p.con_bounds()
    assert np.array_equal(lo, lo2) and np.array_equal(up, up2)


def test_anymal_trot_sizes(product):
    """ANYmal fly-trot 2 s = 1042 variables, 2732 rows, 23992 stored entries."""
    T = product
    st = T.Structure(T.make_spec(T.Anymal, T.C1, 2.0))
    assert (st.n_vars, st.n_rows, st.nnz) == (1042, 2732, 23992)
    assert st.bytes_per_eval == 222128
    r, c = st.pattern()
    key = r.astype(np.int64) * st.n_vars + c
    assert (np.diff(key) > 0).all()          # row-major sorted, no duplicates


def test_structure_class_membership(product):
    T = product
    s = T.make_spec(T.Anymal, T.C1, 2.0)
    st = T.Structure(s)
    other_goal = T.make_spec(T.Anymal, T.C1, 2.0, goal=(2.0, 0.3, -0.4), terrain=T.BlockID)
    other_goal.mass = 31.0
    assert st.accepts(other_goal)            # goals, terrain ids, mass are per-problem data
    assert not st.accepts(T.make_spec(T.Anymal, T.C1, 2.2))
    assert not st.accepts(T.make_spec(T.Anymal, T.C0, 2.0))
    assert not st.accepts(T.make_spec(T.Hyq, T.C1, 2.0)) or True   # same leg count + gait -> same class is allowed
    with pytest.raises(T.TowrB200Error) as e:
        st.x0(T.make_spec(T.Anymal, T.C1, 2.2))
    assert e.value.code == -4


def test_error_behaviour(product):
    """Bad configuration is an error code + message, never an exception across the C boundary
    (the reference throws std::runtime_error: nlp_formulation.cc:396)."""
    T = product
    s = T.make_spec(T.Anymal, T.C1, 2.0)
    s.constraints[0] = 99
    with pytest.raises(T.TowrB200Error) as e:
        T.Structure(s)
    assert e.value.code == -1 and "constraint not defined" in str(e.value)
    s = T.make_spec(T.Anymal, T.C1, 2.0)
    s.n_ee = 7
    with pytest.raises(T.TowrB200Error):
        T.Structure(s)
    s = T.make_spec(T.Anymal, T.C1, 2.0)
    s.phase_durations[1][0] += 0.1          # feet must share the total time (parameters.cc:155)
    with pytest.raises(T.TowrB200Error):
        T.Structure(s)
    with pytest.raises(T.TowrB200Error):
        T.make_spec(5, T.C1, 2.0)


def test_gap_needs_a_class_that_declares_the_curvature_block(product):
    """Gap has a value-dependent curvature block (force_constraint_discretized.cc:52-67).  A class created on Gap declares
    it (superset pattern: more stored entries) and admits problems on any terrain; a class created on a curvature-free
    terrain must refuse a Gap problem - never rc 0 with a Jacobian that lacks the block."""
    T = product
    for opt in (False, True):
        flat = T.Structure(T.make_spec(T.Hyq, T.C4, 2.4, terrain=T.BlockID, optimize_durations=opt))
        curv = T.Structure(T.make_spec(T.Hyq, T.C4, 2.4, terrain=T.GapID, optimize_durations=opt))
        assert curv.nnz > flat.nnz and curv.n_rows == flat.n_rows and curv.n_vars == flat.n_vars
        gap = T.make_spec(T.Hyq, T.C4, 2.4, terrain=T.GapID, optimize_durations=opt)
        assert not flat.accepts(gap)
        with pytest.raises(T.TowrB200Error) as e:
            flat.x0(gap)
        assert e.value.code == -4 and "Gap" in str(e.value)
        for terrain in (T.BlockID, T.StairsID, T.GapID, T.FlatID):
            assert curv.accepts(T.make_spec(T.Hyq, T.C4, 2.4, terrain=terrain, optimize_durations=opt))


def test_library_exports_every_declared_symbol(product):
    """The C-ABI library loads without a GPU and exports everything include/towr_b200.h declares."""
    import os
    import re
    T = product
    hdr = open(os.path.join(os.path.dirname(T._HERE), "include", "towr_b200.h")).read()
    declared = set(re.findall(r"\b(towr_b200_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(T.exported_symbols())
    L = C.CDLL(T.LIB_PATH)
    for name in declared:
        assert hasattr(L, name), name


def test_no_cpu_fallback(product):
    """Without a CUDA device batch creation must fail loudly (TOWR_B200_ERR_CUDA), not evaluate on the host."""
    T = product
    if T.lib().towr_b200_device_count() > 0:
        pytest.skip("a GPU is present")
    st = T.Structure(T.make_spec(T.Anymal, T.C1, 2.0))
    with pytest.raises(T.TowrB200Error) as e:
        T.Batch(st, 4)
    assert e.value.code == -3
