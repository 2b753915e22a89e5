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
    # a fresh problem: the node-based TorqueConstraint's friction-moment bound reads the CURRENT force nodes
    # (torque_constraint.cc:94-112); Ipopt asks for the bounds once, at the initial guess
    lo2, up2 = O.Problem(to_oracle_spec(O, s)).con_bounds()
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


def test_table_terrain_class_rules_and_validation(product, oracle):
    """terrain_ as data (terrain id 8): a class created from a table spec has the SAME pattern as the class of an analytic
    terrain (a piecewise-linear profile has no curvature) and admits analytic and table problems; a class created on an
    analytic terrain refuses a table problem (its kernels do not read the table); bad tables are argument errors; x0 and
    bounds follow the table's heights (final foothold on the terrain, nlp_formulation.cc:199-203)."""
    from util import set_terrain_table
    T, O = product, oracle
    flat = T.make_spec(T.Biped, T.C0, 2.4, terrain=T.FlatID)
    tab = T.make_spec(T.Biped, T.C0, 2.4)
    set_terrain_table(tab, T.TABLE_RAMPS, [0.3, 0.9], [0.0, 0.25])
    st_flat, st_tab = T.Structure(flat), T.Structure(tab)
    assert st_tab.nnz == st_flat.nnz and np.array_equal(st_tab.pattern()[0], st_flat.pattern()[0])
    assert st_tab.accepts(flat) and st_tab.accepts(tab)
    assert not st_flat.accepts(tab)
    with pytest.raises(T.TowrB200Error) as e:
        st_flat.x0(tab)
    assert e.value.code == -4 and "table" in str(e.value)
    po = O.Problem(to_oracle_spec(O, tab))
    assert np.array_equal(st_tab.x0(tab), po.x0())
    for a, b in zip(st_tab.var_bounds(tab), po.var_bounds()):
        assert np.array_equal(np.clip(a, -1e20, 1e20), np.clip(b, -1e20, 1e20))
    assert not np.array_equal(st_tab.x0(tab), st_tab.x0(flat))          # the final footholds sit on the ramp
    for bad in ((T.TABLE_STEPS, [0.5, 0.5], [0.1, 0.2]), (T.TABLE_STEPS, [], []), (7, [0.1], [0.1]),
                (T.TABLE_RAMPS, [0.1, 0.2], [0.0, float("nan")])):
        s = T.make_spec(T.Biped, T.C0, 2.4)
        set_terrain_table(s, *bad)
        with pytest.raises(T.TowrB200Error) as e:
            T.Structure(s)
        assert e.value.code == -1


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


def test_ee_linear_constraint_rules(product):
    """EELinearConstraint definitions (parameters.h:315-327) are part of the class; malformed ones are argument errors."""
    from util import set_ee_linear
    T = product
    base = [c for c in CASES if c[0] == "biped_walk_ee_linear"][0]
    ok = mk(T, base)
    st = T.Structure(ok)
    assert [n for n, _ in st.con_sets()][-2:] == ["ee-linear", "ee-linear"]      # every definition's set has this name (:14)
    assert not T.Structure(mk(T, [c for c in CASES if c[0] == "biped_walk_stairs"][0])).accepts(ok)
    for bad in ([(0, 0, [(2, 0, 1.0)], 0.1, 0.1)], [(0, 0, [(0, 3, 1.0)], 0.1, 0.1)], [(0, 0, [(0, 0, 1.0)], 0.1, 0.0)],
                [(0, 2, [(0, 0, 1.0)], 0.1, 0.1)], [(2, 0, [(0, 0, 1.0)], 0.1, 0.1)]):
        s = mk(T, base)
        set_ee_linear(s, bad)
        with pytest.raises(T.TowrB200Error):
            T.Structure(s)
