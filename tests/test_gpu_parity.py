"""GPU parity (-m gpu): the CUDA path, called through the C ABI, against the oracle on the same inputs.
Tolerance (BASELINE.json north_star): sparsity indices bit-exact, values within 1e-12 relative / 1e-14 absolute."""
import numpy as np
import pytest

from util import CASES, GAP_CASES, check_against_oracle, is_rt, mk, perturbed, tol_ratio

pytestmark = pytest.mark.gpu


def _problem_batch(T, O, case, n, seed):
    """n problems of one structure class with per-problem goals, terrain kept, and distinct x."""
    goal = case[4]
    rng = np.random.default_rng(seed)
    specs, xs = [], []
    s0 = mk(T, case)
    st = T.Structure(s0)
    for i in range(n):
        g = (goal[0] + rng.uniform(-0.3, 0.3), goal[1] + rng.uniform(-0.2, 0.2), goal[2] + rng.uniform(-0.3, 0.3))
        s = mk(T, case, goal=g)
        s.mass *= 1.0 + 0.05 * rng.uniform(-1, 1)
        specs.append(s)
        xs.append(perturbed(st.x0(s), seed * 1000 + i, (0.02 if is_rt(case) else 0.1) if i % 2 else 0.0))
    return st, specs, np.stack(xs)


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_values_and_jacobian_match_oracle(product, oracle, case):
    T, O = product, oracle
    from util import to_oracle_spec
    n = 6
    st, specs, x = _problem_batch(T, O, case, n, 3)
    b = T.Batch(st, n)
    for i, s in enumerate(specs):
        b.set_problem(i, s)
    g, v = b.eval(x)
    assert (b.status() == 0).all()
    for i, s in enumerate(specs):
        check_against_oracle(O, to_oracle_spec(O, s), st, x[i], g[i], v[i], is_rt(case))


@pytest.mark.parametrize("case", GAP_CASES, ids=[c[0] for c in GAP_CASES])
def test_gap_superset_pattern_matches_oracle(product, oracle, case):
    """Gap (curvature): the class's pattern is a superset of what the oracle stores at any x; values equal on the oracle's
    entries, exact zeros on the others (force_constraint_discretized.cc:52-67, torque_constraint_discretized.cc:176-197)."""
    T, O = product, oracle
    from util import to_oracle_spec
    n = 6
    st, specs, x = _problem_batch(T, O, case, n, 7)
    b = T.Batch(st, n)
    for i, s in enumerate(specs):
        b.set_problem(i, s)
    g, v = b.eval(x)
    assert (b.status() == 0).all()
    extras = set()
    for i, s in enumerate(specs):
        extras.add(check_against_oracle(O, to_oracle_spec(O, s), st, x[i], g[i], v[i], is_rt(case), superset=True))
    assert len(extras) > 1


@pytest.mark.parametrize("which", ["anymal_dt0.02", "biped_dt0.01_durations"])
def test_multi_part_classes_match_oracle(product, oracle, which, monkeypatch):
    """Classes split into parts (pack.cc; BASELINE config 5's fine sampling steps): several launches per evaluation, values
    against the oracle; and a small class forced into parts is bit-identical to its single-part evaluation."""
    from test_emu import fine_dt_spec
    from util import to_oracle_spec
    T, O = product, oracle
    if which == "anymal_dt0.02":
        s, rt = fine_dt_spec(T, T.Anymal, T.C1, 2.0, 0.02), False
    else:
        s, rt = fine_dt_spec(T, T.Biped, T.C0, 2.4, 0.01, optimise=True), True
    st = T.Structure(s)
    n = 5
    x = np.stack([perturbed(st.x0(s), 60 + i, 0.02 if rt else 0.08) for i in range(n)])
    b = T.Batch(st, n)
    b.set_all(s)
    assert "parts" in b.kernel_info
    g, v = b.eval(x)
    assert (b.status() == 0).all()
    for i in (0, n - 1):
        check_against_oracle(O, to_oracle_spec(O, s), st, x[i], g[i], v[i], rt)
    small = T.make_spec(T.Anymal, T.C1, 2.0, optimize_durations=rt)
    st1 = T.Structure(small)
    xs = np.stack([perturbed(st1.x0(small), 80 + i, 0.02) for i in range(7)])
    b1 = T.Batch(st1, 7)
    b1.set_all(small)
    ref = b1.eval(xs)
    monkeypatch.setenv("TOWR_B200_MAX_BLOB", "24000")
    st2 = T.Structure(small)
    b2 = T.Batch(st2, 7)
    b2.set_all(small)
    assert "parts" in b2.kernel_info
    alt = b2.eval(xs)
    assert np.array_equal(ref[0], alt[0]) and np.array_equal(ref[1], alt[1])


def test_table_terrain_reproduces_the_five_step_stairs(product):
    """The hopper example's own HeightMap (FiveStepStairs, hopper_example.cc:53-91) written as a table of five steps gives
    the same g and Jacobian values as the built-in terrain id 7, and per-problem tables differ within one batch."""
    from util import set_terrain_table
    T = product
    a = T.make_spec(T.Monoped, T.C0, 2.4, goal=(1.0, 0.0, 0.0), terrain=T.FiveStepStairsID, extra_constraints=(T.Torque,))
    b = T.make_spec(T.Monoped, T.C0, 2.4, goal=(1.0, 0.0, 0.0), extra_constraints=(T.Torque,))
    set_terrain_table(b, T.TABLE_STEPS, [0.5, 0.8, 1.1, 1.4, 1.7], [0.15, 0.3, 0.45, 0.6, 0.75])
    c = T.make_spec(T.Monoped, T.C0, 2.4, goal=(1.0, 0.0, 0.0), extra_constraints=(T.Torque,))
    set_terrain_table(c, T.TABLE_STEPS, [0.5, 0.8], [0.3, 0.6])
    st = T.Structure(b)
    x0 = st.x0(a)
    rng = np.random.default_rng(4)
    x = x0 + rng.uniform(-0.013, 0.013, size=x0.shape)         # no node sits exactly on a step edge (there int((x - 0.5) / 0.3) and x >= knot may round differently)
    batch = T.Batch(st, 3)
    for i, s in enumerate((a, b, c)):
        batch.set_problem(i, s)
    g, v = batch.eval(np.stack([x, x, x]))
    assert (batch.status() == 0).all()
    assert np.array_equal(g[0], g[1]) and np.array_equal(v[0], v[1])
    assert not np.array_equal(g[1], g[2])


@pytest.mark.parametrize("optimise", [False, True], ids=["fixed", "durations"])
def test_mixed_terrain_batch_block_gap_stairs(product, oracle, optimise):
    """BASELINE config 4 as written: HyQ gallop, terrain ids cycling Block / Gap / Stairs inside ONE batch.  The class is
    created on Gap (declares the curvature block); Block and Stairs problems write exact zeros there."""
    T, O = product, oracle
    from util import to_oracle_spec
    terrains = (T.BlockID, T.GapID, T.StairsID)
    rng = np.random.default_rng(11)
    st = T.Structure(T.make_spec(T.Hyq, T.C4, 2.4, terrain=T.GapID, optimize_durations=optimise))
    n = 9
    specs, xs = [], []
    for i in range(n):
        s = T.make_spec(T.Hyq, T.C4, 2.4, goal=(rng.uniform(1.0, 2.0), rng.uniform(-0.2, 0.2), rng.uniform(-0.3, 0.3)),
                        terrain=terrains[i % 3], optimize_durations=optimise)
        specs.append(s)
        xs.append(perturbed(st.x0(s), 300 + i, (0.02 if optimise else 0.08) if i >= 3 else 0.0))
    x = np.stack(xs)
    b = T.Batch(st, n)
    for i, s in enumerate(specs):
        b.set_problem(i, s)
    g, v = b.eval(x)
    assert (b.status() == 0).all()
    n_extra = []
    for i, s in enumerate(specs):
        n_extra.append(check_against_oracle(O, to_oracle_spec(O, s), st, x[i], g[i], v[i], optimise, superset=True))
    # Block / Stairs: the whole declared curvature block is extra (never stored by the reference); Gap: part of it
    assert n_extra[0] == n_extra[2] and n_extra[1] < n_extra[0]


def test_eval_modes_and_ids(product, oracle):
    """EVAL_G / EVAL_JAC alone give the same numbers as EVAL_BOTH; problem_ids selects per-problem constants."""
    T = product
    st, specs, x = _problem_batch(T, oracle, CASES[0], 8, 5)
    b = T.Batch(st, 8)
    for i, s in enumerate(specs):
        b.set_problem(i, s)
    g, v = b.eval(x)
    g1, _ = b.eval(x, T.EVAL_G)
    _, v2 = b.eval(x, T.EVAL_JAC)
    assert np.array_equal(g, g1) and np.array_equal(v, v2)
    ids = np.array([5, 2, 7], dtype=np.int32)
    gs, vs = b.eval(x[ids], ids=ids)
    assert np.array_equal(gs, g[ids]) and np.array_equal(vs, v[ids])
    assert b.launches == 4


@pytest.mark.parametrize("name", ["anymal_trot", "anymal_trot_durations", "monoped_hop_torque", "anymal_walk_slope",
                                  "anymal_trot_force_nodes_slope", "biped_walk_stairs"])
def test_dynamic_only_mode_equals_full_mode(product, name):
    """TOWR_B200_EVAL_JAC_DYNAMIC: constants once (towr_b200_batch_constant_values) + the dynamic entries of every eval,
    written in place, are bit-identical to the full n x nnz result; chunked evals equal unchunked ones."""
    import os
    T = product
    case = [c for c in CASES if c[0] == name][0]
    n = 37
    st, specs, x = _problem_batch(T, None, case, n, 9)
    b = T.Batch(st, n)
    for i, s in enumerate(specs):
        b.set_problem(i, s)
    g, v = b.eval(x)
    idx = b.dynamic_index()
    assert 0 < b.n_dynamic < st.nnz and (np.diff(idx) > 0).all()
    if name == "anymal_trot":                     # flat ground: the force constraint's basis entries are constants too
        assert b.n_dynamic < 0.55 * st.nnz
    const_mask = np.ones(st.nnz, bool)
    const_mask[idx] = False
    full = np.stack([b.constant_values(i) for i in range(n)])
    assert np.array_equal(full[:, const_mask], v[:, const_mask])      # the constants are what the kernel stores
    assert not full[:, ~const_mask].any()
    g2, _ = b.eval(x, T.EVAL_G | T.EVAL_JAC_DYNAMIC, jac_out=full)      # rewrites the dynamic positions in place
    assert np.array_equal(g2, g) and np.array_equal(full, v)
    poison = np.full((n, st.nnz), np.nan)                               # constants: untouched, or rewritten with themselves
    b.eval(x, T.EVAL_JAC_DYNAMIC, jac_out=poison)
    assert np.array_equal(poison[:, ~const_mask], v[:, ~const_mask])
    touched = ~np.isnan(poison)
    assert np.array_equal(poison[touched], v[touched])
    assert int(touched[0].sum()) == b.dynamic_transfer_entries
    os.environ["TOWR_B200_CHUNK"] = "8"                                 # 5 chunks, alternating streams, ragged last one
    try:
        g3, v3 = b.eval(x)
        full3 = np.stack([b.constant_values(i) for i in range(n)])
        b.eval(x, T.EVAL_JAC_DYNAMIC, jac_out=full3)
    finally:
        del os.environ["TOWR_B200_CHUNK"]
    assert np.array_equal(g3, g) and np.array_equal(v3, v) and np.array_equal(full3, v)


def test_full_batch_properties(product, oracle):
    """BASELINE config 3 at full size (ANYmal fly-trot, batch 4096): properties that need no oracle at scale,
    plus a spot check of 8 problems against the oracle."""
    T, O = product, oracle
    n = 4096
    s = T.make_spec(T.Anymal, T.C1, 2.0)
    st = T.Structure(s)
    b = T.Batch(st, n)
    b.set_all(s)
    x0 = st.x0(s)
    rng = np.random.default_rng(20260103)
    x = x0[None, :] + rng.uniform(-1, 1, size=(n, st.n_vars)) * 0.05
    x[n // 2:] = x[: n // 2]                       # second half repeats the first: outputs must repeat bit-exactly
    g, v = b.eval(x)
    assert np.isfinite(g).all() and np.isfinite(v).all()
    assert np.array_equal(g[: n // 2], g[n // 2:]) and np.array_equal(v[: n // 2], v[n // 2:])
    # determinism
    g2, v2 = b.eval(x)
    assert np.array_equal(g, g2) and np.array_equal(v, v2)
    # the Jacobian is the derivative of g: directional finite difference through the GPU path itself
    rows, cols = st.pattern()
    d = rng.uniform(-1, 1, size=st.n_vars)
    h = 1e-6
    gp, _ = b.eval(x[:4] + h * d, T.EVAL_G)
    gm, _ = b.eval(x[:4] - h * d, T.EVAL_G)
    for i in range(4):
        jd = np.zeros(st.n_rows)
        np.add.at(jd, rows, v[i] * d[cols])
        fd = (gp[i] - gm[i]) / (2 * h)
        assert np.abs(jd - fd).max() <= 1e-5 * max(1.0, np.abs(fd).max())
    p = O.Problem(O.make_spec(O.Anymal, 1, 2.0))
    for i in (0, 1, 777, 2047, 2048, 4095, 1234, 3000):
        assert tol_ratio(g[i], p.eval_g(x[i])) <= 1.0
        assert tol_ratio(v[i], p.eval_jac(x[i])[2]) <= 1.0


def test_device_pointer_path(product):
    """eval_device on caller-owned device memory (torch tensors) equals the host-buffer path."""
    import torch
    T = product
    s = T.make_spec(T.Anymal, T.C1, 2.0)
    st = T.Structure(s)
    n = 64
    b = T.Batch(st, n)
    b.set_all(s)
    x = st.x0(s)[None, :] + np.random.default_rng(1).uniform(-1, 1, size=(n, st.n_vars)) * 0.05
    g, v = b.eval(x)
    dx = torch.from_numpy(x).cuda()
    dg = torch.empty((n, st.n_rows), dtype=torch.float64, device="cuda")
    dv = torch.empty((n, st.nnz), dtype=torch.float64, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    b.eval_device(n, T.EVAL_BOTH, dx.data_ptr(), dg.data_ptr(), dv.data_ptr(), stream)
    torch.cuda.synchronize()
    assert np.array_equal(dg.cpu().numpy(), g) and np.array_equal(dv.cpu().numpy(), v)


def test_non_finite_input_flags_only_its_problem(product):
    T = product
    s = T.make_spec(T.Anymal, T.C1, 2.0)
    st = T.Structure(s)
    b = T.Batch(st, 4)
    b.set_all(s)
    x = np.tile(st.x0(s), (4, 1))
    x[2, 10] = np.nan
    g, v = b.eval(x)
    status = b.status()
    assert status[2] != 0 and (status[[0, 1, 3]] == 0).all()
    assert np.isfinite(v[[0, 1, 3]]).all()


def test_gpu_reproduces_committed_fixtures(product):
    """tests/golden/*.npz through the C ABI: the GPU box has neither /root/reference nor needs the oracle here."""
    import glob
    import os
    T = product
    by_name = {c[0]: c for c in CASES}
    files = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))
    assert files
    for f in files:
        z = np.load(f)
        s = mk(T, by_name[os.path.basename(f)[:-4]])
        st = T.Structure(s)
        b = T.Batch(st, 3)
        b.set_all(s)
        g, v = b.eval(np.tile(z["x"], (3, 1)))
        rows, cols = st.pattern()
        assert np.array_equal(rows[z["jac_idx"]], z["jac_rows"]) and np.array_equal(cols[z["jac_idx"]], z["jac_cols"])
        for i in range(3):
            assert tol_ratio(g[i], z["g"]) <= 1.0 and tol_ratio(v[i][z["jac_idx"]], z["jac_vals"]) <= 1.0


@pytest.mark.parametrize("name", ["anymal_trot", "anymal_trot_durations"])
def test_ragged_and_empty_batches(product, name):
    """n = 0 is a no-op; odd n (a last problem group with one lane only) and n = 1 give the rows of the full eval;
    the same for the cost and trajectory entry points."""
    T = product
    case = [c for c in CASES if c[0] == name][0]
    s = mk(T, case)
    st = T.Structure(s)
    nb = 7
    x = np.stack([perturbed(st.x0(s), 900 + i, 0.02) for i in range(nb)])
    b = T.Batch(st, nb)
    b.set_all(s)
    g, v = b.eval(x)
    rows = b.sample_trajectory(x, 0.05)
    f, gr = b.eval_cost(x)
    for n in (0, 1, 3, 7):
        gn, vn = b.eval(x[:n])
        assert gn.shape == (n, st.n_rows) and vn.shape == (n, st.nnz)
        assert np.array_equal(gn, g[:n]) and np.array_equal(vn, v[:n])
        assert np.array_equal(b.sample_trajectory(x[:n], 0.05), rows[:n])
        fn, grn = b.eval_cost(x[:n])
        assert np.array_equal(fn, f[:n]) and np.array_equal(grn, gr[:n])
    with pytest.raises(T.TowrB200Error):
        b.eval(np.tile(x, (2, 1)))              # more problems than the batch holds
