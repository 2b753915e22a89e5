"""The C++ oracle (oracle/*.cc) against the second, independent restatement (oracle/oracle_np.py: numpy + scipy.sparse,
written from the reference's Eigen expressions).  The reference ships no golden vectors (its tests are empty stubs,
SURVEY.md F5), so agreement of two separately derived restatements — values within 1e-12 / 1e-14, stored-entry pattern
identical, also where the pattern depends on the values (Gap terrain) — is what pins the oracle the GPU parity tests use.

CPU only; no product code is involved."""
import numpy as np
import pytest

from oracle import oracle_np as N
from util import CASES, GAP_CASES, mk, perturbed, tol_ratio


def _clip(a):
    return np.clip(a, -1e20, 1e20)     # ifopt::inf = 1e20; the C++ oracle reports +-infinity where the reference uses it


@pytest.mark.parametrize("case", CASES + GAP_CASES, ids=lambda c: c[0])
def test_cpp_oracle_matches_numpy_restatement(oracle, case):
    O = oracle
    s = mk(O, case)
    p = O.Problem(s)
    f = N.Formulation(s)
    # structure: variable sets (column order), constraint sets (row order)
    assert [(v.name, v.rows()) for v in f.var_sets] == p.var_sets()
    assert [(c.name, c.n_rows) for c in f.con_sets] == p.con_sets()
    # initial guess and bounds (NlpFormulation::Make*Variables, GetBounds of every set)
    x0 = p.x0()
    assert tol_ratio(f.x0(), x0) <= 1.0
    for a, b in zip(f.var_bounds(), p.var_bounds()):
        assert tol_ratio(_clip(a), _clip(b)) <= 1.0
    for a, b in zip(f.con_bounds(), p.con_bounds()):
        assert tol_ratio(_clip(a), _clip(b)) <= 1.0
    assert tol_ratio(f.eval_g(x0), p.eval_g(x0)) <= 1.0
    # a Gap class changes its stored entries with x: check two points there, one elsewhere
    two_points = case[0] in ("hyq_gallop_gap", "monoped_hop_gap_torque_durations")      # the others: one x, to keep the suite short
    points = [perturbed(x0, 1, 0.05)] + ([x0] if two_points else [])
    if case[0] == "anymal_trot_very_fine_dt":
        points = []        # the same code as anymal_trot on 4 x the samples: values only (above), to keep the CPU suite short
    nnz = []
    for x in points:
        assert tol_ratio(f.eval_g(x), p.eval_g(x)) <= 1.0
        r, c, v = f.eval_jac(x)
        ro, co, vo = p.eval_jac(x)
        assert len(r) == len(ro) and np.array_equal(r, ro) and np.array_equal(c, co), "stored-entry pattern differs"
        assert tol_ratio(v, vo) <= 1.0
        nnz.append(len(r))
    if two_points:
        assert nnz[0] != nnz[1], "the Gap case should exercise a value-dependent pattern"


@pytest.mark.parametrize("robot,optimise,costs", [
    (0, True, ((0, 1e-3), (1, 1e-2))),                       # hopper_example.cc:147-148 (ForcesCostID, EEMotionCostID)
    (0, True, ((2, 1e-3), (3, 0.7))),                        # EnergyCostID, AngMomCostID with the ee-schedule gradient
    (1, False, ((2, 1e-3), (3, 0.7), (0, 1e-3))),
    (3, True, ((2, 1e-3), (3, 0.7))),
    (0, True, ("swing",)),                                   # EEBasePosCost alone, then with a listed cost
    (3, False, ((2, 1e-3), "swing")),
], ids=["monoped_nodecost", "monoped_sampled_durations", "biped_sampled", "anymal_sampled_durations", "monoped_swing_durations",
        "anymal_swing"])
def test_costs_match_numpy_restatement(oracle, robot, optimise, costs):
    """eval_f / eval_grad_f: NodeCost (node_cost.cc), EnergyCost (energy_cost.cc), AngularMomentumCost
    (angular_momentum_cost.cc), EEBasePosCost (ee_base_pos_cost.cc)."""
    O = oracle
    s = O.make_spec(robot, 0, 2.4, optimize_durations=optimise, extra_constraints=(O.Torque,))
    if "swing" in costs:
        s.enable_swing_ee_base_pos_tracking = 1
        s.swing_ee_base_pos_tracking_weight = 0.3
        s.initial_base_ang[1], s.initial_base_ang[2] = -0.1, 0.2
        costs = tuple(c for c in costs if c != "swing")
    s.n_costs = len(costs)
    for i, (cid, w) in enumerate(costs):
        s.cost_ids[i], s.cost_weights[i] = cid, w
    p, f = O.Problem(s), N.Formulation(s)
    x = perturbed(p.x0(), 3, 0.05)
    assert tol_ratio(np.array([f.eval_f(x)]), np.array([p.eval_f(x)])) <= 1.0
    assert tol_ratio(f.eval_grad_f(x), p.eval_grad_f(x)) <= 1.0
    assert np.abs(p.eval_grad_f(x)).max() > 0


def test_stance_tracking_bounds_match(oracle):
    """nlp_formulation.cc:218-231: the x, y of every stance are pinned to ee_stance_position_ (hopper_example.cc:114-122)."""
    O = oracle
    s = O.make_spec(O.Monoped, 0, 2.4)
    s.enable_stance_tracking = 1
    n_stances = (int(s.n_phases[0]) + (1 if s.in_contact_at_start[0] else 0)) // 2
    s.n_stance_positions[0] = n_stances
    for i in range(n_stances):
        s.ee_stance_position[0][i][0] = 0.3 * i
        s.ee_stance_position[0][i][1] = -0.01 * i
    p, f = O.Problem(s), N.Formulation(s)
    for a, b in zip(f.var_bounds(), p.var_bounds()):
        assert np.array_equal(_clip(a), _clip(b))
    s.n_stance_positions[0] = n_stances - 1            # one too few: std::out_of_range in the reference
    with pytest.raises(IndexError):
        N.Formulation(s)
    with pytest.raises(RuntimeError):
        O.Problem(s)


def test_building_blocks_match(oracle):
    """Hermite basis, d pos / d T, terrain bases and sample times: the unit hooks of the C++ oracle against oracle_np."""
    import ctypes as C
    O = oracle
    L = O.lib()
    dp = lambda a: a.ctypes.data_as(C.POINTER(C.c_double))
    rng = np.random.default_rng(5)
    for _ in range(20):
        T = float(rng.uniform(0.05, 1.2))
        t = float(rng.uniform(0.0, T))
        out = np.zeros(12)
        L.orc_hermite_basis(T, t, dp(out))
        poly = N.CubicHermite(3)
        poly.T = T
        mine = [fn(dxdt, nd, t) for dxdt in (0, 1, 2) for fn in (poly.d_wrt_start, poly.d_wrt_end) for nd in (0, 1)]
        assert tol_ratio(np.array(mine), out) <= 1.0
        nodes = rng.uniform(-1, 1, size=(4, 3))
        poly.n0 = nodes[:2].copy()
        poly.n1 = nodes[2:].copy()
        poly.update_coeff()
        pos, vel, acc, dT = np.zeros(3), np.zeros(3), np.zeros(3), np.zeros(3)
        p0, v0, p1, v1 = (np.ascontiguousarray(nodes[i]) for i in range(4))
        L.orc_hermite_point(dp(p0), dp(v0), dp(p1), dp(v1), T, t, dp(pos), dp(vel), dp(acc))
        L.orc_hermite_dpos_dT(dp(p0), dp(v0), dp(p1), dp(v1), T, t, dp(dT))
        pt = poly.point(t)
        assert tol_ratio(pt[0], pos) <= 1.0 and tol_ratio(pt[1], vel) <= 1.0 and tol_ratio(pt[2], acc) <= 1.0
        assert tol_ratio(poly.dpos_dT(t), dT) <= 1.0
    for tid in range(8):
        terrain = N.TERRAINS[tid]()
        for _ in range(40):
            x, y = float(rng.uniform(-0.5, 3.5)), float(rng.uniform(-1.0, 1.0))
            h, dh, basis, dbasis = np.zeros(1), np.zeros(2), np.zeros(9), np.zeros(18)
            L.orc_terrain(tid, x, y, dp(h), dp(dh), dp(basis), dp(dbasis))
            assert h[0] == terrain.height(x, y)
            assert dh[0] == terrain.d_height(0, x, y) and dh[1] == terrain.d_height(1, x, y)
            mine = np.concatenate([terrain.normalized_basis(b, x, y) for b in range(3)])
            assert tol_ratio(mine, basis) <= 1.0
            mine = np.concatenate([terrain.d_normalized_basis(b, d, x, y) for b in range(3) for d in range(2)])
            assert tol_ratio(mine, dbasis) <= 1.0
    for T, dt in ((2.0, 0.1), (2.4, 0.08), (2.0, 0.02), (1.9, 0.1), (3.0, 0.005)):
        out, n = np.zeros(2000), C.c_int(0)
        L.orc_sample_times(T, dt, dp(out), C.byref(n))
        assert np.array_equal(out[:n.value], np.array(N.sample_times(T, dt)))
