"""eval_f / eval_grad_f: the cost terms of NlpFormulation::GetCosts (nlp_formulation.cc:605-681): NodeCost
(costs/node_cost.cc) as test/hopper_example.cc:147-148 enables it, and the time-discretised EnergyCost
(costs/energy_cost.cc) and AngularMomentumCost (costs/angular_momentum_cost.cc).  Oracle known answers and the kernel
arithmetic replayed on the host (tests/emu) on the CPU, CUDA path against the oracle on the GPU."""
import ctypes as C

import numpy as np
import pytest

from test_hopper_example import hopper_spec
from util import CASES, check_cost_against_oracle, cost_gradient_ok, is_rt, mk, perturbed, to_oracle_spec, tol_ratio

FORCES, EE_MOTION, ENERGY, ANG_MOM = 0, 1, 2, 3


def with_costs(s, costs):
    s.n_costs = len(costs)
    for i, (cid, w) in enumerate(costs):
        s.cost_ids[i] = cid
        s.cost_weights[i] = w
    return s


def test_oracle_cost_known_answers(product, oracle):
    """f is the weighted sum of squared node values; the gradient matches central differences; no cost = 0."""
    T, O = product, oracle
    s = with_costs(mk(T, CASES[4]), [(FORCES, 1e-9), (EE_MOTION, 1e-4)])      # hopper_example.cc:147-148 weights
    st = T.Structure(s)
    p = O.Problem(to_oracle_spec(O, s))
    x = perturbed(st.x0(s), 2, 0.1)
    f = p.eval_f(x)
    g = p.eval_grad_f(x)
    assert f > 0.0 and g.shape == (st.n_vars,)
    names = dict((n, i) for i, (n, _) in enumerate(st.var_sets()))
    sizes = [r for _, r in st.var_sets()]
    off = np.concatenate([[0], np.cumsum(sizes)])
    touched = np.zeros(st.n_vars, bool)
    for n in ("ee-motion_0", "ee-force_0", "ee-torque_0"):
        touched[off[names[n]]:off[names[n] + 1]] = True
    assert (g[~touched] == 0.0).all() and (g[touched] != 0.0).any()
    rng = np.random.default_rng(0)
    for c in rng.choice(np.nonzero(g)[0], 6, replace=False):
        h = 1e-4 * max(1.0, abs(x[c]))
        xp, xm = x.copy(), x.copy()
        xp[c] += h
        xm[c] -= h
        fd = (p.eval_f(xp) - p.eval_f(xm)) / (2 * h)          # f is quadratic: central differences are exact up to rounding
        assert abs(fd - g[c]) <= 1e-6 * abs(g[c]) + 1e-12
    p0 = O.Problem(to_oracle_spec(O, mk(T, CASES[4])))
    assert p0.eval_f(x) == 0.0 and (p0.eval_grad_f(x) == 0.0).all()


def test_cost_spec_errors(product):
    T = product
    with pytest.raises(T.TowrB200Error):
        T.Structure(with_costs(mk(T, CASES[4]), [(7, 1.0)]))   # "cost not defined!" (nlp_formulation.cc:644)
    a = T.Structure(with_costs(mk(T, CASES[4]), [(FORCES, 1.0)]))
    assert not a.accepts(mk(T, CASES[4]))                      # the cost list is part of the structure class
    e = T.Structure(with_costs(mk(T, CASES[4]), [(ENERGY, 1.0)]))
    other = with_costs(mk(T, CASES[4]), [(ENERGY, 1.0)])
    assert e.accepts(other)
    other.dt_cost_energy = 0.05                                # ... and so are the sampling period and the torque weight
    assert not e.accepts(other)
    s = with_costs(mk(T, CASES[4]), [(ENERGY, 1.0)])
    s.dt_cost_energy = float("nan")
    with pytest.raises(T.TowrB200Error):
        T.Structure(s)
    s = mk(T, CASES[4])
    assert s.dt_cost_energy == 0.02 and s.energy_cost_torque_weight == 1.0 and s.dt_cost_ang_mom == 0.02   # parameters.cc:69-71
    assert (s.enable_swing_ee_base_pos_tracking, s.swing_ee_base_pos_tracking_weight, s.dt_cost_swing_ee_base_pos_tracking) == \
        (0, 1e-2, 0.05)                                        # parameters.h:233-235
    plain = T.Structure(mk(T, CASES[4]))
    swing = sampled_spec(T, "monoped_hop", [], None, None, None, (0.3, None))
    assert not plain.accepts(swing) and T.Structure(swing).accepts(swing)
    swing.swing_ee_base_pos_tracking_weight = 0.0              # weight 0: no cost term is made (nlp_formulation.cc:614)
    assert plain.accepts(swing)


# (case, cost list, dt_cost_energy, energy_cost_torque_weight, dt_cost_ang_mom)
SAMPLED = [
    ("monoped_hop", [(ENERGY, 1e-3)], None, None, None),
    ("monoped_hop", [(ANG_MOM, 0.5)], None, None, None),
    ("anymal_trot", [(FORCES, 1e-9), (EE_MOTION, 1e-4), (ENERGY, 1e-3), (ANG_MOM, 0.7)], None, None, None),
    ("biped_walk_durations", [(ENERGY, 1e-3), (ANG_MOM, 0.7)], None, None, None),
    ("anymal_trot_durations", [(ENERGY, 1e-3), (ANG_MOM, 0.7), (FORCES, 1e-5)], 0.03, 0.0, 0.05),   # no torque term (:119)
    ("monoped_hop_torque", [(ENERGY, 2.0), (ANG_MOM, 0.0)], 0.0, 0.5, None),   # dt <= 0: t = 0 and T only (:55-59); weight 0: no term
    # EEBasePosCost (params_.enable_swing_ee_base_pos_tracking, nlp_formulation.cc:612-626): 6th element = (weight, dt)
    ("monoped_hop", [], None, None, None, (0.3, None)),
    ("anymal_trot", [(ENERGY, 1e-3), (ANG_MOM, 0.7)], None, None, None, (0.3, None)),
    ("biped_walk_durations", [(FORCES, 1e-5)], None, None, None, (1e-2, 0.03)),
    ("anymal_trot_durations", [], None, None, None, (0.3, 0.03)),
]


def sampled_spec(T, name, costs, dt_energy, torque_weight, dt_ang_mom, swing=None):
    s = with_costs(mk(T, [c for c in CASES if c[0] == name][0]), costs)
    if swing is not None:
        s.enable_swing_ee_base_pos_tracking = 1
        s.swing_ee_base_pos_tracking_weight = swing[0]
        if swing[1] is not None:
            s.dt_cost_swing_ee_base_pos_tracking = swing[1]
        s.initial_base_ang[1], s.initial_base_ang[2] = -0.1, 0.2      # a reference frame that is not the world frame
    if dt_energy is not None:
        s.dt_cost_energy = dt_energy
    if torque_weight is not None:
        s.energy_cost_torque_weight = torque_weight
    if dt_ang_mom is not None:
        s.dt_cost_ang_mom = dt_ang_mom
    return s


def with_node_costs(s, terms):
    """Individual NodeCost(nodes_id, deriv, dim, weight) terms: (set, foot, deriv, dim, weight)."""
    s.n_node_costs = len(terms)
    for i, (k, ee, deriv, dim, w) in enumerate(terms):
        s.node_cost_set[i], s.node_cost_ee[i], s.node_cost_deriv[i], s.node_cost_dim[i], s.node_cost_weight[i] = k, ee, deriv, dim, w
    return s


def squat_jump_costs(s, w=1e-3):
    """test/squat_jump_example.cc:205-211: NodeCost on the force and the torque node positions, every dimension."""
    EE_FORCE, EE_TORQUE = 4, 5
    return with_node_costs(s, [(k, 0, 0, d, w) for d in range(3) for k in (EE_FORCE, EE_TORQUE)])


def test_individual_node_costs(product, oracle, emu):
    """NodeCost terms named one by one (the reference's hand-built examples) in the oracle and in the kernel arithmetic;
    a cost family equals its expansion into individual terms (nlp_formulation.cc:648-681)."""
    T, O = product, oracle
    DP = C.POINTER(C.c_double)
    case = [c for c in CASES if c[0] == "monoped_squat_jump_durations"][0]
    s = squat_jump_costs(mk(T, case))
    st = T.Structure(s)
    assert not T.Structure(mk(T, case)).accepts(s)              # the terms are part of the class
    p = O.Problem(to_oracle_spec(O, s))
    x = perturbed(st.x0(s), 4, 0.05)
    f, g = np.zeros(1), np.zeros(st.n_vars)
    assert emu.emu_eval_cost(C.byref(s), x.ctypes.data_as(DP), f.ctypes.data_as(DP), g.ctypes.data_as(DP)) == 0
    assert f[0] > 0 and tol_ratio(f, np.array([p.eval_f(x)])) <= 1.0 and tol_ratio(g, p.eval_grad_f(x)) <= 1.0
    # ForcesCostID(w) = force / torque positions at w and velocities at 0.1 w; EEMotionCostID(w) = velocities at w, w, 0.5 w
    w = 1e-3
    fam = O.Problem(to_oracle_spec(O, with_costs(mk(T, case), [(FORCES, w), (EE_MOTION, w)])))
    terms = [(k, 0, deriv, d, w if deriv == 0 else 0.1 * w) for d in range(3) for deriv in (0, 1) for k in (4, 5)]
    terms += [(2, 0, 1, 0, w), (2, 0, 1, 1, w), (2, 0, 1, 2, 0.5 * w)]
    one = O.Problem(to_oracle_spec(O, with_node_costs(mk(T, case), terms)))
    assert tol_ratio(np.array([one.eval_f(x)]), np.array([fam.eval_f(x)])) <= 1.0
    assert tol_ratio(one.eval_grad_f(x), fam.eval_grad_f(x)) <= 1.0
    bad = with_node_costs(mk(T, case), [(6, 0, 0, 0, 1.0)])
    with pytest.raises(T.TowrB200Error):
        T.Structure(bad)
    bad = with_node_costs(mk(T, case), [(4, 1, 0, 0, 1.0)])    # the monoped has one foot
    with pytest.raises(T.TowrB200Error):
        T.Structure(bad)


def test_oracle_sampled_cost_gradient_is_consistent(product, oracle):
    """EnergyCost / AngularMomentumCost: the gradient (incl. the ee-schedule columns) matches central differences of f."""
    T, O = product, oracle
    s = sampled_spec(T, "biped_walk_durations", [(ENERGY, 1e-3), (ANG_MOM, 0.7)], None, None, None)
    st = T.Structure(s)
    p = O.Problem(to_oracle_spec(O, s))
    x = perturbed(st.x0(s), 2, 0.05)
    g = p.eval_grad_f(x)
    off = np.concatenate([[0], np.cumsum([r for _, r in st.var_sets()])])
    names = [n for n, _ in st.var_sets()]
    for n in ("base-ang", "ee-force_0", "ee-torque_1", "ee-schedule0"):
        i = names.index(n)
        assert np.abs(g[off[i]:off[i + 1]]).max() > 0, n
    for n in ("base-lin", "ee-motion_0"):
        i = names.index(n)
        assert (g[off[i]:off[i + 1]] == 0.0).all(), n
    rng = np.random.default_rng(0)
    for c in rng.choice(np.nonzero(g)[0], 12, replace=False):
        h = 1e-6
        xp, xm = x.copy(), x.copy()
        xp[c] += h
        xm[c] -= h
        fd = (p.eval_f(xp) - p.eval_f(xm)) / (2 * h)
        assert abs(fd - g[c]) <= 1e-6 * max(1.0, abs(g[c]))


@pytest.mark.parametrize("case", SAMPLED, ids=lambda c: c[0] + "".join("_%d" % cid for cid, _ in c[1]) + ("_swing" if len(c) > 5 else ""))
def test_emulated_cost_kernels_match_oracle(product, oracle, emu, case):
    """The device code of both cost kernels (NodeCost table, sampled_cost.cuh) replayed serially on the host."""
    T, O = product, oracle
    s = sampled_spec(T, *case)
    st = T.Structure(s)
    p = O.Problem(to_oracle_spec(O, s))
    DP = C.POINTER(C.c_double)
    for seed in (0, 5):
        x = perturbed(st.x0(s), 70 + seed, 0.05 if seed else 0.0)
        f, g = np.zeros(1), np.zeros(st.n_vars)
        rc = emu.emu_eval_cost(C.byref(s), x.ctypes.data_as(DP), f.ctypes.data_as(DP), g.ctypes.data_as(DP))
        assert rc == 0, emu.emu_last_error()
        check_cost_against_oracle(O, p, st, x, f[0], g, is_rt([c for c in CASES if c[0] == case[0]][0]))
        assert f[0] > 0.0 or seed == 0


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["hopper_example", "anymal_trot", "biped_walk_durations", "monoped_squat_jump_durations"])
def test_gpu_cost_matches_oracle(product, oracle, name):
    T, O = product, oracle
    if name == "hopper_example":
        s = hopper_spec(T)
    else:
        s = mk(T, [c for c in CASES if c[0] == name][0])
    if name == "monoped_squat_jump_durations":
        squat_jump_costs(s)                                    # individual NodeCost terms, squat_jump_example.cc:205-211
    else:
        with_costs(s, [(FORCES, 1e-9), (EE_MOTION, 1e-4)])
    st = T.Structure(s)
    p = O.Problem(to_oracle_spec(O, s))
    n = 6
    x = np.stack([perturbed(st.x0(s), 70 + i, 0.05 if i else 0.0) for i in range(n)])
    b = T.Batch(st, n)
    f, g = b.eval_cost(x)
    f1, _ = b.eval_cost(x, 1)
    _, g2 = b.eval_cost(x, 2)
    assert np.array_equal(f, f1) and np.array_equal(g, g2)
    for i in range(n):
        assert tol_ratio(np.array([f[i]]), np.array([p.eval_f(x[i])])) <= 1.0
        assert tol_ratio(g[i], p.eval_grad_f(x[i])) <= 1.0
    # the constraint path is unchanged by the cost list
    gg, vv = b.eval(x)
    assert np.isfinite(gg).all() and np.isfinite(vv).all()


@pytest.mark.gpu
@pytest.mark.parametrize("case", SAMPLED, ids=lambda c: c[0] + "".join("_%d" % cid for cid, _ in c[1]) + ("_swing" if len(c) > 5 else ""))
def test_gpu_sampled_costs_match_oracle(product, oracle, case):
    """EnergyCost / AngularMomentumCost on the GPU; problem 1 has its own inertia (AngularMomentumCost reads the
    batch problem's I_b)."""
    T, O = product, oracle
    s = sampled_spec(T, *case)
    st = T.Structure(s)
    n = 5
    x = np.stack([perturbed(st.x0(s), 70 + i, 0.05 if i else 0.0) for i in range(n)])
    b = T.Batch(st, n)
    b.set_all(s)
    s1 = sampled_spec(T, *case)
    s1.inertia[0] *= 1.5
    s1.inertia[4] += 0.05
    s1.initial_base_ang[2] = -0.3                  # ... and its own EEBasePosCost reference positions
    s1.initial_ee_W[0][1] += 0.02
    b.set_problem(1, s1)
    f, g = b.eval_cost(x)
    f1, _ = b.eval_cost(x, 1)
    _, g2 = b.eval_cost(x, 2)
    assert np.array_equal(f, f1) and np.array_equal(g, g2)
    for i in range(n):
        p = O.Problem(to_oracle_spec(O, s1 if i == 1 else s))
        check_cost_against_oracle(O, p, st, x[i], f[i], g[i], is_rt([c for c in CASES if c[0] == case[0]][0]))
    # device entry point on the caller's buffers and stream
    import torch
    dx = torch.from_numpy(x).cuda()
    df = torch.full((n,), float("nan"), dtype=torch.float64, device="cuda")
    dg = torch.full((n, st.n_vars), float("nan"), dtype=torch.float64, device="cuda")
    b.eval_cost_device(n, 3, dx.data_ptr(), df.data_ptr(), dg.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert np.array_equal(df.cpu().numpy(), f) and np.array_equal(dg.cpu().numpy(), g)


@pytest.mark.gpu
def test_gpu_no_costs_is_zero(product):
    T = product
    s = mk(T, CASES[0])
    st = T.Structure(s)
    b = T.Batch(st, 3)
    f, g = b.eval_cost(np.tile(st.x0(s), (3, 1)))
    assert (f == 0.0).all() and (g == 0.0).all()
