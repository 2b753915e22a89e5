"""eval_f / eval_grad_f: the NodeCost terms of NlpFormulation::GetCosts (nlp_formulation.cc:605-681, costs/node_cost.cc),
as test/hopper_example.cc:147-148 enables them.  Oracle known answers on the CPU, CUDA path against the oracle on the GPU."""
import numpy as np
import pytest

from test_hopper_example import hopper_spec
from util import CASES, mk, perturbed, to_oracle_spec, tol_ratio

FORCES, EE_MOTION, ENERGY = 0, 1, 2


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
    with pytest.raises(T.TowrB200Error) as e:
        T.Structure(with_costs(mk(T, CASES[4]), [(ENERGY, 1.0)]))
    assert e.value.code == -2                                  # unsupported
    with pytest.raises(T.TowrB200Error):
        T.Structure(with_costs(mk(T, CASES[4]), [(7, 1.0)]))   # "cost not defined!" (nlp_formulation.cc:644)
    a = T.Structure(with_costs(mk(T, CASES[4]), [(FORCES, 1.0)]))
    assert not a.accepts(mk(T, CASES[4]))                      # the cost list is part of the structure class


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["hopper_example", "anymal_trot", "biped_walk_durations"])
def test_gpu_cost_matches_oracle(product, oracle, name):
    T, O = product, oracle
    if name == "hopper_example":
        s = hopper_spec(T)
    else:
        s = mk(T, [c for c in CASES if c[0] == name][0])
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
def test_gpu_no_costs_is_zero(product):
    T = product
    s = mk(T, CASES[0])
    st = T.Structure(s)
    b = T.Batch(st, 3)
    f, g = b.eval_cost(np.tile(st.x0(s), (3, 1)))
    assert (f == 0.0).all() and (g == 0.0).all()
