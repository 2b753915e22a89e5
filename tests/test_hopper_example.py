"""BASELINE config 0: the formulation of the reference's own example, towr/test/hopper_example.cc:98-150, stated through the
spec fields: Monoped on the FiveStepStairs terrain of :53-86, 13 phases :105-113, ee_stance_position_ :114-122 with
enable_stance_tracking left at the fork's default (ON, parameters.h:225), initial / final states :125-139, +Torque :145,
ForcesCostID / EEMotionCostID :147-148, OptimizePhaseDurations :150."""
import numpy as np
import pytest

from util import check_against_oracle, parity_ok, perturbed, pow_ok, to_oracle_spec

PHASES = [0.5, 0.3, 0.4, 0.3, 0.4, 0.3, 0.4, 0.3, 0.4, 0.3, 0.4, 0.3, 0.4]   # hopper_example.cc:105-113
STANCES = [(0.0, 0.0), (0.4, 0.0), (0.7, -0.0), (0.7, 0.0), (1.3, -0.0), (1.3, 0.0), (0.0, 0.0)]     # :114-122
TORQUE = 6


def hopper_spec(M):
    s = M.make_spec(M.Monoped, 0, 2.4)          # robot model and defaults; the gait is replaced below
    s.terrain_id = 7                            # FiveStepStairs
    s.n_phases[0] = len(PHASES)
    for i, d in enumerate(PHASES):
        s.phase_durations[0][i] = d
    s.in_contact_at_start[0] = 1                # :141
    for i, v in enumerate([0.0, 0.0, 0.6, 0.0, 0.0, 0.0]):     # :125-126
        s.initial_base_lin[i] = v
        s.final_base_lin[i] = v                 # :132: z = 0.6 + GetHeight(0, 0) = 0.6
    for i in range(6):
        s.initial_base_ang[i] = 0.0
        s.final_base_ang[i] = 0.0
    for i in range(3):
        s.initial_ee_W[0][i] = 0.0              # :129
    s.constraints[s.n_constraints] = TORQUE     # :145
    s.n_constraints += 1
    s.enable_stance_tracking = 1                # the fork's default, parameters.h:225
    s.n_stance_positions[0] = len(STANCES)      # :114-122
    for k, (px, py) in enumerate(STANCES):
        s.ee_stance_position[0][k][0] = px
        s.ee_stance_position[0][k][1] = py
    s.n_costs = 2                               # :147-148
    s.cost_ids[0], s.cost_weights[0] = 0, 1e-9  # ForcesCostID
    s.cost_ids[1], s.cost_weights[1] = 1, 1e-4  # EEMotionCostID
    if hasattr(M, "lib") and hasattr(M.lib(), "towr_b200_spec_optimize_durations"):
        M._check(M.lib().towr_b200_spec_optimize_durations(__import__("ctypes").byref(s)))
    return s


def test_hopper_example_structure(product, oracle):
    """Sizes of SURVEY.md section 8 (888 variables, 2 026 + 948 rows) and the pattern, bit-exact against the oracle."""
    T, O = product, oracle
    s = hopper_spec(T)
    st = T.Structure(s)
    p = O.Problem(to_oracle_spec(O, s))
    assert st.n_vars == p.n_vars == 888
    assert st.n_rows == p.n_rows == 2026 + 948
    x0 = st.x0(s)
    assert np.array_equal(x0, p.x0())
    r, c, _ = p.eval_jac(x0)
    rows, cols = st.pattern()
    assert np.array_equal(r, rows) and np.array_equal(c, cols)
    r2, c2, _ = p.eval_jac(perturbed(x0, 3, 0.02))
    assert np.array_equal(r2, rows) and np.array_equal(c2, cols)
    lo, up = st.con_bounds(s)
    olo, oup = p.con_bounds()
    assert np.array_equal(lo, olo) and np.array_equal(up, oup)
    # variable bounds with the stance-tracking branch (nlp_formulation.cc:218-231): 7 stance footholds fixed in x, y
    vlo, vup = st.var_bounds(s)
    ovlo, ovup = p.var_bounds()
    assert np.array_equal(vlo, ovlo) and np.array_equal(vup, ovup)
    names = st.var_sets()
    c0 = sum(n for nm, n in names[:2])
    fixed = [(i, vlo[i]) for i in range(c0, c0 + names[2][1]) if vlo[i] == vup[i]]
    assert len(fixed) == 14 and sorted(set(v for _, v in fixed)) == [0.0, 0.4, 0.7, 1.3]
    # fewer positions than stance phases: the reference throws std::out_of_range (.at()), here an error code
    bad = hopper_spec(T)
    bad.n_stance_positions[0] = 5
    with pytest.raises(T.TowrB200Error) as e:
        st.var_bounds(bad)
    assert e.value.code == -1 and "ee_stance_position_" in str(e.value)
    assert st.trajectory_samples(0.01) == 471 and st.trajectory_width() == 44      # T = 4.7 s


@pytest.mark.gpu
def test_hopper_example_gpu_parity(product, oracle):
    T, O = product, oracle
    s = hopper_spec(T)
    st = T.Structure(s)
    p = O.Problem(to_oracle_spec(O, s))
    n = 4
    x = np.stack([perturbed(st.x0(s), 40 + i, 0.02 if i else 0.0) for i in range(n)])
    b = T.Batch(st, n)
    b.set_all(s)
    g, v = b.eval(x)
    assert (b.status() == 0).all()
    for i in range(n):
        check_against_oracle(O, to_oracle_spec(O, s), st, x[i], g[i], v[i], True)
    rows = b.sample_trajectory(x, 0.02)
    for i in range(n):
        with O.pow_mode(1):
            assert parity_ok(rows[i], p.sample_trajectory(x[i], 0.02))
        assert pow_ok(rows[i], p.sample_trajectory(x[i], 0.02))
