"""The reference's second example, towr/test/biped_example.cc:50-180, stated through the spec fields: Biped on the
FiveStepStairs terrain (:14-48, the same profile as the hopper example's), 4 force / torque polynomials per stance (:59-60),
the stepping schedule built by the loop of :70-99 (fixed durations: OptimizePhaseDurations is commented out, :179), stance
tracking with the footholds of that loop, initial / final states :156-167, +Torque +TerrainHard :173-174, ForcesCostID /
EEMotionCostID :176-177."""
import numpy as np
import pytest

from util import check_against_oracle, perturbed, to_oracle_spec

TORQUE, TERRAIN_HARD = 6, 4


def five_step_stairs(x):                        # biped_example.cc:24-41
    if x < 0.5:
        return 0.0
    n = int((x - 0.5) / 0.3)
    return 5 * 0.15 if n >= 5 else (n + 1) * 0.15


def biped_spec(M):
    s = M.make_spec(M.Biped, 0, 2.4)            # robot model and defaults; gait and states are replaced below
    s.terrain_id = 7
    s.force_polynomials_per_stance_phase = 4    # :59-61
    s.torque_polynomials_per_stance_phase = 4
    s.ee_polynomials_per_swing_phase = 2
    z0, x0, step_t, stand_t, ds_rate, phase_diff_rate = 0.65, 0.0, 0.60, 0.5, 0.2, 1.0      # :70-79
    st_, ft = step_t * (1 + ds_rate), step_t * (1 - ds_rate)
    xend, step_len = 2.0 + x0, 0.5 * step_t
    phase_l, phase_r = [stand_t, ft], [stand_t + step_t * phase_diff_rate, ft]             # :81-82
    stance_l, stance_r = [(x0, 0.1)], [(x0, -0.1)]                                           # :83-84
    stance_x = x0
    while stance_x < xend:                                                                   # :90-98
        phase_l += [st_, ft]
        phase_r += [st_, ft]
        stance_l.append((max(stance_x - step_len, x0), 0.1))
        stance_r.append((stance_x, -0.1))
        stance_x += 2 * step_len
    phase_l.append(stand_t + step_t * phase_diff_rate)                                       # :99-102
    phase_r.append(stand_t)
    stance_l.append((xend, 0.1))
    stance_r.append((xend, -0.1))
    for e, (ph, stn) in enumerate(((phase_l, stance_l), (phase_r, stance_r))):
        s.n_phases[e] = len(ph)
        for i, d in enumerate(ph):
            s.phase_durations[e][i] = d
        s.in_contact_at_start[e] = 1            # :169-170
        s.n_stance_positions[e] = len(stn)
        for k, (px, py) in enumerate(stn):
            s.ee_stance_position[e][k][0] = px
            s.ee_stance_position[e][k][1] = py
    s.enable_stance_tracking = 1                # the fork's default (parameters.h:225)
    for i, v in enumerate([x0, 0.0, z0 + five_step_stairs(x0), 0.0, 0.0, 0.0]):             # :156-157
        s.initial_base_lin[i] = v
    for i, v in enumerate([xend, 0.0, z0 + five_step_stairs(xend), 0.0, 0.0, 0.0]):         # :164-165
        s.final_base_lin[i] = v
    for i in range(6):
        s.initial_base_ang[i] = 0.0
        s.final_base_ang[i] = 0.0
    for e, y in enumerate((0.1, -0.1)):                                                      # :160-161
        s.initial_ee_W[e][0], s.initial_ee_W[e][1], s.initial_ee_W[e][2] = x0, y, five_step_stairs(x0)
    for c in (TORQUE, TERRAIN_HARD):                                                         # :173-174
        s.constraints[s.n_constraints] = c
        s.n_constraints += 1
    s.n_costs = 2                                                                            # :176-177
    s.cost_ids[0], s.cost_weights[0] = 0, 5 * 1e-9
    s.cost_ids[1], s.cost_weights[1] = 1, 1e-4
    return s


def test_biped_example_structure(product, oracle):
    """Set sizes, x0, bounds and the Jacobian pattern of the example, bit-exact against the oracle."""
    T, O = product, oracle
    s = biped_spec(T)
    st = T.Structure(s)
    p = O.Problem(to_oracle_spec(O, s))
    assert (st.n_vars, st.n_rows) == (p.n_vars, p.n_rows)
    assert st.var_sets() == p.var_sets() and st.con_sets() == p.con_sets()
    assert any(name.startswith("terrainhard-") for name, _ in st.con_sets())
    x0 = st.x0(s)
    assert np.array_equal(x0, p.x0())
    for x in (x0, perturbed(x0, 3, 0.05)):
        r, c, _ = p.eval_jac(x)
        rows, cols = st.pattern()
        assert np.array_equal(r, rows) and np.array_equal(c, cols)
    for a, b in zip(st.var_bounds(s) + st.con_bounds(s), p.var_bounds() + p.con_bounds()):
        assert np.array_equal(np.clip(a, -1e20, 1e20), np.clip(b, -1e20, 1e20))


def test_biped_example_emulated(product, oracle, emu):
    from test_emu import emu_eval
    T, O = product, oracle
    s = biped_spec(T)
    st = T.Structure(s)
    x = perturbed(st.x0(s), 5, 0.05)
    g, v = emu_eval(emu, s, st, x)
    check_against_oracle(O, to_oracle_spec(O, s), st, x, g, v, False)


@pytest.mark.gpu
def test_biped_example_on_gpu(product, oracle):
    """All four Ipopt callbacks of the example for a batch of perturbed iterates, against the oracle."""
    from util import tol_ratio
    T, O = product, oracle
    s = biped_spec(T)
    st = T.Structure(s)
    n = 4
    x = np.stack([perturbed(st.x0(s), 20 + i, 0.05 if i else 0.0) for i in range(n)])
    b = T.Batch(st, n)
    b.set_all(s)
    g, v = b.eval(x)
    f, grad = b.eval_cost(x)
    assert (b.status() == 0).all()
    p = O.Problem(to_oracle_spec(O, s))
    for i in range(n):
        check_against_oracle(O, to_oracle_spec(O, s), st, x[i], g[i], v[i], False)
        assert tol_ratio(np.array([f[i]]), np.array([p.eval_f(x[i])])) <= 1.0
        assert tol_ratio(grad[i], p.eval_grad_f(x[i])) <= 1.0
