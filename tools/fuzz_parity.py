"""Random formulations against the oracle: hand-made phase timings (2-6 phases per foot, a first phase in stance or in
swing, feet that differ), random terrain, constraint subset, sampling steps, EELinearConstraint definitions and costs.
TEST INFRASTRUCTURE.  Default: the host emulation of the device arithmetic (tests/emu, no GPU); `gpu`: the CUDA path.

    python tools/fuzz_parity.py [gpu | cross] [n_cases] [seed]

`cross`: the same random formulations through the two oracle legs instead (oracle/*.cc against oracle/oracle_np.py):
structure, bounds, g, the stored-entry pattern and values, f and its gradient.
"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import towr2025_b200 as T                      # noqa: E402
from oracle import oracle_py as O              # noqa: E402  (checker)
from util import check_against_oracle, check_cost_against_oracle, cost_gradient_ok, parity_ok, pow_ok, perturbed, to_oracle_spec, tol_ratio, set_ee_linear, set_terrain_table   # noqa: E402

DP = C.POINTER(C.c_double)


def random_spec(rng):
    robot = int(rng.integers(0, 4))
    n_ee = (1, 2, 4, 4)[robot]
    total = float(rng.choice([1.2, 1.7, 2.0, 2.4]))
    optimise = bool(rng.integers(0, 2))
    s = T.make_spec(robot, 0, total, goal=(float(rng.uniform(0.2, 1.2)), float(rng.uniform(-0.2, 0.2)), float(rng.uniform(-0.5, 0.5))),
                    terrain=0, optimize_durations=optimise)
    # SwingConstraint reads the neighbours of every swing node (swing_constraint.cc:66-70): a motion that starts or ends
    # in swing runs off the node vector there, in the reference too.  Mostly stance-first-and-last; else without Swing.
    ends_in_stance = rng.random() < 0.8
    for e in range(n_ee):                                   # params_.ee_phase_durations_, ee_in_contact_at_start_
        n = int(rng.choice([3, 5])) if ends_in_stance else int(rng.integers(2, 7))
        d = rng.uniform(0.25, 1.0, size=n)
        d = d / d.sum() * total
        s.n_phases[e] = n
        s.in_contact_at_start[e] = 1 if ends_in_stance else int(rng.integers(0, 2))
        for i in range(n):
            s.phase_durations[e][i] = float(d[i])
    terrain = int(rng.integers(0, 9))
    if terrain == 8:
        xs = np.unique(np.round(np.sort(rng.uniform(0.1, 1.5, size=int(rng.integers(1, 5)))), 2))
        set_terrain_table(s, int(rng.integers(0, 2)), [float(v) for v in xs], [float(v) for v in rng.uniform(0.0, 0.3, size=len(xs))])
    else:
        s.terrain_id = terrain
    cons = [T.Terrain, T.Dynamic, T.BaseAcc, T.EndeffectorRom, T.Force, T.Swing, T.BaseHeight]
    keep = [c for c in cons if rng.random() < 0.85 and (ends_in_stance or c != T.Swing)]
    for extra in (T.Torque, T.BaseRom, T.TerrainHard):
        if rng.random() < 0.3:
            keep.append(extra)
    rng.shuffle(keep)
    if optimise:
        keep.append(T.TotalTime)
    s.n_constraints = len(keep)
    for i, c in enumerate(keep):
        s.constraints[i] = int(c)
    s.dt_constraint_dynamic = float(rng.choice([0.1, 0.07, 0.05]))
    s.dt_constraint_range_of_motion = float(rng.choice([0.08, 0.11, 0.04]))
    s.dt_constraint_force = float(rng.choice([0.02, 0.05, -1.0]))
    s.dt_constraint_torque = float(rng.choice([0.02, 0.06, -1.0]))
    s.dt_constraint_base_motion = float(rng.choice([0.025, 0.04]))
    if rng.random() < 0.4:                                   # spline resolution (parameters.cc:61-66)
        s.duration_base_polynomial = float(rng.choice([0.1, 0.15, 0.07]))
        s.force_polynomials_per_stance_phase = int(rng.choice([2, 3, 4]))
        s.torque_polynomials_per_stance_phase = int(rng.choice([2, 3, 4]))
        s.ee_polynomials_per_swing_phase = int(rng.choice([2, 3]))
    if rng.random() < 0.3:                                   # stance tracking: bounds only (nlp_formulation.cc:218-231)
        s.enable_stance_tracking = 1
        for e in range(n_ee):
            s.n_stance_positions[e] = 4
            for k in range(4):
                s.ee_stance_position[e][k][0] = float(rng.uniform(-0.3, 1.2))
                s.ee_stance_position[e][k][1] = float(rng.uniform(-0.3, 0.3))
    if rng.random() < 0.4:
        defs = []
        for _ in range(int(rng.integers(1, 3))):
            terms = [(int(rng.integers(0, n_ee)), int(rng.integers(0, 3)), float(rng.uniform(-2, 2))) for _ in range(int(rng.integers(1, 5)))]
            defs.append((int(rng.integers(0, 2)), int(rng.integers(0, 2)), terms, float(rng.uniform(0.01, 0.5)), float(rng.choice([0.1, 0.06]))))
        set_ee_linear(s, defs)
    costs = [(int(c), float(rng.uniform(1e-4, 1.0))) for c in range(4) if rng.random() < 0.4]
    if rng.random() < 0.15 and len(costs) < 4:               # a family listed twice: its columns get two writers
        costs.append((int(rng.choice([2, 3])), float(rng.uniform(1e-3, 0.5))))
    s.dt_cost_energy = float(rng.choice([0.02, 0.05, 0.0]))
    s.energy_cost_torque_weight = float(rng.choice([1.0, 0.0, 0.3]))
    s.dt_cost_ang_mom = float(rng.choice([0.02, 0.07]))
    s.dt_cost_swing_ee_base_pos_tracking = float(rng.choice([0.05, 0.03]))
    if rng.random() < 0.3:                                   # individual NodeCost terms (node_cost.cc:36-45)
        terms = [(int(rng.integers(0, 6)), int(rng.integers(0, n_ee)), int(rng.integers(0, 2)), int(rng.integers(0, 3)),
                  float(rng.uniform(1e-4, 1.0))) for _ in range(int(rng.integers(1, 6)))]
        s.n_node_costs = len(terms)
        for i, (k, ee, deriv, dim, w) in enumerate(terms):
            s.node_cost_set[i], s.node_cost_ee[i], s.node_cost_deriv[i], s.node_cost_dim[i], s.node_cost_weight[i] = k, ee, deriv, dim, w
    s.n_costs = len(costs)
    for i, (c, w) in enumerate(costs):
        s.cost_ids[i], s.cost_weights[i] = c, w
    if rng.random() < 0.3:
        s.enable_swing_ee_base_pos_tracking = 1
        s.initial_base_ang[2] = float(rng.uniform(-0.5, 0.5))
    return s


def second_robot(s):
    """The same formulation for a robot with its own mass, inertia and kinematic box (per-problem constants of a class)."""
    s_b = T.Spec()
    C.memmove(C.byref(s_b), C.byref(s), C.sizeof(s))
    s_b.mass *= 1.3
    s_b.inertia[0] *= 1.2
    s_b.inertia[5] += 0.01
    for e in range(s.n_ee):
        s_b.nominal_stance[e][0] += 0.02
        s_b.max_dev[e][2] *= 1.1
    return s_b


def note(ok, passed, what):
    if not passed:
        print("  failed check: " + what, flush=True)
    return ok and passed


def describe(s):
    n_ee = s.n_ee
    return "robot_ee=%d T-phases=%s start=%s terrain=%d cons=%s dur=%s eelin=%d costs=%s swing=%d" % (
        n_ee, [int(s.n_phases[e]) for e in range(n_ee)], [int(s.in_contact_at_start[e]) for e in range(n_ee)], s.terrain_id,
        [int(s.constraints[i]) for i in range(s.n_constraints)], any(int(s.constraints[i]) == int(T.TotalTime) for i in range(s.n_constraints)),
        s.n_ee_linear, [int(s.cost_ids[i]) for i in range(s.n_costs)], s.enable_swing_ee_base_pos_tracking)


def cross_check(n_cases, seed):
    from oracle import oracle_np as N
    rng = np.random.default_rng(seed)
    clip = lambda a: np.clip(a, -1e20, 1e20)
    for case in range(n_cases):
        s = to_oracle_spec(O, random_spec(rng))
        p, f = O.Problem(s), N.Formulation(s)
        why = []
        if [(v.name, v.rows()) for v in f.var_sets] != p.var_sets() or [(c.name, c.n_rows) for c in f.con_sets] != p.con_sets():
            why.append("sets")
        x0 = p.x0()
        if tol_ratio(f.x0(), x0) > 1.0:
            why.append("x0")
        for a, b in zip(f.var_bounds() + f.con_bounds(), p.var_bounds() + p.con_bounds()):
            if tol_ratio(clip(a), clip(b)) > 1.0:
                why.append("bounds")
        x = perturbed(x0, 100 + case, 0.03)
        if tol_ratio(f.eval_g(x), p.eval_g(x)) > 1.0:
            why.append("g")
        r, c, v = f.eval_jac(x)
        ro, co, vo = p.eval_jac(x)
        if len(r) != len(ro) or not (np.array_equal(r, ro) and np.array_equal(c, co)):
            why.append("pattern")
        elif tol_ratio(v, vo) > 1.0:
            why.append("jacobian values")
        if tol_ratio(np.array([f.eval_f(x)]), np.array([p.eval_f(x)])) > 1.0 or tol_ratio(f.eval_grad_f(x), p.eval_grad_f(x)) > 1.0:
            why.append("costs")
        if why:
            raise SystemExit("cross case %d FAILED (%s): %s" % (case, ", ".join(why), describe(s)))
    print("fuzz_parity (oracle legs): the C++ and the numpy restatement agree on %d random formulations; seed %d" % (n_cases, seed))


def main():
    if "cross" in sys.argv:
        nums = [a for a in sys.argv[1:] if a.lstrip("-").isdigit()]
        return cross_check(int(nums[0]) if nums else 20, int(nums[1]) if len(nums) > 1 else 1)
    gpu = "gpu" in sys.argv
    nums = [a for a in sys.argv[1:] if a.lstrip("-").isdigit()]
    n_cases = int(nums[0]) if nums else 60
    seed = int(nums[1]) if len(nums) > 1 else 1
    rng = np.random.default_rng(seed)
    E = None
    if not gpu:
        import subprocess
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "tests", "emu")], stdout=subprocess.DEVNULL)
        E = C.CDLL(os.path.join(ROOT, "tests", "emu", "_build", "libemu.so"))
        E.emu_last_error.restype = C.c_char_p
    done = skipped = 0
    worst = 0.0
    for case in range(n_cases):
        s = random_spec(rng)
        try:
            st = T.Structure(s)
        except T.TowrB200Error as e:
            skipped += 1
            print("case %d refused (%s): %s" % (case, str(e)[:90], describe(s)))
            continue
        try:
            p = O.Problem(to_oracle_spec(O, s))
        except RuntimeError as e:
            raise SystemExit("case %d: the product accepts what the oracle refuses (%s): %s" % (case, e, describe(s)))
        ok = (st.n_vars, st.n_rows) == (len(p.x0()), sum(n for _, n in p.con_sets()))
        gap = s.terrain_id == 3
        optimise = any(int(s.constraints[i]) == int(T.TotalTime) for i in range(s.n_constraints))
        x0 = st.x0(s)
        ok = ok and tol_ratio(x0, p.x0()) <= 1.0
        clip = lambda a: np.clip(a, -1e20, 1e20)           # ifopt::inf
        for mine, ref in zip(st.var_bounds(s) + st.con_bounds(s), p.var_bounds() + p.con_bounds()):
            ok = ok and tol_ratio(clip(mine), clip(ref)) <= 1.0
        xs = [perturbed(x0, 100 + case, float(os.environ.get("FUZZ_SCALE", "0.03"))), x0]   # FUZZ_SCALE: size of the perturbation
        r_o, c_o, _ = p.eval_jac(xs[0])
        rows, cols = st.pattern()
        if not gap:
            ok = ok and len(rows) == len(r_o) and np.array_equal(rows, r_o) and np.array_equal(cols, c_o)
        if gpu:
            b = T.Batch(st, len(xs))
            b.set_all(s)
            g_all, v_all = b.eval(np.stack(xs))
            f_all, gr_all = b.eval_cost(np.stack(xs))
            # the dynamic-only transfer (constants once, x-dependent entries per eval) equals the full result bit for bit
            full = np.stack([b.constant_values(i) for i in range(len(xs))])
            b.eval(np.stack(xs), T.EVAL_JAC_DYNAMIC, jac_out=full)
            ok = note(ok, np.array_equal(full, v_all), 'dynamic-only transfer')
            # a second problem of the class with its own robot (mass, inertia, kinematic box) and terrain of the same
            # kind: per-problem constants, checked against its own oracle problem
            s_b = second_robot(s)
            other = int(rng.integers(0, 8))                  # ... on another terrain, where the class admits it
            keep_terrain = s_b.terrain_id
            s_b.terrain_id = other
            if not st.accepts(s_b):
                s_b.terrain_id = keep_terrain
            gap_b = gap or s_b.terrain_id == 3
            if st.accepts(s_b):
                b.set_problem(1, s_b)
                g_b, v_b = b.eval(np.stack(xs))
                f_b, gr_b = b.eval_cost(np.stack(xs))
                try:
                    check_against_oracle(O, to_oracle_spec(O, s_b), st, xs[1], g_b[1], v_b[1], optimise, superset=gap_b)
                except AssertionError as err:
                    print("case %d (second robot): %s" % (case, err))
                    ok = False
                p_b = O.Problem(to_oracle_spec(O, s_b))
                try:
                    check_cost_against_oracle(O, p_b, st, xs[1], float(f_b[1]), gr_b[1], optimise)
                except AssertionError as err:
                    ok = note(ok, False, 'second problem: %s' % err)
                ok = note(ok, np.array_equal(g_b[0], g_all[0]) and np.array_equal(v_b[0], v_all[0]), 'first problem changed by the second')
                b.set_problem(1, s)
            dt = float(rng.choice([0.05, 0.013]))            # SaveTrajectoryToCSV's sampling loop (save_data.cpp:54-119)
            rows = b.sample_trajectory(np.stack(xs), dt)
            for i, x in enumerate(xs):                       # as tests/test_trajectory.py: strict against exact powers
                if optimise:
                    with O.pow_mode(1):
                        ok = note(ok, parity_ok(rows[i], p.sample_trajectory(x, dt)), 'trajectory sampler (exact-power oracle)')
                    ok = note(ok, pow_ok(rows[i], p.sample_trajectory(x, dt)), 'trajectory sampler (std::pow oracle)')
                else:
                    ok = note(ok, parity_ok(rows[i], p.sample_trajectory(x, dt)), 'trajectory sampler')
                if not ok:                                   # where: sample, column (save_data.cpp:68-112 order), values
                    ref = p.sample_trajectory(x, dt)
                    bad = np.argwhere(np.abs(rows[i] - ref) > 1e-14 + 1e-12 * np.abs(ref))
                    print("  dt %g, %d entries differ; first: %s" % (dt, len(bad), [(int(a), int(c), float(rows[i][a, c]), float(ref[a, c])) for a, c in bad[:6]]))
                    print("  columns that differ:", sorted(set(int(c) for _, c in bad)))
        if not gpu:                                          # the second robot as a class of its own (the emulation has no batch)
            s_b = second_robot(s)
            g, v = np.zeros(st.n_rows), np.zeros(st.nnz)
            assert E.emu_eval(C.byref(s_b), x0.ctypes.data_as(DP), g.ctypes.data_as(DP), v.ctypes.data_as(DP)) == 0, E.emu_last_error()
            try:
                check_against_oracle(O, to_oracle_spec(O, s_b), st, x0, g, v, optimise, superset=gap)
            except AssertionError as err:
                print("case %d (second robot): %s" % (case, err))
                ok = False
        for i, x in enumerate(xs):
            if gpu:
                g, v, f, gr = g_all[i], v_all[i], f_all[i], gr_all[i]
            else:
                g, v = np.zeros(st.n_rows), np.zeros(st.nnz)
                assert E.emu_eval(C.byref(s), x.ctypes.data_as(DP), g.ctypes.data_as(DP), v.ctypes.data_as(DP)) == 0, E.emu_last_error()
                fa, gr = np.zeros(1), np.zeros(st.n_vars)
                assert E.emu_eval_cost(C.byref(s), x.ctypes.data_as(DP), fa.ctypes.data_as(DP), gr.ctypes.data_as(DP)) == 0, E.emu_last_error()
                f = fa[0]
            try:                                             # strict band; optimised durations as in the parity tests
                check_against_oracle(O, to_oracle_spec(O, s), st, x, g, v, optimise, superset=gap)
                jac_ok = g_ok = True
            except AssertionError as err:
                print("case %d: %s" % (case, err))
                jac_ok = g_ok = False
            try:
                check_cost_against_oracle(O, p, st, x, float(f), gr, optimise)
                c_ok = True
            except AssertionError as err:
                print("case %d: %s" % (case, err))
                c_ok = False
            if not (ok and jac_ok and g_ok and c_ok):
                raise SystemExit("case %d x%d FAILED (structure %s, jac %s, g %s, cost %s): %s" % (case, i, ok, jac_ok, g_ok, c_ok, describe(s)))
        done += 1
    print("fuzz_parity (%s): %d random formulations match the oracle, %d refused as unsupported; seed %d" %
          ("CUDA path" if gpu else "host emulation", done, skipped, seed))


if __name__ == "__main__":
    main()
