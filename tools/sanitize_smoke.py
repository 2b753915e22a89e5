#!/usr/bin/env python
"""Small invocation of every kernel through the C ABI, meant to run under compute-sanitizer:
    compute-sanitizer --tool memcheck python tools/sanitize_smoke.py [kernel]      (kernel: lockstep | ring | streams | streamsring)"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
if len(sys.argv) > 1:
    os.environ["TOWR_B200_KERNEL"] = sys.argv[1]
import numpy as np  # noqa: E402
import towr2025_b200 as T  # noqa: E402


def main():
    fixed_only = os.environ.get("TOWR_B200_KERNEL", "lockstep") == "streams"
    for dur in ([False] if fixed_only else [False, True]):
        for robot, combo, Tt in ((T.Anymal, T.C1, 2.0), (T.Monoped, 0, 2.4)):
            s = T.make_spec(robot, combo, Tt, optimize_durations=dur)
            s.n_costs = 1
            s.cost_ids[0], s.cost_weights[0] = 0, 1e-3
            st = T.Structure(s)
            n = 5
            b = T.Batch(st, n, 0)
            b.set_all(s)
            x = np.tile(st.x0(s), (n, 1)) + np.random.default_rng(1).uniform(-0.01, 0.01, (n, st.n_vars))
            for m in (n, 3, 1):
                g, v = b.eval(x[:m])
                assert np.isfinite(g).all() and np.isfinite(v).all()
            f, gr = b.eval_cost(x)
            rows = b.sample_trajectory(x, 0.05)
            assert np.isfinite(f).all() and np.isfinite(gr).all() and np.isfinite(rows).all()
            print("ok", robot, dur, st.nnz, flush=True)
    if fixed_only or os.environ.get("TOWR_B200_KERNEL", "lockstep") != "lockstep":
        return
    # the extended instantiations and the multi-part path: Gap, a table terrain, TerrainHard, a fine-dt class, the
    # dynamic-only transfer, arbitrary-time sampling, the allocation-free cost entry point's batch twin
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from util import set_terrain_table
    extra = []
    s = T.make_spec(T.Hyq, T.C4, 2.4, terrain=T.GapID, optimize_durations=True)
    extra.append(("gap", s))
    s = T.make_spec(T.Biped, T.C0, 2.4, extra_constraints=(T.Torque, T.TerrainHard))
    set_terrain_table(s, T.TABLE_RAMPS, [0.4, 0.9, 1.3], [0.0, 0.2, 0.1])
    extra.append(("table+terrainhard", s))
    s = T.make_spec(T.Monoped, T.C0, 2.4)
    s.dt_constraint_dynamic, s.dt_constraint_range_of_motion, s.dt_constraint_force = 0.01, 0.008, 0.002
    extra.append(("parts", s))
    for name, s in extra:
        st = T.Structure(s)
        n = 5
        b = T.Batch(st, n, 0)
        b.set_all(s)
        x = np.tile(st.x0(s), (n, 1)) + np.random.default_rng(2).uniform(-0.01, 0.01, (n, st.n_vars))
        g, v = b.eval(x)
        base = np.stack([b.constant_values(i) for i in range(n)])
        g2, v2 = b.eval(x, T.EVAL_G | T.EVAL_JAC_DYNAMIC, jac_out=base)
        assert np.array_equal(v, v2) and np.array_equal(g, g2) and (b.status() == 0).all()
        rows = b.sample_points(x, np.array([0.0, 0.37, 1.234]))
        assert np.isfinite(rows).all()
        print("ok", name, st.nnz, b.kernel_info, flush=True)


if __name__ == "__main__":
    main()
