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


if __name__ == "__main__":
    main()
