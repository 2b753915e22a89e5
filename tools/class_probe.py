#!/usr/bin/env python
"""One point of the mixed sweep with the batch's kernel choice printed (TOWR_B200_DEBUG=1 and the TOWR_B200_LANES /
TOWR_B200_GCAP / TOWR_B200_KERNEL knobs apply):   python tools/class_probe.py <monoped|biped|anymal> <dt> [batch]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import towr2025_b200 as T  # noqa: E402


def main():
    name, dt = sys.argv[1], float(sys.argv[2])
    batch = int(sys.argv[3]) if len(sys.argv) > 3 else 4096
    robot, combo, Tt = {"monoped": (T.Monoped, 0, 2.4), "biped": (T.Biped, 0, 2.4), "anymal": (T.Anymal, 1, 2.0)}[name]
    s = T.make_spec(robot, combo, Tt)
    s.dt_constraint_dynamic, s.dt_constraint_range_of_motion, s.dt_constraint_force = dt, 0.8 * dt, 0.2 * dt
    st = T.Structure(s)
    b = T.Batch(st, batch, 0)
    b.set_all(s)
    x = torch.from_numpy(st.x0(s)[None, :] + np.random.default_rng(0).uniform(-0.02, 0.02, (batch, st.n_vars))).cuda()
    g = torch.empty((batch, st.n_rows), dtype=torch.float64, device="cuda")
    v = torch.empty((batch, st.nnz), dtype=torch.float64, device="cuda")
    for _ in range(3):
        b.eval_device(batch, T.EVAL_BOTH, x.data_ptr(), g.data_ptr(), v.data_ptr())
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        b.eval_device(batch, T.EVAL_BOTH, x.data_ptr(), g.data_ptr(), v.data_ptr())
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print("%s dt %.3f batch %d: nnz %d, %.3f ms, %.1f GB/s" % (name, dt, batch, st.nnz, ms, st.bytes_per_eval * batch / ms / 1e6))


if __name__ == "__main__":
    main()
