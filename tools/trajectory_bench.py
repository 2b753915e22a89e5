#!/usr/bin/env python
"""Device-resident throughput of the trajectory sampler (towr_b200_batch_sample_trajectory_device) on one B200.

    python tools/trajectory_bench.py [batch] [dt]
Prints one JSON line: trajectories/s, rows/s and the HBM roofline fraction (algorithmic bytes = 8 (n_vars + samples * width))."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import towr2025_b200 as T  # noqa: E402


def main():
    batch = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
    dt = float(sys.argv[2]) if len(sys.argv) > 2 else 0.01
    out = []
    for name, dur in (("anymal_trot", False), ("anymal_trot_durations", True)):
        s = T.make_spec(T.Anymal, T.C1, 2.0, optimize_durations=dur)
        st = T.Structure(s)
        b = T.Batch(st, batch, 0)
        x = np.tile(st.x0(s), (batch, 1)) + np.random.default_rng(0).uniform(-0.01, 0.01, (batch, st.n_vars))
        dx = torch.from_numpy(x).cuda()
        K, W = st.trajectory_samples(dt), st.trajectory_width()
        rows = torch.empty((batch, K, W), dtype=torch.float64, device="cuda")
        stream = torch.cuda.current_stream().cuda_stream
        for _ in range(3):
            b.sample_trajectory_device(batch, dt, dx.data_ptr(), rows.data_ptr(), stream)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        steps = 20
        e0.record()
        for _ in range(steps):
            b.sample_trajectory_device(batch, dt, dx.data_ptr(), rows.data_ptr(), stream)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        bytes_ = 8.0 * batch * (st.n_vars + K * W)
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
        out.append({"workload": name, "batch": batch, "dt": dt, "samples": K, "width": W, "ms_per_launch": ms,
                    "trajectories_per_s": batch / (ms * 1e-3), "rows_per_s": batch * K / (ms * 1e-3),
                    "achieved_gbs": bytes_ / ms / 1e6, "hbm_frac": bytes_ / ms / 1e6 / peak})
    print(json.dumps(out))


if __name__ == "__main__":
    main()
