#!/usr/bin/env python
"""BASELINE.json configs[4] as written: mixed-robot sweep (monoped / biped / quadruped), batch
256 -> 65536, dt_constraint_dynamic in {0.1, 0.05, 0.02, 0.01} with the other sampling steps scaled along, 16 distinct x
per problem rotated between launches: device-resident evals/s against the HBM roofline (algorithmic bytes
8*(n_vars+n_rows+nnz) per eval).  Points whose buffers would not fit 60 % of the free HBM are skipped and say so.

    python tools/mixed_sweep.py [out.json]        (needs a B200; one JSON line per point)
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import towr2025_b200 as T  # noqa: E402

ROBOTS = [("monoped", T.Monoped, 0, 2.4), ("biped", T.Biped, 0, 2.4), ("anymal", T.Anymal, 1, 2.0)]
BATCHES = [256, 1024, 4096, 16384, 65536]
DTS = [0.1, 0.05, 0.02, 0.01]       # dt_constraint_dynamic; range-of-motion 0.8 dt, force 0.2 dt (the default ratios)
X_VARIANTS = 16


def peak_gbs():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        return 6650.0


def main():
    out = open(sys.argv[1], "w") if len(sys.argv) > 1 else None
    peak = peak_gbs()
    free = torch.cuda.mem_get_info()[0]
    for name, robot, combo, Tt in ROBOTS:
        for dt in DTS:
            s = T.make_spec(robot, combo, Tt)
            s.dt_constraint_dynamic = dt
            s.dt_constraint_range_of_motion = 0.8 * dt
            s.dt_constraint_force = 0.2 * dt
            try:
                st = T.Structure(s)
            except T.TowrB200Error as e:                # class beyond the compact tables' 16-bit offsets at this dt
                line = {"robot": name, "dt_dynamic": dt, "batch": None, "skipped": str(e)}
                print(json.dumps(line), flush=True)
                if out:
                    out.write(json.dumps(line) + "\n")
                continue
            x0 = st.x0(s)
            for batch in BATCHES:
                nx = X_VARIANTS if batch * st.n_vars * 8 * X_VARIANTS < 0.1 * free else 4
                need = batch * 8 * (st.n_vars * nx + st.n_rows + st.nnz)
                if need > 0.6 * free:
                    line = {"robot": name, "dt_dynamic": dt, "batch": batch, "skipped": "needs %.1f GB" % (need / 1e9)}
                    print(json.dumps(line), flush=True)
                    if out:
                        out.write(json.dumps(line) + "\n")
                    continue
                try:
                    b = T.Batch(st, batch, 0)
                except T.TowrB200Error as e:            # class tables beyond the SM's shared memory at this dt
                    line = {"robot": name, "dt_dynamic": dt, "batch": batch, "skipped": str(e)}
                    print(json.dumps(line), flush=True)
                    if out:
                        out.write(json.dumps(line) + "\n")
                    break
                b.set_all(s)
                rng = np.random.default_rng(batch)
                xs = [torch.from_numpy(x0[None, :] + rng.uniform(-0.05, 0.05, (batch, st.n_vars))).cuda() for _ in range(nx)]
                g = torch.empty((batch, st.n_rows), dtype=torch.float64, device="cuda")
                v = torch.empty((batch, st.nnz), dtype=torch.float64, device="cuda")
                steps = 20 if batch <= 4096 else 6
                for k in range(3):
                    b.eval_device(batch, T.EVAL_BOTH, xs[k % nx].data_ptr(), g.data_ptr(), v.data_ptr())
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for k in range(steps):
                    b.eval_device(batch, T.EVAL_BOTH, xs[k % nx].data_ptr(), g.data_ptr(), v.data_ptr())
                e1.record()
                torch.cuda.synchronize()
                ms = e0.elapsed_time(e1) / steps
                assert (b.status() == 0).all() and bool(torch.isfinite(v[:1024]).all())   # status covers every value; the slice is a spot check
                line = {"robot": name, "dt_dynamic": dt, "batch": batch, "x_variants": nx, "kernel": b.kernel_info, "n_vars": st.n_vars, "n_rows": st.n_rows,
                        "nnz": st.nnz, "ms_per_eval_batch": ms, "evals_per_s": batch / (ms * 1e-3),
                        "achieved_gbs": st.bytes_per_eval * batch / (ms * 1e-3) / 1e9,
                        "roofline_frac": st.bytes_per_eval * batch / (ms * 1e-3) / 1e9 / peak}
                print(json.dumps(line), flush=True)
                if out:
                    out.write(json.dumps(line) + "\n")
                del b, xs, g, v
                torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
