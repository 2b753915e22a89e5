"""Device time of eval_f + eval_grad_f for a batch (CUDA events on the launching stream): NodeCost only, and with the
time-discretised EnergyCost / AngularMomentumCost terms.  Prints one JSON line per configuration."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import towr2025_b200 as T


def run(name, robot, combo, total_time, optimise, costs, n=4096, steps=20, warmup=5, swing=False):
    s = T.make_spec(robot, combo, total_time, optimize_durations=optimise)
    if swing:                                    # one EEBasePosCost per foot
        s.enable_swing_ee_base_pos_tracking = 1
        name += " + swing tracking"
    s.n_costs = len(costs)
    for i, (cid, w) in enumerate(costs):
        s.cost_ids[i], s.cost_weights[i] = cid, w
    st = T.Structure(s)
    b = T.Batch(st, n)
    b.set_all(s)
    rng = np.random.default_rng(0)
    x = torch.from_numpy(st.x0(s)[None, :] + 0.02 * rng.standard_normal((n, st.n_vars))).cuda()
    f = torch.empty((n,), dtype=torch.float64, device="cuda")
    g = torch.empty((n, st.n_vars), dtype=torch.float64, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    for _ in range(warmup):
        b.eval_cost_device(n, 3, x.data_ptr(), f.data_ptr(), g.data_ptr(), stream)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        b.eval_cost_device(n, 3, x.data_ptr(), f.data_ptr(), g.data_ptr(), stream)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    assert torch.isfinite(f).all() and torch.isfinite(g).all()
    if "--clocks" in sys.argv and hasattr(T.lib(), "tb_cost_clocks"):      # make cost_timing build, TOWR_B200_LIB
        import ctypes as C
        c = (C.c_longlong * 16)()
        T.lib().tb_cost_clocks(c)
        names = ["nodes", "poly records", "coefficients", "items", "summands + phase sums", "term sums", "gradient jobs", "tail barrier"]
        d = [c[i + 1] - c[i] for i in range(8)]
        print(json.dumps({"workload": name, "costs": costs, "clocks_cta0_problem0": dict(zip(names, d)), "total": c[8] - c[0]}))
    print(json.dumps({"workload": name, "costs": costs, "batch": n, "n_vars": st.n_vars, "ms_per_eval": round(ms, 4),
                      "cost_evals_per_s": round(n / ms * 1e3)}))


if __name__ == "__main__":
    run("anymal_trot", T.Anymal, T.C1, 2.0, False, [(0, 1e-3), (1, 1e-2)])
    run("anymal_trot", T.Anymal, T.C1, 2.0, False, [(0, 1e-3), (1, 1e-2), (2, 1e-3), (3, 0.5)])
    run("anymal_trot_durations", T.Anymal, T.C1, 2.0, True, [(0, 1e-3), (1, 1e-2), (2, 1e-3), (3, 0.5)])
    run("monoped_hop_durations", T.Monoped, T.C0, 2.4, True, [(2, 1e-3), (3, 0.5)])
    run("anymal_trot", T.Anymal, T.C1, 2.0, False, [(0, 1e-3), (1, 1e-2), (2, 1e-3), (3, 0.5)], swing=True)
    run("anymal_trot_durations", T.Anymal, T.C1, 2.0, True, [(0, 1e-3), (1, 1e-2), (2, 1e-3), (3, 0.5)], swing=True)
    if "--split" in sys.argv:
        run("anymal_trot", T.Anymal, T.C1, 2.0, False, [(2, 1e-3)])
        run("anymal_trot", T.Anymal, T.C1, 2.0, False, [(3, 0.5)])
        run("anymal_trot_durations", T.Anymal, T.C1, 2.0, True, [(2, 1e-3)])
        run("anymal_trot_durations", T.Anymal, T.C1, 2.0, True, [(3, 0.5)])
