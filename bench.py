#!/usr/bin/env python
"""bench.py — constraint + Jacobian evals/s of the batched TOWR backend (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--batch B]

Workload (BASELINE.json configs[2]): ANYmal fly-trot 2 s (RobotModel::Anymal, quadruped combo C1, default
Parameters, fixed durations): 1042 variables, 2732 constraint rows, 23992 stored Jacobian values per problem,
batch 4096 per GPU, random goals, 16 distinct x per problem rotated between steps.  A step = one batched eval
of all g and all Jacobian values for the rank's batch.  Problems are independent: the batch shards across
ranks with no collective on the data path (weak scaling: 4096 problems per GPU); after the timed region the
ranks all_gather one solution vector each (the only NCCL use, as in north_star).

`value` is device-resident throughput (CUDA events on the launching stream, max over ranks); `e2e` is the same
metric through towr_b200_batch_eval with pinned HOST buffers (H2D of x and D2H of g + values inside the timed
region).  `--impl reference` times the reference's CPU evaluation instead: the reference itself cannot be
compiled on this image (Eigen/ifopt/Ipopt absent), so it is the oracle port, OpenMP over
problems on all host cores.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "ANYmal-trot constraint+Jacobian evals/s"
# name -> (robot, gait combo, total time, optimise durations, terrain, description).  The default is the configuration
# BASELINE.json's metric is quoted on (configs[2]); the others are the remaining BASELINE configs, measured on request.
WORKLOADS = {
    "anymal_trot": (3, 1, 2.0, False, 0, "ANYmal fly-trot 2 s (Anymal, combo C1, default Parameters, fixed durations)"),
    "anymal_trot_durations": (3, 1, 2.0, True, 0, "ANYmal fly-trot 2 s with optimised phase durations"),
    "hyq_gallop_durations": (2, 4, 2.4, True, 3, "HyQ gallop 2.4 s, terrains cycling Block / Gap / Stairs inside the batch, optimised "
                                                 "step durations (class declares Gap's curvature block)"),
    "biped_walk_durations": (1, 0, 2.4, True, 0, "biped walk 2.4 s with gait/phase-duration optimisation"),
    "monoped_hop": (0, 0, 2.4, False, 0, "monoped hop 2.4 s on flat ground"),
}
WL = WORKLOADS["anymal_trot"]


def wl_spec(M, goal=(1.0, 0.0, 0.0), index=None):
    """The workload's spec; BASELINE config 4 (terrain id 3 = Gap class) cycles Block / Gap / Stairs over the problem index."""
    terrain = WL[4]
    if terrain == 3 and index is not None:
        terrain = (1, 3, 2)[index % 3]
    return M.make_spec(WL[0], WL[1], WL[2], goal=goal, terrain=terrain, optimize_durations=WL[3])


def metric_name(name):
    return METRIC if name == "anymal_trot" else name + " constraint+Jacobian evals/s"
UNIT = "evals/s"
GOAL_SEED = 20260101 + 2      # seed = 20260101 + config index
X_VARIANTS = 16


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=4096, help="problems per GPU")
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the extra duration-optimised workload of the default run")
    ap.add_argument("--workload", default="anymal_trot", choices=sorted(WORKLOADS))
    a = ap.parse_args()
    global WL
    WL = WORKLOADS[a.workload]
    return a


def workload_config(st, batch, n_gpus):
    return {"workload": WL[5],
            "batch_per_gpu": batch, "global_batch": batch * n_gpus, "n_vars": st["n_vars"], "n_rows": st["n_rows"],
            "nnz": st["nnz"], "bytes_per_eval": st["bytes"], "x_variants": X_VARIANTS,
            "l2": "outputs written per step (%.0f MB) and x rotated over %d variants exceed the 126 MB L2"
                  % (batch * 8 * (st["n_rows"] + st["nnz"]) / 1e6, X_VARIANTS),
            "parallelism": "problem batch sharded per GPU, no data-path collective"}


class ClockSampler:
    """Polls NVML for SM clocks and throttle reasons while the timed regions run."""

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nv = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self._nv = None

    def _run(self):
        nv = self._nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20,
                 "hw_power_brake": 0x80}
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(self._h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.02)

    def start(self):
        if self._nv:
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()

    def stop(self):
        self._stop.set()
        if self._thr:
            self._thr.join()
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def x_scales(var_sets, mass):
    """Per-variable perturbation scale of the evaluation points, from the variable-set names."""
    import numpy as np
    mg = mass * 9.80665
    out = []
    for name, n in var_sets:
        sc = 0.05
        if name.startswith("base-ang") or name.startswith("ee-ang"):
            sc = 0.02
        elif name.startswith("ee-force"):
            sc = 0.05 * mg
        elif name.startswith("ee-torque"):
            sc = 0.5
        elif name.startswith("ee-schedule"):
            sc = 0.02
        out.append(np.full(n, sc))
    return np.concatenate(out)


def draw_goal(rng):
    return (rng.uniform(0.5, 2.5), rng.uniform(-0.3, 0.3), rng.uniform(-0.4, 0.4))


def make_inputs(T, batch, rank):
    """Structure class + per-problem specs (random goals) + X_VARIANTS perturbed x per problem."""
    import numpy as np
    rng = np.random.default_rng(GOAL_SEED + 7919 * rank)
    s0 = wl_spec(T)
    st = T.Structure(s0)
    specs = []
    x0 = np.empty((batch, st.n_vars))
    for i in range(batch):
        s = wl_spec(T, draw_goal(rng), index=i)
        specs.append(s)
        x0[i] = st.x0(s)
    scale = x_scales(st.var_sets(), s0.mass)
    xs = []
    for v in range(X_VARIANTS):
        xs.append(x0 + rng.uniform(-1.0, 1.0, size=x0.shape) * scale)
    return s0, st, specs, xs


def oracle_inputs(O, sample, variants):
    """The same distribution of evaluation points as make_inputs, built with the oracle only (reference arm and
    cpu_baseline): random goals, x = the goal's initial guess + U[-1,1] * scale.  Goals only move x0 and bounds, not g / J
    (SURVEY.md H7), so one oracle object graph per thread serves every problem of the sample."""
    import numpy as np
    rng = np.random.default_rng(GOAL_SEED)
    so = wl_spec(O)
    probe = O.Problem(so)
    x0 = np.empty((sample, probe.n_vars))
    for i in range(sample):
        x0[i] = O.Problem(wl_spec(O, draw_goal(rng))).x0()
    scale = x_scales(probe.var_sets(), so.mass)
    xs = [x0 + rng.uniform(-1.0, 1.0, size=x0.shape) * scale for _ in range(variants)]
    return so, probe, xs


def run_reference(args, rank, world):
    """The reference's CPU evaluation of the path on a bounded sample: the oracle port (the reference itself needs Eigen /
    ifopt / Ipopt, absent), OpenMP over problems on all host cores, one object graph per thread built before the timed
    region.  This arm imports nothing of the product: sizes, x0 and the pattern size come from the oracle."""
    if rank != 0:
        return
    from oracle import oracle_py as O
    cores = os.cpu_count() or 1
    sample = max(cores * 16, 32)
    so, probe, xs = oracle_inputs(O, sample, 4)
    pool = O.Pool(so, cores)
    nnz = pool.nnz
    import numpy as np
    g = np.empty((sample, pool.n_rows))
    v = np.empty((sample, nnz))
    for k in range(args.warmup):
        pool.eval(xs[k % len(xs)], g, v)
    t0 = time.perf_counter()
    for k in range(args.steps):
        pool.eval(xs[k % len(xs)], g, v)
    dt = time.perf_counter() - t0
    value = sample * args.steps / dt
    stinfo = {"n_vars": pool.n_vars, "n_rows": pool.n_rows, "nnz": nnz, "bytes": 8 * (pool.n_vars + pool.n_rows + nnz)}
    line = {"impl": "reference", "metric": metric_name(args.workload), "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(stinfo, args.batch, args.gpus),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": "%d problems per step (bounded sample of the %d-problem batch, same goal and x "
                                       "distribution), g + Jacobian values, OpenMP over problems, object graphs built before "
                                       "the timed region; the reference itself needs Eigen/ifopt/Ipopt (absent)"
                                       % (sample, args.batch)},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def cpu_baseline(n_rows, nnz):
    """Oracle port on the host cores, bounded sample (reported next to the GPU number, rank 0, N=1 only)."""
    import numpy as np
    from oracle import oracle_py as O
    cores = os.cpu_count() or 1
    sample = max(cores * 16, 32)
    so, probe, xs = oracle_inputs(O, sample, 2)
    pool = O.Pool(so, cores)
    assert pool.nnz == nnz and pool.n_rows == n_rows
    g = np.empty((sample, pool.n_rows))
    v = np.empty((sample, nnz))
    pool.eval(xs[0], g, v)
    reps, t0 = 0, time.perf_counter()
    while True:
        pool.eval(xs[reps % 2], g, v)
        reps += 1
        dt = time.perf_counter() - t0
        if dt > 10.0:
            break
    return {"value": sample * reps / dt, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": "%d problems x %d repetitions of the same workload (g + Jacobian values), OpenMP over problems, "
                      "one object graph per thread built before the timed region, %.1f s of CPU work" % (sample, reps, dt)}


def extra_workload(T, torch, name, batch, steps, device):
    """Device-resident evals/s and roofline fraction of another BASELINE workload (same timing method, fewer steps)."""
    global WL
    saved = WL
    WL = WORKLOADS[name]
    try:
        s0, st, specs, xs_host = make_inputs(T, batch, 0)
        b = T.Batch(st, batch, device=device)
        for i, s in enumerate(specs):
            b.set_problem(i, s)
        dev = torch.device("cuda", device)
        xs = [torch.from_numpy(x).to(dev) for x in xs_host[:4]]
        g = torch.empty((batch, st.n_rows), dtype=torch.float64, device=dev)
        vals = torch.empty((batch, st.nnz), dtype=torch.float64, device=dev)
        stream = torch.cuda.current_stream()
        for k in range(3):
            b.eval_device(batch, T.EVAL_BOTH, xs[k % 4].data_ptr(), g.data_ptr(), vals.data_ptr(), stream.cuda_stream)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for k in range(steps):
            b.eval_device(batch, T.EVAL_BOTH, xs[k % 4].data_ptr(), g.data_ptr(), vals.data_ptr(), stream.cuda_stream)
        e1.record(stream)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        assert bool(torch.isfinite(vals).all())
        try:
            peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
        except Exception:
            peak = 6650.0
        achieved = st.bytes_per_eval * batch / (ms * 1e-3) / 1e9
        return {name: {"workload": WL[5], "batch": batch, "steps": steps, "ms_per_step": ms, "value": batch / (ms * 1e-3), "unit": UNIT,
                       "n_vars": st.n_vars, "n_rows": st.n_rows, "nnz": st.nnz, "bytes_per_eval": st.bytes_per_eval,
                       "roofline_frac": achieved / peak, "achieved_gbs": achieved, "kernel": b.kernel_info}}
    finally:
        WL = saved


def run_ours(args, rank, world, local_rank):
    import numpy as np
    import torch
    import torch.distributed as dist
    import towr2025_b200 as T

    if not torch.cuda.is_available() or T.lib().towr_b200_device_count() < 1:
        raise SystemExit("bench.py: no CUDA device - the backend has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    # one rank per GPU: keep the rank's threads and its pinned staging buffers on the GPU's own NUMA node, so that N ranks
    # do not share one socket's memory bandwidth (single-GPU runs stay unbound: the CPU baseline wants every core)
    numa = {"node": T.device_numa_node(local_rank), "cpus_bound": 0}
    if world > 1:
        numa["cpus_bound"] = T.bind_thread_to_device_node(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    batch = args.batch
    s0, st, specs, xs_host = make_inputs(T, batch, rank)
    for v in range(X_VARIANTS):
        for i in range(0, batch, max(1, batch // 64)):      # clip a sample of problems to their own bounds
            lo, up = st.var_bounds(specs[i])
            xs_host[v][i] = np.clip(xs_host[v][i], lo, up)
    b = T.Batch(st, batch, device=local_rank)
    for i, s in enumerate(specs):
        b.set_problem(i, s)
    dev = torch.device("cuda", local_rank)
    xs = [torch.from_numpy(x).to(dev) for x in xs_host]
    g = torch.empty((batch, st.n_rows), dtype=torch.float64, device=dev)
    vals = torch.empty((batch, st.nnz), dtype=torch.float64, device=dev)
    stream = torch.cuda.current_stream()

    def step(k):
        b.eval_device(batch, T.EVAL_BOTH, xs[k % X_VARIANTS].data_ptr(), g.data_ptr(), vals.data_ptr(), stream.cuda_stream)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for k in range(args.warmup):
        step(k)
    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    launches0 = b.launches
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    ev[0].record(stream)
    for k in range(args.steps):
        step(k)
        ev[k + 1].record(stream)
    barrier()
    total_ms = ev[0].elapsed_time(ev[-1])
    launches = b.launches - launches0
    per_launch_ms = [ev[k].elapsed_time(ev[k + 1]) for k in range(args.steps)]
    assert bool(torch.isfinite(g).all()) and bool(torch.isfinite(vals).all())

    # ---- end to end through the C ABI with pinned HOST buffers: x goes up, g and the Jacobian values come back as full
    # n x nnz arrays.  The x-independent entries of every problem are fetched once at set-up
    # (towr_b200_batch_constant_values) into the pinned n x nnz array; each timed step moves x (H2D), g and the
    # x-dependent entries (D2H, TOWR_B200_EVAL_JAC_DYNAMIC: strided copies straight to their places in that array).  The
    # plain full transfer is timed next to it.
    hx = [torch.from_numpy(x).pin_memory() for x in xs_host[:2]]
    hg = torch.empty((batch, st.n_rows), dtype=torch.float64).pin_memory()
    hv = torch.empty((batch, st.nnz), dtype=torch.float64).pin_memory()
    nd, nd_moved = b.n_dynamic, b.dynamic_transfer_entries
    b.eval(hx[0].numpy(), T.EVAL_BOTH, g_out=hg.numpy(), jac_out=hv.numpy())
    barrier()
    t0 = time.perf_counter()
    for k in range(args.e2e_steps):
        b.eval(hx[k % 2].numpy(), T.EVAL_BOTH, g_out=hg.numpy(), jac_out=hv.numpy())
    barrier()
    e2e_full_s = time.perf_counter() - t0
    v_check = hv.numpy()[:: max(1, batch // 16)].copy()          # full-transfer result of the last step, a sample of rows
    hvn = hv.numpy()
    for i in range(batch):                                        # set-up: constants of every problem, once
        hvn[i] = b.constant_values(i)
    b.eval(hx[0].numpy(), T.EVAL_G | T.EVAL_JAC_DYNAMIC, g_out=hg.numpy(), jac_out=hvn)
    barrier()
    t0 = time.perf_counter()
    for k in range(args.e2e_steps):
        b.eval(hx[k % 2].numpy(), T.EVAL_G | T.EVAL_JAC_DYNAMIC, g_out=hg.numpy(), jac_out=hvn)
    barrier()
    e2e_s = time.perf_counter() - t0
    # both loops end on the same x: the merged arrays must be bit-identical to the full transfer
    assert np.array_equal(hvn[:: max(1, batch // 16)], v_check), "dynamic-only transfer differs from the full transfer"
    clocks = sampler.stop()
    other_workloads = {}
    if world == 1 and args.workload == "anymal_trot" and not args.no_extra:
        # one duration-optimised class in the same record (BASELINE config 2 / 4 regime: 8.5 x the stored entries),
        # device-resident only, smaller batch so that it stays a few seconds
        del hv, hvn, hg, vals, g
        torch.cuda.empty_cache()
        other_workloads = extra_workload(T, torch, "anymal_trot_durations", 2048, min(args.steps, 10), local_rank)

    t = torch.tensor([total_ms, e2e_s, e2e_full_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        # the only collective of the job: gather the solution vectors of all ranks (north_star "gather converged
        # solutions"); here the last evaluated x stands in for the converged iterate
        from towr2025_b200.sharding import gather_solutions
        sol = gather_solutions(xs[0], batch * world)
        assert sol.shape == (batch * world, st.n_vars)
    total_ms, e2e_s, e2e_full_s = float(t[0]), float(t[1]), float(t[2])
    if rank == 0:
        peaks, peak_src = None, "fallback"
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
            peak, peak_src = float(peaks["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            peak = 6650.0
        kern_ms = sorted(per_launch_ms)[len(per_launch_ms) // 2]
        avg_ms = sum(per_launch_ms) / len(per_launch_ms)
        achieved = st.bytes_per_eval * batch / (avg_ms * 1e-3) / 1e9
        # DRAM bytes per launch come from an ncu capture (never measured inside a bench run): profiles/traffic.json says
        # which capture (round tag, commit, kernel); it is only quoted for the workload it was taken on
        traffic, traffic_src = None, None
        try:
            tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
            if args.workload == tj.get("workload", "anymal_trot"):
                traffic = tj["dram_bytes_per_launch"]
                traffic_src = "ncu --set full capture %s (commit %s, %s)" % (tj.get("tag"), tj.get("commit", "?"), tj.get("kernel", "?"))
        except Exception:
            pass
        stinfo = {"n_vars": st.n_vars, "n_rows": st.n_rows, "nnz": st.nnz, "bytes": st.bytes_per_eval}
        line = {"metric": metric_name(args.workload), "value": world * batch * args.steps / (total_ms * 1e-3), "unit": UNIT, "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": workload_config(stinfo, batch, world),
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "traffic": traffic, "traffic_source": traffic_src, "kernel": b.kernel_info, "avg_launch_ms": avg_ms,
                             "median_launch_ms": kern_ms, "peak_source": peak_src,
                             "algorithmic_bytes_per_launch": st.bytes_per_eval * batch},
                "e2e": {"value": world * batch * args.e2e_steps / e2e_s, "unit": UNIT,
                        "h2d_bytes_per_step": batch * st.n_vars * 8, "d2h_bytes_per_step": batch * (st.n_rows + nd_moved) * 8,
                        "steps": args.e2e_steps, "ms_per_step": e2e_s / args.e2e_steps * 1e3,
                        "api": "towr_b200_batch_eval(EVAL_G | EVAL_JAC_DYNAMIC) (C ABI), pinned host buffers; result = full n x nnz "
                               "value arrays on the host (their constant entries filled once at set-up)",
                        "dynamic_entries": nd, "entries_copied": nd_moved, "stored_entries": st.nnz,
                        "d2h_gbs_per_gpu": batch * (st.n_rows + nd_moved) * 8 * args.e2e_steps / e2e_s / 1e9,
                        "host_placement": numa,
                        "full_transfer": {"value": world * batch * args.e2e_steps / e2e_full_s, "ms_per_step": e2e_full_s / args.e2e_steps * 1e3,
                                          "d2h_bytes_per_step": batch * (st.n_rows + st.nnz) * 8,
                                          "api": "towr_b200_batch_eval(EVAL_BOTH): every stored value crosses PCIe"}},
                "gpu_launches": launches, "clocks": clocks}
        if world == 1 and args.workload == "anymal_trot" and not args.no_extra:
            line["other_workloads"] = other_workloads
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(st.n_rows, st.nnz)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if world == 1 and args.gpus > 1:
        # launched without torchrun: re-exec under it so that one process drives each GPU
        import subprocess
        port = 29500 + os.getpid() % 1000
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(args.gpus),
               "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.abspath(__file__)] + sys.argv[1:]
        raise SystemExit(subprocess.call(cmd))
    run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
