#!/usr/bin/env python
"""Summarise an ncu report (run here, no GPU needed) into profiles/: key metrics of the evaluation kernel,
per-launch DRAM traffic (profiles/traffic.json, read by bench.py) and the launch list shares.

    python tools/ncu_summary.py gpurun_out/prof_r01.ncu-rep gpurun_out/launches_r01.csv r01 [commit of the captured build]
"""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__inst_executed.sum", "sm__inst_executed_pipe_fp64.sum.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem",
        "launch__occupancy_limit_registers", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum",
        "lts__t_sectors_srcunit_tex_op_read.sum", "lts__t_sectors_srcunit_tex_op_write.sum",
        "smsp__sass_inst_executed_op_shared.sum", "smsp__sass_inst_executed_op_global_st.sum",
        "smsp__sass_inst_executed_op_global_ld.sum"]


def main():
    rep, launches, tag = sys.argv[1], sys.argv[2], sys.argv[3]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, vals = rows[0], rows[1], rows[2]
    out = ["# ncu summary %s — %s" % (tag, vals[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "kernel"), "",
           "Source: `ncu --set full --clock-control none --import-source on -k regex:EvalKernel -s 3 -c 1` on "
           "`python bench.py --steps 3 --warmup 3 --e2e-steps 1 --no-cpu-baseline --no-extra` (B200, batch 4096 ANYmal trot; "
           "tools/run_round_gpu.sh).", "",
           "| metric | value | unit |", "|---|---|---|"]
    got = {}
    for k in KEYS:
        if k in hdr:
            i = hdr.index(k)
            got[k] = vals[i]
            out.append("| %s | %s | %s |" % (k, vals[i], units[i]))
    stalls = []
    for i, h in enumerate(hdr):
        if "issue_stalled" in h and h.endswith("_per_issue_active.ratio"):
            try:
                stalls.append((float(vals[i]), h))
            except ValueError:
                pass
    out += ["", "Top warp stall reasons (warps stalled per issue-active cycle):", ""]
    for v, h in sorted(stalls, reverse=True)[:6]:
        out.append("* %.2f %s" % (v, h.replace("smsp__average_warp_latency_issue_stalled_", "").replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", "")))

    def to_bytes(k):
        i = hdr.index(k)
        mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[units[i]]
        return float(vals[i]) * mult
    traffic = to_bytes("dram__bytes_read.sum") + to_bytes("dram__bytes_write.sum")
    out += ["", "DRAM traffic per launch: %.1f MB (read %.1f + write %.1f); algorithmic bytes per launch 909.8 MB "
            "(8*(1042+2732+23992) B x 4096)." % (traffic / 1e6, to_bytes("dram__bytes_read.sum") / 1e6,
                                                   to_bytes("dram__bytes_write.sum") / 1e6)]
    # launch list shares
    if os.path.exists(launches):
        lr = [r for r in csv.reader(open(launches)) if len(r) > 10 and r[0].isdigit()]
        tot, per = 0.0, {}
        for r in lr:
            name, ns = r[4].split("(")[0][-60:], float(r[-1])
            per[name] = per.get(name, [0, 0.0])
            per[name][0] += 1
            per[name][1] += ns
            tot += ns
        out += ["", "Launch list (`ncu --metrics gpu__time_duration.sum`, cold-cache and serialised: compare shares):", "",
                "| kernel | launches | total us | share |", "|---|---|---|---|"]
        for name, (n, ns) in sorted(per.items(), key=lambda kv: -kv[1][1])[:8]:
            out.append("| %s | %d | %.1f | %.1f%% |" % (name, n, ns / 1e3, 100 * ns / tot))
        out.append("")
        out.append("The torch elementwise/reduce kernels are bench.py's own `isfinite` sanity check after the timed "
                   "region and its input upload; inside the timed region only EvalKernel launches.")
    open(os.path.join(ROOT, "profiles", "%s_ncu_summary.md" % tag), "w").write("\n".join(out) + "\n")
    commit = sys.argv[4] if len(sys.argv) > 4 else subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
    json.dump({"dram_bytes_per_launch": traffic, "source": os.path.basename(rep), "tag": tag, "commit": commit, "workload": "anymal_trot",
               "kernel": vals[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "EvalKernel"},
              open(os.path.join(ROOT, "profiles", "traffic.json"), "w"))
    print("\n".join(out))


if __name__ == "__main__":
    main()
