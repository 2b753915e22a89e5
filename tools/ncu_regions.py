#!/usr/bin/env python
"""Split the kernel's SASS at barriers and print, per region, instruction share, stall-sample share, the stall
reasons and the shared-memory wavefront excess (bank conflicts).

    python tools/ncu_regions.py gpurun_out/prof.ncu-rep
"""
import csv
import io
import subprocess
import sys


def main():
    txt = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    h = rows[1]
    rows = rows[2:]
    stall_cols = [(i, x) for i, x in enumerate(h) if x.startswith("stall_") and "Not Issued" not in x]
    wf, wfi = h.index("L1 Wavefronts Shared"), h.index("L1 Wavefronts Shared Ideal")

    def num(r, i):
        try:
            return int(r[i])
        except (ValueError, IndexError):
            return 0
    regs = []
    cur = None
    for i, r in enumerate(rows):
        if cur is None:
            cur = dict(start=i, inst=0, samp=0, st={}, wf=0, wfi=0)
        cur["inst"] += num(r, 5)
        cur["samp"] += num(r, 2)
        cur["wf"] += num(r, wf)
        cur["wfi"] += num(r, wfi)
        for c, name in stall_cols:
            cur["st"][name] = cur["st"].get(name, 0) + num(r, c)
        s = r[1]
        if "BAR.SYNC" in s or s.strip().startswith("EXIT"):
            cur["end"] = i
            regs.append(cur)
            cur = None
    if cur:
        cur["end"] = len(rows) - 1
        regs.append(cur)
    ti = sum(r["inst"] for r in regs) or 1
    ts = sum(r["samp"] for r in regs) or 1
    print("total warp instructions %d, samples %d" % (ti, ts))
    for k, r in enumerate(regs):
        top = sorted(r["st"].items(), key=lambda x: -x[1])[:5]
        print("region %d sass[%d:%d] inst %.1f%% samples %.1f%% smem wavefronts %d (ideal %d) | %s" %
              (k, r["start"], r["end"], 100.0 * r["inst"] / ti, 100.0 * r["samp"] / ts, r["wf"], r["wfi"],
               ", ".join("%s %.1f%%" % (a.replace("stall_", ""), 100.0 * b / ts) for a, b in top)))


if __name__ == "__main__":
    main()
