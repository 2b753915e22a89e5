#!/usr/bin/env python
"""Per-source-line instruction / stall-sample totals from an ncu report captured with --import-source on.

    python tools/ncu_lines.py gpurun_out/prof.ncu-rep [top_n]
"""
import csv
import io
import os
import subprocess
import sys


def main():
    rep = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"],
                         capture_output=True, text=True).stdout
    fname = None
    lines = []
    for r in csv.reader(io.StringIO(txt)):
        if not r:
            continue
        if r[0] == "File Path":
            fname = os.path.basename(r[1])
        elif r[0].isdigit() and len(r) > 8:
            try:
                lines.append((fname, int(r[0]), r[1].strip(), int(r[4] or 0), int(r[7] or 0)))
            except ValueError:
                pass
    ti = sum(l[4] for l in lines) or 1
    ts = sum(l[3] for l in lines) or 1
    print("total warp instructions %d, samples %d" % (ti, ts))
    per_file = {}
    for f, n, s, sm, ins in lines:
        a = per_file.setdefault(f, [0, 0])
        a[0] += ins
        a[1] += sm
    for f, (i, s) in per_file.items():
        print("%-22s inst %5.1f%%  samples %5.1f%%" % (f, 100.0 * i / ti, 100.0 * s / ts))
    print("--- top lines by instructions")
    for f, n, s, sm, ins in sorted(lines, key=lambda l: -l[4])[:top]:
        print("%5.1f%% i %5.1f%% s  %s:%d  %s" % (100.0 * ins / ti, 100.0 * sm / ts, f, n, s[:90]))
    print("--- top lines by samples")
    for f, n, s, sm, ins in sorted(lines, key=lambda l: -l[3])[:top]:
        print("%5.1f%% i %5.1f%% s  %s:%d  %s" % (100.0 * ins / ti, 100.0 * sm / ts, f, n, s[:90]))


if __name__ == "__main__":
    main()
