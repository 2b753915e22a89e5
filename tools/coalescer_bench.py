#!/usr/bin/env python
"""ms per solver iteration through the callback coalescer on one B200 (second half of BASELINE.json's metric; Ipopt itself
is absent, so stub solver threads stand in: tests/cpp/coalescer.cc `bench`).

    python tools/coalescer_bench.py [out.json]        one JSON line per thread count (64, 256, 1024)"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from test_cpp_surface import build  # noqa: E402


def main():
    exe = build("coalescer")
    out = open(sys.argv[1], "w") if len(sys.argv) > 1 else None
    for n, iters in ((64, 40), (256, 30), (1024, 12)):
        r = subprocess.run([exe, "bench", str(n), str(iters)], capture_output=True, text=True, timeout=600)
        line = [l for l in r.stdout.splitlines() if l.startswith("{")]
        text = line[-1] if line else '{"solver_threads": %d, "failed": %r}' % (n, (r.stdout + r.stderr)[-300:])
        print(text, flush=True)
        if out:
            out.write(text + "\n")


if __name__ == "__main__":
    main()
