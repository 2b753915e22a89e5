#!/usr/bin/env python
"""Per-phase clocks of CTA 0 from the instrumented build (make -C towr2025_b200/csrc timing).

    python tools/phase_timing.py [batch]
Not a benchmark: the instrumented library is never the one bench.py or the tests load.
"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
# TIMING_CODE = <lanes><rt><streams><ring> of the instrumented variant (make -C towr2025_b200/csrc timing TIMING_CODE=...)
CODE = os.environ.get("TIMING_CODE", "2010")
os.environ["TOWR_B200_LIB"] = os.path.join(ROOT, "towr2025_b200", "libtowr_b200_timing_%s.so" % CODE)
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import towr2025_b200 as T  # noqa: E402


def main():
    batch = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
    s0 = T.make_spec(T.Anymal, T.C1, 2.0, optimize_durations=len(sys.argv) > 2 and sys.argv[2] == "durations")
    st = T.Structure(s0)
    b = T.Batch(st, batch, 0)
    b.set_all(s0)
    x = np.tile(st.x0(s0), (batch, 1))
    x += np.random.default_rng(0).uniform(-0.01, 0.01, x.shape)
    dx = torch.from_numpy(x).cuda()
    dg = torch.empty((batch, st.n_rows), dtype=torch.float64, device="cuda")
    dv = torch.empty((batch, st.nnz), dtype=torch.float64, device="cuda")
    L = T.lib()
    out = (C.c_ulonglong * 64)()
    for it in range(3):
        b.eval_device(batch, T.EVAL_BOTH, dx.data_ptr(), dg.data_ptr(), dv.data_ptr())
        torch.cuda.synchronize()
        L.towr_b200_debug_phase_clocks(out, 64)
    v = list(out)
    tot = sum(v)
    print("raw:", v)
    names = {0: "N work", 1: "N wait", 2: "C work", 3: "C wait"}
    print("total clocks of CTA 0 (thread 0 view): %d" % tot)
    for k in range(4):
        print("%-10s %9d %5.1f%%" % (names[k], v[k], 100.0 * v[k] / tot))
    s = 0
    while 8 + 6 * s < 64 and sum(v[8 + 6 * s: 14 + 6 * s]):
        w = v[8 + 6 * s: 14 + 6 * s]
        print("segment %d: L0 work %d wait %d | L1 work %d wait %d | B work %d wait %d  (%.1f%%)" %
              (s, w[0], w[1], w[2], w[3], w[4], w[5], 100.0 * sum(w) / tot))
        s += 1


if __name__ == "__main__":
    main()
