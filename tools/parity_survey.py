#!/usr/bin/env python
"""Per-block parity survey: worst |a-b| / (1e-14 + 1e-12 |b|) and entries outside the band, for every Jacobian block
(constraint family x variable family) and every constraint family's g rows, over all test classes and 8 x per class.

    python tools/parity_survey.py            host emulation of the kernel arithmetic (tests/emu) against the oracle
    python tools/parity_survey.py gpu        the CUDA path through the C ABI against the oracle (needs a GPU)

TEST INFRASTRUCTURE: the table in DESIGN.md section 5 and profiles/r02_parity_survey*.txt come from here."""
import collections
import ctypes as C
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import oracle_py as O  # noqa: E402
import towr2025_b200 as T  # noqa: E402
from util import ATOL, CASES, RTOL, is_rt, mk, perturbed  # noqa: E402

GPU = len(sys.argv) > 1 and sys.argv[1] == "gpu"
SEEDS = range(20, 28)


def block_ids(sets, idx):
    edges = np.cumsum([0] + [k for _, k in sets])
    return np.searchsorted(edges, idx, side="right") - 1


def short(n):
    return n.split("_")[0].rstrip("0123456789").replace("-ee-force", "").replace("-ee-motion", "").replace("-ee-torque", "")


def main():
    if not GPU:
        d = os.path.join(ROOT, "tests", "emu")
        subprocess.check_call(["make", "-C", d], stdout=subprocess.DEVNULL)
        E = C.CDLL(os.path.join(d, "_build", "libemu.so"))
        DP = C.POINTER(C.c_double)
    worst = collections.defaultdict(lambda: [0.0, 0, 0])
    gw = collections.defaultdict(lambda: [0.0, 0, 0])
    for case in CASES:
        s = mk(T, case)
        st = T.Structure(s)
        p = O.Problem(mk(O, case))
        cs, vs = st.con_sets(), st.var_sets()
        rows, cols = st.pattern()
        keys = np.array([short(cs[a][0]) + " x " + short(vs[b][0]) for a, b in zip(block_ids(cs, rows), block_ids(vs, cols))])
        gk = np.array([short(cs[a][0]) for a in block_ids(cs, np.arange(st.n_rows))])
        xs = np.stack([perturbed(st.x0(s), seed, (0.01, 0.03, 0.05, 0.1)[seed % 4] if is_rt(case) else (0.03, 0.1, 0.2, 0.3)[seed % 4])
                       for seed in SEEDS])
        if GPU:
            b = T.Batch(st, len(xs))
            b.set_all(s)
            gs, vsl = b.eval(xs)
        for i, x in enumerate(xs):
            if GPU:
                g, v = gs[i], vsl[i]
            else:
                g, v = np.full(st.n_rows, np.nan), np.full(st.nnz, np.nan)
                assert E.emu_eval(C.byref(s), x.ctypes.data_as(DP), g.ctypes.data_as(DP), v.ctypes.data_as(DP)) == 0
            vo, go = p.eval_jac(x)[2], p.eval_g(x)
            r = np.abs(v - vo) / (ATOL + RTOL * np.abs(vo))
            rg = np.abs(g - go) / (ATOL + RTOL * np.abs(go))
            for k in np.unique(keys):
                m = keys == k
                w = worst[k]
                w[0] = max(w[0], float(r[m].max())); w[1] += int((r[m] > 1).sum()); w[2] += int(m.sum())
            for k in np.unique(gk):
                m = gk == k
                w = gw[k]
                w[0] = max(w[0], float(rg[m].max())); w[1] += int((rg[m] > 1).sum()); w[2] += int(m.sum())
    print("# %s against the oracle (std::pow, the reference's arithmetic): %d classes x %d x"
          % ("CUDA path through the C ABI" if GPU else "host emulation of the kernel arithmetic", len(CASES), len(SEEDS)))
    print("# Jacobian blocks: worst |a-b|/(1e-14+1e-12|b|), entries outside the band / entries compared")
    for k, w in sorted(worst.items(), key=lambda kv: -kv[1][0]):
        print("  %-40s %8.3g  %d / %d" % (k, w[0], w[1], w[2]))
    print("# g rows")
    for k, w in sorted(gw.items(), key=lambda kv: -kv[1][0]):
        print("  %-40s %8.3g  %d / %d" % (k, w[0], w[1], w[2]))


if __name__ == "__main__":
    main()
