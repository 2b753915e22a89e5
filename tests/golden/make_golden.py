#!/usr/bin/env python
"""Generates tests/golden/*.npz from the ORACLE (not from the reference: hexb66/towr2025 cannot be compiled or
imported here — Eigen/ifopt/Ipopt absent — and holds no golden vectors of its own).  The fixtures
freeze the oracle's outputs so that drift of the oracle or of the GPU path is caught by both the CPU and the GPU
suites; they do not pin the oracle to the reference ("parity unpinned").

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import oracle_py as O   # noqa: E402
from util import CASES, mk, perturbed   # noqa: E402

GOLDEN = list(CASES[:5]) + [c for c in CASES if c[0] in ("monoped_hop_torque", "biped_walk_torque_baserom_block")]
for case in GOLDEN:
    name = case[0]
    p = O.Problem(mk(O, case))
    x0 = p.x0()
    x = perturbed(x0, 2026, 0.1)
    r, c, v = p.eval_jac(x)
    idx = np.linspace(0, len(v) - 1, 400).astype(np.int64)          # keep the fixtures small
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", name + ".npz"), x=x, g=p.eval_g(x), jac_idx=idx,
                        jac_rows=r[idx], jac_cols=c[idx], jac_vals=v[idx], nnz=len(v), sum_vals=v.sum(), sum_abs=np.abs(v).sum())
    print(name, p.n_vars, p.n_rows, len(v))
