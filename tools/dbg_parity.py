import sys, os
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(),'tests'))
import numpy as np
import towr2025_b200 as T
from oracle import oracle_py as O
from util import CASES, perturbed, to_oracle_spec
import test_gpu_parity as tg
case = [c for c in CASES if c[0] == (sys.argv[1] if len(sys.argv) > 1 else 'biped_walk_stairs')][0]
st, specs, x = tg._problem_batch(T, O, case, 6, 3)
b = T.Batch(st, 6)
for i, s in enumerate(specs): b.set_problem(i, s)
g, v = b.eval(x)
rows, cols = st.pattern()
cs = st.con_sets()
for i, s in enumerate(specs):
    p = O.Problem(to_oracle_spec(O, s))
    go = p.eval_g(x[i]); vo = p.eval_jac(x[i])[2]
    bg = np.nonzero(np.abs(g[i]-go) > 1e-14+1e-12*np.abs(go))[0]
    bv = np.nonzero(np.abs(v[i]-vo) > 1e-14+1e-12*np.abs(vo))[0]
    print("problem", i, "bad g", len(bg), "bad J", len(bv), "terrain", s.terrain_id)
    for k in bg[:6]: print("   g row", k, g[i][k], go[k])
    for k in bv[:6]: print("   J", k, "row", rows[k], "col", cols[k], v[i][k], vo[k])
if len(sys.argv) > 2:
    off=0
    for n,r in cs:
        print(n, off, off+r); off+=r
