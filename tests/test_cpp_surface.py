"""include/towr_b200_ifopt.hpp: the C++ mirror of the ifopt component surface over the C ABI."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "tests", "cpp", "_build", "ifopt_surface")


def build(name="ifopt_surface"):
    import towr2025_b200 as T
    T.build()
    exe = os.path.join(os.path.dirname(EXE), name)
    os.makedirs(os.path.dirname(EXE), exist_ok=True)
    src = os.path.join(ROOT, "tests", "cpp", name + ".cc")
    lib_dir = os.path.join(ROOT, "towr2025_b200")
    hdrs = [os.path.join(ROOT, "include", h) for h in ("towr_b200_ifopt.hpp", "towr_b200_coalescer.hpp", "towr_b200.h",
                                                       "towr_b200_formulation.hpp", "towr_b200_spline_holder.hpp")]
    newest = max([os.path.getmtime(src), os.path.getmtime(os.path.join(lib_dir, "libtowr_b200.so"))] +
                 [os.path.getmtime(h) for h in hdrs])
    if not os.path.exists(exe) or os.path.getmtime(exe) < newest:
        from oracle import oracle_py as O      # TEST INFRASTRUCTURE: the coalescer test checks values against the oracle
        O.lib()
        orc_dir = os.path.join(ROOT, "oracle", "_build")
        subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-pthread", "-I", os.path.join(ROOT, "include"), src,
                               "-o", exe, "-L", lib_dir, "-ltowr_b200", "-Wl,-rpath," + lib_dir,
                               "-L", orc_dir, "-loracle", "-Wl,-rpath," + orc_dir, "-fopenmp"])
    return exe


def run(*args):
    out = subprocess.run([build(), *args], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    return {l.split()[0]: l.split()[1:] for l in out.stdout.splitlines() if l.strip()}


@pytest.mark.parametrize("mode", ["fixed", "durations"])
def test_surface_host_side(product, oracle, mode):
    T, O = product, oracle
    r = run("host", mode)
    opt = mode == "durations"
    p = O.Problem(O.make_spec(O.Anymal, 1, 2.0, goal=(0.7, 0.0, 0.0), optimize_durations=opt))
    st = T.Structure(T.make_spec(T.Anymal, T.C1, 2.0, optimize_durations=opt))
    assert [int(v) for v in r["sizes"]] == [st.n_vars, st.n_rows, st.nnz]
    assert float(r["x0sum"][0]) == pytest.approx(float(np.sum(p.x0())), rel=1e-13)
    assert "ok" in r


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["fixed", "durations"])
def test_surface_batched_eval(product, oracle, mode):
    T, O = product, oracle
    r = run("gpu", mode)
    opt = mode == "durations"
    p = O.Problem(O.make_spec(O.Anymal, 1, 2.0, goal=(1.5, 0.0, 0.0), optimize_durations=opt))
    x = p.x0()
    x = x + 0.01 * np.sin(0.37 * np.arange(len(x)))
    g = p.eval_g(x)
    rows, cols, v = p.eval_jac(x)
    assert float(r["gsum"][0]) == pytest.approx(float(g.sum()), rel=1e-9, abs=1e-9)
    assert float(r["gsum"][2]) == pytest.approx(float(v.sum()), rel=1e-9, abs=1e-7)
    dyn = dict(p.con_sets())["dynamic"]
    assert int(r["dynamic"][1]) == dyn
    off = 0
    for name, n in p.con_sets():
        if name == "dynamic":
            break
        off += n
    m = (rows >= off) & (rows < off + dyn)
    assert int(r["dynamic"][3]) == int(m.sum())
    assert float(r["dynamic"][5]) == pytest.approx(float(v[m].sum()), rel=1e-9, abs=1e-7)
    assert float(r["dynamic"][7]) == pytest.approx(float(g[off]), rel=1e-9, abs=1e-12)
    # eval_f / eval_grad_f and the trajectory rows (costs as in hopper_example.cc:147-148)
    so = O.make_spec(O.Anymal, 1, 2.0, goal=(1.5, 0.0, 0.0), optimize_durations=opt)
    so.n_costs = 2
    so.cost_ids[0], so.cost_weights[0] = 0, 1e-9
    so.cost_ids[1], so.cost_weights[1] = 1, 1e-4
    pc = O.Problem(so)
    assert float(r["cost"][0]) == pytest.approx(pc.eval_f(x), rel=1e-10, abs=1e-14)
    assert float(r["cost"][2]) == pytest.approx(float(pc.eval_grad_f(x).sum()), rel=1e-9, abs=1e-12)
    traj = pc.sample_trajectory(x, 0.05)
    assert [int(r["trajectory"][0]), int(r["trajectory"][1])] == list(traj.shape)
    assert float(r["trajectory"][2]) == pytest.approx(float(traj.sum()), rel=1e-9, abs=1e-7)


def test_callback_coalescer_policy():
    """Solver threads with diverging iteration / line-search counts: every request answered, no deadlock when
    solvers finish early, and far fewer launches than requests (stub evaluator, no device needed)."""
    out = subprocess.run([build("coalescer"), "host"], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "ok" in out.stdout


@pytest.mark.gpu
def test_callback_coalescer_on_gpu():
    out = subprocess.run([build("coalescer"), "gpu"], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "ok" in out.stdout


def test_callback_coalescer_survives_failures():
    """An evaluator that throws and a staging allocation that fails reach the waiting solver threads as exceptions; nobody
    stays blocked (the timeout is the deadlock detector)."""
    out = subprocess.run([build("coalescer"), "throw"], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "ok" in out.stdout


@pytest.mark.gpu
def test_callback_coalescer_values_match_oracle_on_gpu():
    """GPU leg: every g row and every stored Jacobian value each solver thread receives is compared with the oracle."""
    out = subprocess.run([build("coalescer"), "gpu"], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    line = [l for l in out.stdout.splitlines() if l.startswith("oracle_checks")][0].split()
    assert int(line[1]) > 100 and float(line[3]) <= 1.0 and float(line[5]) <= 4.0


def _hopper(mode):
    out = subprocess.run([build("hopper_surface"), mode], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "ok" in out.stdout
    return {l.split()[0]: l.split()[1:] for l in out.stdout.splitlines() if l.strip()}


def test_hopper_example_through_reference_types_host(product, oracle):
    """hopper_example.cc:93-150 stated with the reference's member names (towr_b200_formulation.hpp), ToSpec, the ifopt
    mirror: sizes, x0, all bounds and the Jacobian pattern equal the oracle's entry by entry; ToSpec throws what the
    reference throws (std::out_of_range without stance positions) and refuses what is not built."""
    r = _hopper("host")
    assert float(r["host_checks"][5]) <= 1.0


@pytest.mark.gpu
def test_hopper_example_through_reference_types_gpu(product, oracle):
    """The same problem evaluated: g, every Jacobian value, eval_f / eval_grad_f and the SplineHolder view (GetPoint of
    base / ee-motion / force / torque splines, IsContactPhase) against the oracle, entry by entry."""
    r = _hopper("gpu")["gpu_checks"]
    vals = dict(zip(r[0::2], r[1::2]))
    assert float(vals["g"]) <= 1.0 and float(vals["jac"]) <= 4.0 and float(vals["cost"]) <= 1.0
    assert float(vals["splines"]) <= 10.0      # the oracle's trajectory rows use std::pow (pow_ok band, tests/util.py)
    assert int(vals["contact_mismatch"]) == 0 and int(vals["samples"]) == 24 and int(vals["phases"]) == 13
    assert float(vals["duration_sum"]) == pytest.approx(4.7, abs=1e-12) and int(vals["force_v_threw"]) == 1
