"""Random formulations (tools/fuzz_parity.py): hand-made phase timings, constraint orders, sampling steps, terrains,
EELinearConstraint definitions and costs that no named case covers, against the oracle.  A handful per run here; the tool
runs hundreds (DESIGN.md section 5 lists what it found)."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run(*args):
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "fuzz_parity.py"), *args], capture_output=True, text=True,
                         timeout=900)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert "match the oracle" in out.stdout, out.stdout[-2000:]
    return out.stdout


def test_random_formulations_in_the_host_emulation(product, oracle, emu):
    run("30", "11")


def test_random_formulations_through_both_oracle_legs(product, oracle):
    """oracle/*.cc against oracle/oracle_np.py on formulations no named case covers (a few: the numpy leg is slow)."""
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "fuzz_parity.py"), "cross", "4", "13"], capture_output=True,
                         text=True, timeout=900)
    assert out.returncode == 0 and "agree on 4 random formulations" in out.stdout, out.stdout[-2000:] + out.stderr[-2000:]


@pytest.mark.gpu
def test_random_formulations_on_the_gpu(product, oracle):
    run("gpu", "24", "12")
