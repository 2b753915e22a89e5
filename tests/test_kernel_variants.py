"""The experimental kernel variants (TOWR_B200_KERNEL = ring | streams | streamsring | split; built only by
`make -C towr2025_b200/csrc EXPERIMENTS=1` into libtowr_b200_experiments.so, because they measured slower than the default
lock-step kernel) compute the same products as the default kernel: values and g must be bit-identical on the same inputs, including a short last problem group."""
import os

import numpy as np
import pytest

from util import CASES, mk, perturbed

pytestmark = pytest.mark.gpu


def _eval_with(T, kernel, st, s, x):
    old = os.environ.get("TOWR_B200_KERNEL")
    try:
        if kernel:
            os.environ["TOWR_B200_KERNEL"] = kernel
        else:
            os.environ.pop("TOWR_B200_KERNEL", None)
        b = T.Batch(st, x.shape[0])        # the variant is chosen at batch creation
        b.set_all(s)
        g, v = b.eval(x)
        assert (b.status() == 0).all()
        return g, v
    finally:
        if old is None:
            os.environ.pop("TOWR_B200_KERNEL", None)
        else:
            os.environ["TOWR_B200_KERNEL"] = old


@pytest.mark.parametrize("kernel", ["ring", "streams", "streamsring"])
@pytest.mark.parametrize("name", ["anymal_trot", "hyq_gallop_block", "monoped_hop"])
def test_variant_matches_default_kernel(product, kernel, name):
    T = product
    case = [c for c in CASES if c[0] == name][0]
    s = mk(T, case)
    st = T.Structure(s)
    n = 301                                   # odd: the last group holds one problem; > 2 x 148 so CTAs loop
    x = np.stack([perturbed(st.x0(s), 500 + i, 0.05) for i in range(n)])
    g0, v0 = _eval_with(T, "lockstep", st, s, x)
    try:
        g1, v1 = _eval_with(T, kernel, st, s, x)
    except T.TowrB200Error as e:
        if e.code == -2:
            pytest.skip("variant does not fit this class: " + str(e))
        raise
    assert np.array_equal(g0, g1)
    assert np.array_equal(v0, v1)


@pytest.mark.parametrize("kernel", ["streamsring"])
def test_ring_variants_with_optimised_durations(product, kernel):
    T = product
    case = [c for c in CASES if c[0] == "biped_walk_durations"][0]
    s = mk(T, case)
    st = T.Structure(s)
    x = np.stack([perturbed(st.x0(s), 700 + i, 0.02) for i in range(9)])
    g0, v0 = _eval_with(T, "lockstep", st, s, x)
    try:
        g1, v1 = _eval_with(T, kernel, st, s, x)
    except T.TowrB200Error as e:
        if e.code == -2:
            pytest.skip("variant not built or does not fit this class: " + str(e))
        raise
    assert np.array_equal(g0, g1) and np.array_equal(v0, v1)


@pytest.mark.parametrize("name", ["anymal_trot", "hyq_gallop_block", "monoped_hop", "biped_walk_durations",
                                  "anymal_trot_durations", "anymal_trot_very_fine_dt"])
def test_split_kernel_matches_lockstep(product, name):
    """TOWR_B200_KERNEL=split (eval_kernel_split.cuh: producer / consumer warp groups, two G buffers, more and smaller
    segments): bit-identical g and values, status flags included (one problem gets a NaN in x)."""
    T = product
    case = [c for c in CASES if c[0] == name][0]
    s = mk(T, case)
    st = T.Structure(s)
    n = 301
    x = np.stack([perturbed(st.x0(s), 900 + i, 0.03) for i in range(n)])
    g0, v0 = _eval_with(T, "lockstep", st, s, x)
    try:
        g1, v1 = _eval_with(T, "split", st, s, x)
    except T.TowrB200Error as e:
        if e.code == -2:
            pytest.skip("variant does not fit this class: " + str(e))
        raise
    assert np.array_equal(g0, g1)
    assert np.array_equal(v0, v1)
    os.environ["TOWR_B200_KERNEL"] = "split"
    try:
        b = T.Batch(st, n)
        b.set_all(s)
        x[17, 5] = np.nan
        b.eval(x)
        bad = np.nonzero(b.status() & 2)[0]
        assert list(bad) == [17]
    finally:
        os.environ.pop("TOWR_B200_KERNEL", None)
