"""towr_b200_eval_cost_static (include/towr_b200.h): the allocation-free variant of eval_f / eval_grad_f whose work tables
live in a file-scope device array of the library.  No counterpart in the reference.
Collected last on purpose (file name).

Checked through the C ABI against the existing entry point (towr_b200_batch_eval_cost) within the suite's tolerance, on
two calls in a row that use DIFFERENT structure classes, so the second call must overwrite the first call's tables."""
import numpy as np
import pytest

from util import perturbed, tol_ratio

pytestmark = pytest.mark.gpu


def _with_costs(T, robot, combo, total_time, optimise, sampled=False):
    s = T.make_spec(robot, combo, total_time, optimize_durations=optimise)
    s.n_costs = 2                                  # as test/hopper_example.cc:147-148
    s.cost_ids[0], s.cost_weights[0] = 0, 1e-3     # ForcesCostID
    s.cost_ids[1], s.cost_weights[1] = 1, 1e-2     # EEMotionCostID
    if sampled:                                    # the time-discretised terms: their tables live in the static array too
        s.n_costs = 4
        s.cost_ids[2], s.cost_weights[2] = 2, 1e-3     # EnergyCostID
        s.cost_ids[3], s.cost_weights[3] = 3, 0.5      # AngMomCostID
        s.enable_swing_ee_base_pos_tracking = 1        # EEBasePosCost, reference positions of the class spec
        s.swing_ee_base_pos_tracking_weight = 0.3
        s.initial_base_ang[2] = 0.2
    return s


def test_static_cost_matches_batch_entry_point_twice(product):
    import torch
    T = product
    stream = torch.cuda.current_stream().cuda_stream
    for robot, combo, total_time, optimise, n, sampled in ((T.Anymal, T.C1, 2.0, False, 33, False),
                                                           (T.Monoped, T.C0, 2.4, True, 7, False),
                                                           (T.Biped, T.C0, 2.4, True, 9, True),
                                                           (T.Anymal, T.C1, 2.0, False, 5, True)):
        s = _with_costs(T, robot, combo, total_time, optimise, sampled)
        st = T.Structure(s)
        x = np.stack([perturbed(st.x0(s), 40 + i, 0.05) for i in range(n)])
        b = T.Batch(st, n)
        b.set_all(s)
        f_ref, grad_ref = b.eval_cost(x)           # existing entry point (cudaMalloc'ed tables)
        assert np.abs(f_ref).max() > 0 and np.abs(grad_ref).max() > 0
        dx = torch.from_numpy(x).cuda()
        df = torch.full((n,), float("nan"), dtype=torch.float64, device="cuda")
        dg = torch.full((n, st.n_vars), float("nan"), dtype=torch.float64, device="cuda")
        torch.cuda.synchronize()
        st.eval_cost_static(n, 3, dx.data_ptr(), df.data_ptr(), dg.data_ptr(), device=0, stream=stream)
        torch.cuda.synchronize()
        assert tol_ratio(df.cpu().numpy(), f_ref) <= 1.0
        assert tol_ratio(dg.cpu().numpy(), grad_ref) <= 1.0
        # f alone and the gradient alone
        df2 = torch.zeros_like(df)
        st.eval_cost_static(n, 1, dx.data_ptr(), df2.data_ptr(), 0, device=0, stream=stream)
        torch.cuda.synchronize()
        assert tol_ratio(df2.cpu().numpy(), f_ref) <= 1.0


def test_static_cost_rejects_oversized_and_bad_arguments(product):
    import torch
    T = product
    s = _with_costs(T, T.Anymal, T.C1, 2.0, False)
    st = T.Structure(s)
    dx = torch.zeros((1, st.n_vars), dtype=torch.float64, device="cuda")
    df = torch.zeros((1,), dtype=torch.float64, device="cuda")
    with pytest.raises(T.TowrB200Error):
        st.eval_cost_static(1, 3, dx.data_ptr(), df.data_ptr(), 0)          # gradient asked for, no buffer
    with pytest.raises(T.TowrB200Error):
        st.eval_cost_static(1, 7, dx.data_ptr(), df.data_ptr(), 0)
    # a class beyond TOWR_B200_STATIC_COST_MAX_VARS (4096 variables): a long, finely split ANYmal walk
    big = T.make_spec(T.Anymal, T.C0, 8.0)
    big.duration_base_polynomial = 0.01
    big.n_costs = 1
    big.cost_ids[0], big.cost_weights[0] = 0, 1.0
    try:
        stb = T.Structure(big)
    except T.TowrB200Error:
        pytest.skip("the oversized class cannot be built")
    assert stb.n_vars > 4096
    dxb = torch.zeros((1, stb.n_vars), dtype=torch.float64, device="cuda")
    dgb = torch.zeros((1, stb.n_vars), dtype=torch.float64, device="cuda")
    with pytest.raises(T.TowrB200Error) as e:
        stb.eval_cost_static(1, 3, dxb.data_ptr(), df.data_ptr(), dgb.data_ptr())
    assert e.value.code == -1 and "TOWR_B200_STATIC_COST_MAX_VARS" in str(e.value)
