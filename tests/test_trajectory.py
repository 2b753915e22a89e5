"""Trajectory sampling (SaveTrajectoryToCSV's sampling loop, towr/src/utils/save_data.cpp:54-119): the oracle's
restatement against known answers on the CPU, and the CUDA sampler (through the C ABI) against the oracle on the GPU."""
import os

import numpy as np
import pytest

from util import CASES, is_rt, mk, parity_ok, perturbed, pow_ok, to_oracle_spec

TRAJ_CASES = [c for c in CASES if c[0] in ("anymal_trot", "hyq_gallop_block", "biped_walk_stairs", "monoped_hop_torque",
                                           "biped_walk_durations", "anymal_trot_durations", "monoped_hop_stairs_durations")]


def test_oracle_trajectory_known_answers(product, oracle):
    """Rows at the initial guess: t = k dt by repeated addition, the first row is the initial state, the last sample
    of the base is the goal, force z is m g / E in stance, is_contact follows the gait."""
    T, O = product, oracle
    case = CASES[0]
    s = mk(T, case)
    p = O.Problem(to_oracle_spec(O, s))
    st = T.Structure(s)
    x0 = st.x0(s)
    dt = 0.01
    rows = p.sample_trajectory(x0, dt)
    assert rows.shape == (st.trajectory_samples(dt), st.trajectory_width()) == (201, 19 + 25 * 4)
    t = 0.0
    for k in range(rows.shape[0]):
        assert rows[k, 0] == t
        t += dt
    assert np.allclose(rows[0, 1:4], np.array(s.initial_base_lin[:3]), atol=1e-15)
    assert np.allclose(rows[-1, 1:3], [case[4][0], case[4][1]], atol=1e-9)
    # Hermite interpolation: at the sample that coincides with base node k the spline returns the node value
    n_vars_lin = 6 * 21
    assert np.allclose(rows[100, 1:4], x0[6 * 10:6 * 10 + 3], atol=1e-12)
    assert n_vars_lin == 126
    for e in range(4):
        c0 = 19 + 25 * e
        contact = rows[:, c0 + 24]
        assert set(np.unique(contact)) <= {0.0, 1.0}
        assert contact[0] == 1.0 and contact[-1] == 1.0          # the trot starts and ends in stance
        assert (rows[contact == 0.0, c0 + 18:c0 + 21] == 0.0).all()   # no force in swing (initial guess)
        assert np.allclose(rows[0, c0 + 20], s.mass * 9.80665 / 4)    # nlp_formulation.cc:303-307


def test_trajectory_sizes_and_errors(product):
    T = product
    st = T.Structure(mk(T, CASES[4]))
    assert st.trajectory_width() == 19 + 25
    assert st.trajectory_samples(0.1) == 25       # T = 2.4: 0, 0.1, ..., 2.4 (<= T + 1e-9)
    with pytest.raises(T.TowrB200Error):
        st.trajectory_samples(0.0)


def test_csv_format(product, oracle, tmp_path):
    """Header and number format of save_data.cpp:24-46,68-116 (std::fixed, 6 decimals, is_contact as 0/1)."""
    T, O = product, oracle
    s = mk(T, CASES[4])
    st = T.Structure(s)
    rows = O.Problem(to_oracle_spec(O, s)).sample_trajectory(st.x0(s), 0.1)
    f = os.path.join(tmp_path, "traj.csv")
    st.write_csv(rows, f)
    lines = open(f).read().strip().split("\n")
    assert len(lines) == 1 + rows.shape[0]
    hdr = lines[0].split(",")
    assert hdr[0] == "time" and hdr[1] == "base_pos_x" and hdr[10] == "base_euler_roll" and hdr[18] == "base_omegadot_z"
    assert hdr[19] == "ee_pos_x_0" and hdr[28] == "ee_euler_roll_0" and hdr[37] == "contact_force_x_0"
    assert hdr[40] == "contact_torque_x_0" and hdr[43] == "is_contact_phase_0" and len(hdr) == 44
    first = lines[1].split(",")
    assert first[0] == "0.000000" and first[43] in ("0", "1")
    assert first[3] == "%.6f" % rows[0, 3]


@pytest.mark.gpu
@pytest.mark.parametrize("case", TRAJ_CASES, ids=[c[0] for c in TRAJ_CASES])
def test_gpu_trajectory_matches_oracle(product, oracle, case):
    T, O = product, oracle
    n = 5
    s0 = mk(T, case)
    st = T.Structure(s0)
    rng = np.random.default_rng(11)
    specs = [mk(T, case, goal=(case[4][0] + rng.uniform(-0.2, 0.2), case[4][1], case[4][2])) for _ in range(n)]
    x = np.stack([perturbed(st.x0(s), 100 + i, (0.02 if is_rt(case) else 0.1) if i else 0.0) for i, s in enumerate(specs)])
    b = T.Batch(st, n)
    for dt in (0.01, 0.037):
        rows = b.sample_trajectory(x, dt)
        for i, s in enumerate(specs):
            po = O.Problem(to_oracle_spec(O, s))
            ref = po.sample_trajectory(x[i], dt)
            assert rows[i].shape == ref.shape
            assert np.array_equal(rows[i][:, 0], ref[:, 0])                       # sample times: bit-exact
            cc = [19 + 25 * e + 24 for e in range((st.trajectory_width() - 19) // 25)]
            assert np.array_equal(rows[i][:, cc], ref[:, cc])                     # is_contact: exact
            if is_rt(case):          # strict against the exact-power oracle, pow_ok against std::pow (tests/util.py)
                with O.pow_mode(1):
                    assert parity_ok(rows[i], po.sample_trajectory(x[i], dt))
                assert pow_ok(rows[i], ref)
            else:
                assert parity_ok(rows[i], ref)


@pytest.mark.gpu
@pytest.mark.parametrize("case", [TRAJ_CASES[0], TRAJ_CASES[-1]], ids=lambda c: c[0])
def test_gpu_sample_points_at_arbitrary_times(product, oracle, case):
    """towr_b200_batch_sample_points = SplineHolder::...->GetPoint(t) at any t: the rows for an unsorted list of times
    taken from the oracle's grid must be the oracle's rows for those times; times outside [0, T] are refused."""
    T, O = product, oracle
    s = mk(T, case)
    st = T.Structure(s)
    x = np.stack([perturbed(st.x0(s), 300 + i, 0.02 if is_rt(case) else 0.1) for i in range(3)])
    b = T.Batch(st, 3)
    po = O.Problem(to_oracle_spec(O, s))
    for i in range(3):
        ref = po.sample_trajectory(x[i], 0.037)
        pick = np.random.default_rng(i).permutation(ref.shape[0])[:17]
        rows = b.sample_points(x[i:i + 1], ref[pick, 0])[0]
        assert np.array_equal(rows[:, 0], ref[pick, 0])
        assert (pow_ok if is_rt(case) else parity_ok)(rows, ref[pick])
    with pytest.raises(T.TowrB200Error):
        b.sample_points(x[:1], np.array([0.1, case[3] + 0.5]))
    with pytest.raises(T.TowrB200Error):
        b.sample_points(x[:1], np.array([-0.1]))


@pytest.mark.gpu
def test_gpu_trajectory_full_batch_properties(product):
    """Batch 1024: identical x rows give identical trajectories; node interpolation holds on the device rows."""
    import torch
    T = product
    s = mk(T, CASES[0])
    st = T.Structure(s)
    n, dt = 1024, 0.01
    x0 = st.x0(s)
    x = np.tile(x0, (n, 1))
    x[1::2] = perturbed(x0, 5, 0.05)
    b = T.Batch(st, n)
    dx = torch.from_numpy(x).cuda()
    rows = torch.empty((n, st.trajectory_samples(dt), st.trajectory_width()), dtype=torch.float64, device="cuda")
    b.sample_trajectory_device(n, dt, dx.data_ptr(), rows.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    r = rows.cpu().numpy()
    assert np.isfinite(r).all()
    assert np.array_equal(r[0], r[2]) and np.array_equal(r[1], r[3]) and not np.array_equal(r[0], r[1])
    assert np.allclose(r[1, 100, 1:4], x[1, 60:63], atol=1e-12)          # base node 10 at t = 1.0
