"""CEM elite selection on the device (histogram-assisted exact top-k, csrc/qpmpc_kernels.cuh block_topk_hist) and BASELINE
config 3 at full size.  Elite indices must equal a stable ascending argsort of the costs (S:1075 | A:1115) exactly."""
import numpy as np
import pytest

import quadruped_pympc_b200 as qp
from quadruped_pympc_b200 import synthetic
from helpers import LANES, make_pair, seeded_noise, seeded_freqs, run_both, assert_tick_parity

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize('lanes', LANES)
def test_config3_full_size(built, lanes):
    """BASELINE config 3: HyQReal pace, CEM-MPPI with step-frequency adaptation, cubic, N = 50,000, 200 elites."""
    oracle, ctrl, _ = make_pair('hyqreal1', 50000, control_parametrization='cubic_spline', sampling_method='cem_mppi',
                                optimize_step_freq=True, lanes_per_sample=lanes)
    sc = synthetic.scenario('hyqreal1', 'pace', seed=3)
    best0 = np.zeros(oracle.P, np.float32)
    sigma = np.full(oracle.P, 3.0, np.float32)
    noise = seeded_noise(oracle, 3, sigma)
    freqs = seeded_freqs(oracle, 3)
    ref, got = run_both(oracle, ctrl, sc, best0, noise, sigma=sigma, freqs=freqs)
    assert len(got['elite_indices']) == 200
    assert_tick_parity(ref, got, cem=True, noise=noise)


def _tick_cem(ctrl, sc, sigma, optimize_swing=1):
    ctrl.with_newkey()
    P = ctrl.num_control_parameters
    out = ctrl.jitted_compute_control(sc['state'], sc['reference'], sc['contact_sequence'], np.zeros(P, np.float32),
                                      ctrl.master_key, sigma, sc['phase_signal'], sc['nominal_step_freq'], optimize_swing)
    return np.asarray(out[6]), ctrl.last_elite_indices, ctrl.last_best_index


@pytest.mark.parametrize('lanes', LANES)
@pytest.mark.parametrize('N,adaptive,elite_k', [(1337, False, 0), (10000, False, 0), (10000, True, 0), (50000, True, 0),
                                                (200000, False, 256), (300, True, 0), (7, False, 0)])
def test_elites_are_the_stable_argsort_of_the_costs(built, N, adaptive, elite_k, lanes):
    """Production mode (in-kernel Philox), several ticks per handle (the device histogram must come back clean)."""
    cfg = qp.SamplerConfig.for_robot('go2', sampling_method='cem_mppi', control_parametrization='linear_spline',
                                     num_parallel_computations=N, optimize_step_freq=adaptive, elite_k=elite_k,
                                     lanes_per_sample=lanes)
    ctrl = qp.Sampling_MPC(cfg)
    sc = synthetic.scenario('go2', 'trot', seed=5)
    sigma = np.full(ctrl.num_control_parameters, 3.0, np.float32)
    k = min(elite_k or (200 if adaptive else 10), N)
    for _ in range(3):
        costs, elite, best = _tick_cem(ctrl, sc, sigma)
        want = np.argsort(costs, kind='stable')[:k]
        np.testing.assert_array_equal(elite, want)
        assert best == want[0]


@pytest.mark.parametrize('lanes', LANES)
@pytest.mark.parametrize('adaptive', [False, True])
@pytest.mark.parametrize('N', [6000, 40000])
def test_heavy_ties_take_the_exact_slow_path(built, adaptive, lanes, N):
    """All-zero Q: every fixed-gait cost is exactly 0 and the adaptive ones take three values (A:500), so the threshold
    bin overflows the shared-memory candidate list and the selection must still be the stable argsort.  N = 40 000 runs
    the two-stage launch (tick kernel + multi-block tail kernel, csrc/qpmpc_tail.cuh)."""
    cfg = qp.SamplerConfig.for_robot('go2', sampling_method='cem_mppi', control_parametrization='zero_order',
                                     num_parallel_computations=N, optimize_step_freq=adaptive, lanes_per_sample=lanes,
                                     q_diag=(0.0,) * 12)
    ctrl = qp.Sampling_MPC(cfg)
    sc = synthetic.scenario('go2', 'trot', seed=9)
    sigma = np.full(ctrl.num_control_parameters, 3.0, np.float32)
    for _ in range(2):
        costs, elite, best = _tick_cem(ctrl, sc, sigma)
        assert len(np.unique(costs)) <= 3
        want = np.argsort(costs, kind='stable')[:len(elite)]
        np.testing.assert_array_equal(elite, want)
        assert best == want[0]


@pytest.mark.parametrize('lanes', LANES)
def test_moderate_ties_are_resolved_by_index_digits(built, lanes):
    """~N/3 equal costs with N/3 below the candidate capacity: refinement has to walk into the index digits."""
    N = 3000
    cfg = qp.SamplerConfig.for_robot('go2', sampling_method='cem_mppi', control_parametrization='cubic_spline',
                                     num_parallel_computations=N, optimize_step_freq=True, lanes_per_sample=lanes,
                                     q_diag=(0.0,) * 12)
    ctrl = qp.Sampling_MPC(cfg)
    sc = synthetic.scenario('go2', 'pace', seed=1)
    sigma = np.full(ctrl.num_control_parameters, 3.0, np.float32)
    costs, elite, best = _tick_cem(ctrl, sc, sigma)
    assert len(elite) == 200
    np.testing.assert_array_equal(elite, np.argsort(costs, kind='stable')[:200])


@pytest.mark.parametrize('lanes', LANES)
@pytest.mark.parametrize('N,robot,adaptive', [(30000, 'go2', False), (50000, 'hyqreal1', True), (120000, 'aliengo', True)])
def test_two_stage_launch_elites_and_merge(built, N, robot, adaptive, lanes):
    """From 192 block records on the reduction tail runs as a second, multi-block kernel (csrc/qpmpc_tail.cuh): slices of
    the costs filtered against the histogram threshold by 32 blocks, two-level record merge.  The elite set must still be
    the exact stable argsort of the device's costs, tick after tick (counters and histogram come back clean), and the
    MPPI update must equal the one-level merge of a host replay of the weights."""
    cfg = qp.SamplerConfig.for_robot(robot, sampling_method='cem_mppi', control_parametrization='cubic_spline',
                                     num_parallel_computations=N, optimize_step_freq=adaptive, lanes_per_sample=lanes)
    ctrl = qp.Sampling_MPC(cfg)
    sc = synthetic.scenario(robot, 'pace', seed=4)
    P = ctrl.num_control_parameters
    sigma = np.full(P, 3.0, np.float32)
    k = 200 if adaptive else 10
    for _ in range(3):
        ctrl.with_newkey()
        key = ctrl.master_key.copy()
        out = ctrl.jitted_compute_control(sc['state'], sc['reference'], sc['contact_sequence'], np.zeros(P, np.float32), key,
                                          sigma, sc['phase_signal'], sc['nominal_step_freq'], 1)
        costs = np.asarray(out[6])
        want = np.argsort(costs, kind='stable')[:k]
        np.testing.assert_array_equal(ctrl.last_elite_indices, want)
        assert ctrl.last_best_index == want[0]
        noise = ctrl.generate_noise(key, sigma)
        w = np.exp(-(costs.astype(np.float64) - float(costs.min())))
        best = (w[:, None] * noise.astype(np.float64)).sum(0) / w.sum()
        np.testing.assert_allclose(out[3], best, rtol=1e-4, atol=2e-5)


@pytest.mark.parametrize('method', ['mppi', 'cem_mppi'])
def test_two_stage_launch_inside_a_captured_graph(built, method):
    """io_mode='memcpy' replays a captured CUDA graph (H2D, kernels, D2H).  At 30 000 samples the graph holds the (tick,
    tail) pair with its programmatic dependency; the replayed ticks must equal the direct-I/O handle bit for bit."""
    outs = []
    sc = synthetic.scenario('go2', 'trot', seed=8)
    for io in ('direct', 'memcpy'):
        cfg = qp.SamplerConfig.for_robot('go2', sampling_method=method, control_parametrization='cubic_spline',
                                         num_parallel_computations=30000, io_mode=io)
        ctrl = qp.Sampling_MPC(cfg)
        P = ctrl.num_control_parameters
        best, res = np.zeros(P, np.float32), []
        for _ in range(3):
            ctrl.with_newkey()
            args = (np.full(P, 3.0, np.float32),) if method == 'cem_mppi' else ()
            out = ctrl.jitted_compute_control(sc['state'], sc['reference'], sc['contact_sequence'], best, ctrl.master_key,
                                              *args, sc['phase_signal'], sc['nominal_step_freq'], 0)
            best = out[3]
            res.append([np.asarray(x) for x in out] + [ctrl.last_best_index])
        outs.append(res)
    for a, b in zip(*outs):
        for x, y in zip(a, b):
            np.testing.assert_array_equal(x, y)
