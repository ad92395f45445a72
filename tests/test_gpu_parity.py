"""GPU parity tests proper: the CUDA path (through the C ABI) against the oracle on identical seeded inputs.

Tolerances are BASELINE.json's: sample costs rel 1e-5; returned GRFs / knots rel 1e-4; argmin / elite indices exact.
"""
import numpy as np
import pytest

from quadruped_pympc_b200 import synthetic
from helpers import (PARAMS, METHODS, LANES, make_pair, seeded_noise, seeded_freqs, run_both, assert_tick_parity,
                     COST_RTOL)

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize('lanes', LANES)
@pytest.mark.parametrize('adaptive', [False, True])
@pytest.mark.parametrize('method', METHODS)
@pytest.mark.parametrize('param', PARAMS)
def test_tick_parity_injected_noise(built, param, method, adaptive, lanes):
    """3 parametrisations x 3 control laws x {fixed, adaptive}; N deliberately not a multiple of the tile (128)."""
    N = 1337
    oracle, ctrl, _ = make_pair('go2', N, control_parametrization=param, sampling_method=method,
                                optimize_step_freq=adaptive, lanes_per_sample=lanes)
    for seed, gait in [(1, 'trot'), (7, 'pace')]:
        sc = synthetic.scenario('go2', gait, seed=seed)
        rng = np.random.RandomState(seed)
        best0 = (0.5 * rng.randn(oracle.P)).astype(np.float32)
        sigma = np.full(oracle.P, 3.0, np.float32) if method == 'cem_mppi' else None
        noise = seeded_noise(oracle, seed, sigma)
        freqs = seeded_freqs(oracle, seed) if adaptive else None
        ref, got = run_both(oracle, ctrl, sc, best0, noise, sigma=sigma, freqs=freqs)
        assert_tick_parity(ref, got, cem=(method == 'cem_mppi'), noise=noise)


@pytest.mark.parametrize('lanes', LANES)
@pytest.mark.parametrize('adaptive', [False, True])
@pytest.mark.parametrize('param', PARAMS)
def test_rollout_seam(built, param, adaptive, lanes):
    """jit_vectorized_rollout (S:176-177 / A:186-187): explicit parameters in, raw costs out."""
    N = 600
    oracle, ctrl, _ = make_pair('aliengo', N, control_parametrization=param, optimize_step_freq=adaptive,
                                lanes_per_sample=lanes)
    sc = synthetic.scenario('aliengo', 'trot', seed=11)
    rng = np.random.RandomState(5)
    params = (4.0 * rng.randn(N, oracle.P)).astype(np.float32)
    if adaptive:
        freqs = seeded_freqs(oracle, 3)
        want = oracle.compute_rollout_adaptive(sc['state'], sc['reference'], sc['phase_signal'], params, freqs)
        got = ctrl.jit_vectorized_rollout(sc['state'], sc['reference'], sc['phase_signal'], params, freqs)
    else:
        want = oracle.compute_rollout(sc['state'], sc['reference'], params, sc['contact_sequence'])
        got = ctrl.jit_vectorized_rollout(sc['state'], sc['reference'], params, sc['contact_sequence'])
    np.testing.assert_allclose(got, want, rtol=COST_RTOL)


@pytest.mark.parametrize('lanes', LANES)
def test_config2_full_size(built, lanes):
    """BASELINE config 2 at full size: Go2 trot, MPPI, cubic, N = 10,000 with injected noise."""
    oracle, ctrl, _ = make_pair('go2', 10000, control_parametrization='cubic_spline', sampling_method='mppi',
                                lanes_per_sample=lanes)
    sc = synthetic.scenario('go2', 'trot', seed=0)
    best0 = np.zeros(oracle.P, np.float32)
    noise = seeded_noise(oracle, 0)
    ref, got = run_both(oracle, ctrl, sc, best0, noise)
    assert_tick_parity(ref, got)


@pytest.mark.parametrize('lanes', LANES)
def test_config1_full_size(built, lanes):
    """BASELINE config 1: Aliengo trot, random_sampling, zero-order, N = 10,000."""
    oracle, ctrl, _ = make_pair('aliengo', 10000, control_parametrization='zero_order', sampling_method='random_sampling',
                                lanes_per_sample=lanes)
    sc = synthetic.scenario('aliengo', 'trot', seed=2)
    best0 = np.zeros(oracle.P, np.float32)
    noise = seeded_noise(oracle, 2)
    ref, got = run_both(oracle, ctrl, sc, best0, noise)
    assert_tick_parity(ref, got)


@pytest.mark.parametrize('lanes', LANES)
def test_philox_tick_matches_oracle_on_exported_noise(built, lanes):
    """Production mode (in-kernel Philox): the tick must equal the oracle fed with the noise the device exports."""
    for method in METHODS:
        oracle, ctrl, _ = make_pair('go2', 3000, control_parametrization='cubic_spline', sampling_method=method,
                                    lanes_per_sample=lanes)
        sc = synthetic.scenario('go2', 'trot', seed=4)
        ctrl.with_newkey()
        key = ctrl.master_key.copy()
        sigma = np.linspace(0.5, 4.0, oracle.P).astype(np.float32) if method == 'cem_mppi' else None
        noise = ctrl.generate_noise(key, sigma)
        assert np.all(noise[0] == 0.0)
        best0 = np.zeros(oracle.P, np.float32)
        ref = oracle.compute_control(sc['state'], sc['reference'], sc['contact_sequence'], best0, noise, sigma=sigma)
        args = (sigma,) if method == 'cem_mppi' else (sc['phase_signal'], sc['nominal_step_freq'], 0)
        out = ctrl.jitted_compute_control(sc['state'], sc['reference'], sc['contact_sequence'], best0, key, *args)
        got = dict(nmpc_GRFs=out[0], nmpc_predicted_state=out[2], best_control_parameters=out[3], best_cost=out[4],
                   best_freq=out[5], costs=np.asarray(out[6]), best_index=ctrl.last_best_index)
        if method == 'cem_mppi':
            got['new_sigma_cem_mppi'], got['elite_indices'] = out[7], ctrl.last_elite_indices
        assert_tick_parity(ref, got, cem=(method == 'cem_mppi'), noise=noise)


@pytest.mark.parametrize('method', METHODS)
def test_device_noise_matches_host_mirror(built, method):
    """The in-kernel generator (Philox4x32-10 + Box-Muller on the SFU) vs its host mirror in oracle/sampling_oracle.c
    (same integer stream, libm instead of the SFU approximations) and the distribution the reference asks for."""
    from oracle import c_oracle
    N = 6000
    oracle, ctrl, ocfg = make_pair('go2', N, control_parametrization='cubic_spline', sampling_method=method)
    key = np.array([42, 5], np.uint32)
    sigma = np.linspace(0.5, 4.0, oracle.P).astype(np.float32) if method == 'cem_mppi' else None
    dev = ctrl.generate_noise(key, sigma)
    host = c_oracle.COracle(ocfg).noise(42, 5, sigma=sigma)
    np.testing.assert_allclose(dev, host, rtol=0, atol=2e-4)
    assert np.all(dev[0] == 0.0)
    if method == 'mppi':
        from test_oracle import _normal_distribution_checks       # moments, KS distance, tail mass, independence
        big = make_pair('go2', 30001, control_parametrization='cubic_spline', sampling_method=method)[1]
        _normal_distribution_checks(big.generate_noise(key)[1:] / np.float32(3.0))


@pytest.mark.parametrize('method', METHODS)
@pytest.mark.parametrize('N', [1, 2, 127, 129, 1000, 18944, 18945])
def test_kernel_variants_agree_bit_for_bit_on_ragged_sizes(built, method, N):
    """Production mode (in-kernel Philox): sample per thread, the same with noise-helper threads (one tile per block on
    <= 148 blocks: N <= 18 944, otherwise it IS the plain kernel) and four lanes per sample draw the same rows and must
    return the same tick, also when the last tile is ragged or holds only the zero-noise row."""
    import quadruped_pympc_b200 as qp
    from quadruped_pympc_b200 import synthetic
    sc = synthetic.scenario('go2', 'trot', seed=11)
    outs, keep = [], []
    for lanes in (1, 14, 4):
        cfg = qp.SamplerConfig.for_robot('go2', sampling_method=method, control_parametrization='cubic_spline',
                                         num_parallel_computations=N, lanes_per_sample=lanes)
        ctrl = qp.Sampling_MPC(cfg)
        keep.append(ctrl)                                 # costs are fetched lazily from the live handle
        best = np.linspace(-1, 1, ctrl.num_control_parameters).astype(np.float32)
        args = (np.full(ctrl.num_control_parameters, 2.0, np.float32),) if method == 'cem_mppi' else ()
        ctrl.with_newkey()
        outs.append(ctrl.jitted_compute_control(sc['state'], sc['reference'], sc['contact_sequence'], best, ctrl.master_key,
                                                *args, sc['phase_signal'], sc['nominal_step_freq'], 1))
        assert ctrl.last_best_index < N
    for o in outs[1:2]:                                   # helpers: bit-identical to the plain kernel
        for a, b in zip(outs[0], o):
            np.testing.assert_array_equal(np.asarray(a), np.asarray(b))
    # four lanes: a different summation order inside the rollout -> FP32 tolerance of the north_star
    np.testing.assert_allclose(np.asarray(outs[2][6]), np.asarray(outs[0][6]), rtol=3 * COST_RTOL)
    np.testing.assert_allclose(np.asarray(outs[2][0]), np.asarray(outs[0][0]), rtol=1e-4, atol=1e-3)


def test_kernel_variant_selection(built):
    """lanes_per_sample = 0: four lanes per sample up to 6 144 samples per launch, sample per thread + noise helpers for one
    controller of up to 148 tiles, plain sample per thread otherwise (batches, multi-tile launches)."""
    import quadruped_pympc_b200 as qp
    from quadruped_pympc_b200.batched import BatchedSampling_MPC

    def name(N, lanes=0, robots=None):
        cfg = qp.SamplerConfig.for_robot('go2', sampling_method='mppi', num_parallel_computations=N, lanes_per_sample=lanes)
        if robots is None:
            return qp.Sampling_MPC(cfg).kernel_name
        from quadruped_pympc_b200 import _lib
        b = BatchedSampling_MPC(cfg, robots)
        return _lib.load().qpmpc_kernel_name(b._ensure()).decode()

    assert name(4000) == 'tick_kernel_quad'
    assert name(10000) == 'tick_kernel_h'
    assert name(148 * 128) == 'tick_kernel_h' and name(148 * 128 + 1) == 'tick_kernel'
    assert name(10000, lanes=1) == 'tick_kernel' and name(10000, lanes=4) == 'tick_kernel_quad'
    assert name(200000, lanes=14) == 'tick_kernel'                 # does not qualify: falls back
    assert name(10000, robots=2) == 'tick_kernel' and name(10000, lanes=14, robots=2) == 'tick_kernel'
    assert name(1000, robots=4) == 'tick_kernel_quad'


# ------------------------------------------------------------------------------------------------------------------
# Row a15: the device's own step-frequency sampler (csrc/qpmpc_kernels.cuh sample_step_freq) against the oracle.
#   random_sampling A:687-691: choice(where(optimize_swing, step_freq_delta, nominal))
#   mppi            A:835-838: choice(step_freq_delta), optimize_swing ignored
#   cem_mppi        A:994-1000: choice([0, 0.2, 0.4]) * optimize_swing + nominal
#   best_step_frequency = step_frequencies_vec[best_index]   A:705 | 861 | 1022
# The choice indices come from the integer-exact host mirror of the Philox stream (oracle/sampling_oracle.c
# orc_freq_choice), the assembly per method from the oracle (assemble_step_frequencies).
# ------------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize('nominal', [1.4, 1.1])
@pytest.mark.parametrize('optimize_swing', [0, 1])
@pytest.mark.parametrize('method', METHODS)
def test_device_step_freqs_match_oracle(built, method, optimize_swing, nominal):
    from oracle import c_oracle
    N = 5000
    oracle, ctrl, ocfg = make_pair('go2', N, control_parametrization='linear_spline', sampling_method=method,
                                   optimize_step_freq=True)
    mirror = c_oracle.COracle(ocfg)
    for key in ([42, 1], [42, 977], [7, 3]):
        dev = ctrl.generate_step_freqs(np.array(key, np.uint32), nominal, optimize_swing)
        choice = mirror.freq_choice(key[0], key[1])
        assert choice.min() == 0 and choice.max() == 2 and np.all(np.bincount(choice, minlength=3) > N // 4)
        want = oracle.assemble_step_frequencies(choice, nominal, optimize_swing)
        np.testing.assert_array_equal(dev, want)
        assert dev.dtype == np.float32 and want.dtype == np.float32


def _production_adaptive_tick(oracle, ctrl, sc, method, optimize_swing, sigma=None, best0=None):
    """One tick in production mode (in-kernel noise AND in-kernel step frequencies: nothing injected) and the oracle fed
    with the tensors the device exports for the same key."""
    ctrl.with_newkey()
    key = ctrl.master_key.copy()
    nominal = sc['nominal_step_freq']
    noise = ctrl.generate_noise(key, sigma)
    freqs = ctrl.generate_step_freqs(key, nominal, optimize_swing)
    best0 = np.zeros(oracle.P, np.float32) if best0 is None else best0
    ref = oracle.compute_control(sc['state'], sc['reference'], sc['contact_sequence'], best0, noise, sigma=sigma,
                                 timing=sc['phase_signal'], nominal_step_frequency=nominal,
                                 optimize_swing=optimize_swing, step_frequencies=freqs)
    args = ((sigma,) if method == 'cem_mppi' else ()) + (sc['phase_signal'], nominal, optimize_swing)
    out = ctrl.jitted_compute_control(sc['state'], sc['reference'], sc['contact_sequence'], best0, key, *args)
    got = dict(nmpc_GRFs=out[0], nmpc_predicted_state=out[2], best_control_parameters=out[3], best_cost=out[4],
               best_freq=out[5], costs=np.asarray(out[6]), best_index=ctrl.last_best_index)
    if method == 'cem_mppi':
        got['new_sigma_cem_mppi'], got['elite_indices'] = out[7], ctrl.last_elite_indices
    return ref, got, noise, freqs


@pytest.mark.parametrize('lanes', LANES)
@pytest.mark.parametrize('optimize_swing', [0, 1])
@pytest.mark.parametrize('method', METHODS)
def test_production_adaptive_tick_matches_oracle_on_exported_noise_and_freqs(built, method, optimize_swing, lanes):
    """a15 end to end: gait-adaptive tick with NOTHING injected vs the oracle on the exported noise and frequencies;
    best_freq = step_frequencies[best_index], best_index and the elite indices must match exactly."""
    from oracle import c_oracle
    N = 3000
    for param in ('cubic_spline', 'zero_order'):
        oracle, ctrl, ocfg = make_pair('go2', N, control_parametrization=param, sampling_method=method,
                                       optimize_step_freq=True, lanes_per_sample=lanes)
        sc = synthetic.scenario('go2', 'trot', seed=9)
        sigma = np.linspace(0.5, 4.0, oracle.P).astype(np.float32) if method == 'cem_mppi' else None
        for _ in range(2):                                   # two ticks: the key counter advances, the streams change
            ref, got, noise, freqs = _production_adaptive_tick(oracle, ctrl, sc, method, optimize_swing, sigma)
            # the exported frequencies are the oracle's own assembly of the mirrored Philox choices
            choice = c_oracle.COracle(ocfg).freq_choice(int(ctrl.master_key[0]), int(ctrl.master_key[1]))
            np.testing.assert_array_equal(freqs, oracle.assemble_step_frequencies(choice, sc['nominal_step_freq'], optimize_swing))
            assert_tick_parity(ref, got, cem=(method == 'cem_mppi'), noise=noise)
            assert got['best_freq'] == freqs[ref['best_index']]


@pytest.mark.parametrize('lanes', LANES)
def test_config3_full_size_production_mode(built, lanes):
    """BASELINE config 3 as it runs in production: HyQReal pace, CEM-MPPI with step-frequency adaptation, cubic,
    N = 50,000, 200 elites, in-kernel noise and in-kernel step frequencies."""
    oracle, ctrl, _ = make_pair('hyqreal1', 50000, control_parametrization='cubic_spline', sampling_method='cem_mppi',
                                optimize_step_freq=True, lanes_per_sample=lanes)
    sc = synthetic.scenario('hyqreal1', 'pace', seed=3)
    sigma = np.full(oracle.P, 3.0, np.float32)
    ref, got, noise, freqs = _production_adaptive_tick(oracle, ctrl, sc, 'cem_mppi', 1, sigma)
    assert len(got['elite_indices']) == 200
    assert set(np.unique(freqs)) == {np.float32(1.4), np.float32(1.4) + np.float32(0.2), np.float32(1.4) + np.float32(0.4)}
    assert_tick_parity(ref, got, cem=True, noise=noise)
    assert got['best_freq'] == freqs[ref['best_index']]
