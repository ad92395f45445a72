"""GPU: one controller whose samples are split over `world` shards.  The shards are emulated on
ONE GPU as `world` handles ticked one after another (no kernel waits on another, see B200_PROFILING.md); their records
are gathered by the test instead of ncclAllGather, then every shard finishes.  Result must equal the unsharded tick and
the oracle, with global argmin / elite indices."""
import numpy as np
import pytest

import quadruped_pympc_b200 as qp
from quadruped_pympc_b200 import synthetic
from quadruped_pympc_b200 import _lib
from helpers import make_pair, seeded_noise, seeded_freqs, assert_tick_parity

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize('world', [2, 3, 8])
@pytest.mark.parametrize('method,adaptive', [('random_sampling', False), ('mppi', False), ('cem_mppi', False),
                                             ('mppi', True), ('cem_mppi', True)])
def test_sharded_tick_equals_oracle(built, method, adaptive, world):
    N = 2501
    oracle, _, _ = make_pair('go2', N, control_parametrization='cubic_spline', sampling_method=method,
                             optimize_step_freq=adaptive)
    sc = synthetic.scenario('go2', 'trot', seed=9)
    rng = np.random.RandomState(3)
    best0 = (0.5 * rng.randn(oracle.P)).astype(np.float32)
    sigma = np.linspace(1.0, 4.0, oracle.P).astype(np.float32) if method == 'cem_mppi' else None
    noise = seeded_noise(oracle, 5, sigma)
    freqs = seeded_freqs(oracle, 5) if adaptive else None
    ref = oracle.compute_control(sc['state'], sc['reference'], sc['contact_sequence'], best0, noise, sigma=sigma,
                                 timing=sc['phase_signal'], nominal_step_frequency=sc['nominal_step_freq'],
                                 optimize_swing=1, step_frequencies=freqs)
    cfg = qp.SamplerConfig.for_robot('go2', num_parallel_computations=N, control_parametrization='cubic_spline',
                                     sampling_method=method, optimize_step_freq=adaptive)
    shards = [qp.Sampling_MPC(cfg, shard_rank=r, shard_world=world) for r in range(world)]
    key = np.array([42, 1], np.uint32)
    recs = [s.shard_partial(sc['state'], sc['reference'], sc['contact_sequence'], best0, key, sigma=sigma,
                            timing=sc['phase_signal'], nominal=sc['nominal_step_freq'], optimize_swing=1,
                            noise=noise, step_freqs=freqs) for s in shards]
    outs = [s.shard_finish(recs) for s in shards]
    costs = np.concatenate([np.asarray(o[6]) for o in outs])
    for s, o in zip(shards, outs):
        got = dict(nmpc_GRFs=o[0], nmpc_predicted_state=o[2], best_control_parameters=o[3], best_cost=o[4],
                   best_freq=o[5], costs=costs, best_index=s.last_best_index)
        if method == 'cem_mppi':
            got['new_sigma_cem_mppi'], got['elite_indices'] = o[7], s.last_elite_indices
        assert_tick_parity(ref, got, cem=(method == 'cem_mppi'), noise=noise)
    # every shard computed the identical result (deterministic merge in rank order)
    for o in outs[1:]:
        assert np.array_equal(o[3], outs[0][3]) and np.array_equal(o[0], outs[0][0])


def test_sharded_philox_equals_unsharded(built):
    """Production mode: shards draw their slice of the SAME global noise tensor (global row index in the counter)."""
    N, world = 4000, 4
    cfg = qp.SamplerConfig.for_robot('go2', num_parallel_computations=N, sampling_method='mppi')
    sc = synthetic.scenario('go2', 'trot', seed=2)
    key = np.array([42, 7], np.uint32)
    best0 = np.zeros(96, np.float32)
    whole = qp.Sampling_MPC(cfg).compute_control(sc['state'], sc['reference'], sc['contact_sequence'], best0, key,
                                                 sc['phase_signal'], 1.4, 0)
    shards = [qp.Sampling_MPC(cfg, shard_rank=r, shard_world=world) for r in range(world)]
    recs = [s.shard_partial(sc['state'], sc['reference'], sc['contact_sequence'], best0, key) for s in shards]
    out = shards[0].shard_finish(recs)
    assert shards[0].last_best_index == int(np.argmin(np.asarray(whole[6])))
    np.testing.assert_allclose(out[4], whole[4], rtol=1e-6)
    np.testing.assert_allclose(out[3], whole[3], rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(out[0], whole[0], rtol=1e-4, atol=1e-4)


@pytest.mark.parametrize('lanes', [1, 4, 14, 2])
@pytest.mark.parametrize('method,adaptive,world', [('random_sampling', False, 2), ('mppi', False, 3), ('cem_mppi', False, 8),
                                                   ('mppi', True, 2), ('cem_mppi', True, 4)])
def test_fused_peer_exchange_loopback(built, method, adaptive, world, lanes):
    """The collective fused into the tick kernel (qpmpc_tick.cuh peer_exchange), on ONE GPU: each shard in turn runs
    its fused tick with the peers' records delivered to its mailbox beforehand (no kernel waits on a running kernel).
    Every shard must return the oracle's result, bit-identical across shards; repeated for three ticks so that both
    mailbox parities and the stamp sequence are exercised."""
    N = 2501
    oracle, _, _ = make_pair('go2', N, control_parametrization='cubic_spline', sampling_method=method,
                             optimize_step_freq=adaptive)
    cfg = qp.SamplerConfig.for_robot('go2', num_parallel_computations=N, control_parametrization='cubic_spline',
                                     sampling_method=method, optimize_step_freq=adaptive, lanes_per_sample=lanes)
    shards = [qp.Sampling_MPC(cfg, shard_rank=r, shard_world=world).peer_loopback() for r in range(world)]
    sc = synthetic.scenario('go2', 'trot', seed=9)
    rng = np.random.RandomState(3)
    best0 = (0.5 * rng.randn(oracle.P)).astype(np.float32)
    sigma = np.linspace(1.0, 4.0, oracle.P).astype(np.float32) if method == 'cem_mppi' else None
    for tick in range(3):
        noise = seeded_noise(oracle, 5 + tick, sigma)
        freqs = seeded_freqs(oracle, 5 + tick) if adaptive else None
        ref = oracle.compute_control(sc['state'], sc['reference'], sc['contact_sequence'], best0, noise, sigma=sigma,
                                     timing=sc['phase_signal'], nominal_step_frequency=sc['nominal_step_freq'],
                                     optimize_swing=1, step_frequencies=freqs)
        key = np.array([42, 1 + tick], np.uint32)
        recs = [s.shard_partial(sc['state'], sc['reference'], sc['contact_sequence'], best0, key, sigma=sigma,
                                timing=sc['phase_signal'], nominal=sc['nominal_step_freq'], optimize_swing=1,
                                noise=noise, step_freqs=freqs) for s in shards]
        outs = []
        for me, s in enumerate(shards):
            for w in range(world):
                if w != me:
                    s.peer_inject(w, recs[w])
            args = (sigma, sc['phase_signal'], sc['nominal_step_freq'], 1) if method == 'cem_mppi' else \
                   (sc['phase_signal'], sc['nominal_step_freq'], 1)
            outs.append(s.jitted_compute_control(sc['state'], sc['reference'], sc['contact_sequence'], best0, key, *args,
                                                 noise=noise, step_freqs=freqs))
        costs = np.concatenate([np.asarray(o[6]) for o in outs])
        for s, o in zip(shards, outs):
            got = dict(nmpc_GRFs=o[0], nmpc_predicted_state=o[2], best_control_parameters=o[3], best_cost=o[4],
                       best_freq=o[5], costs=costs, best_index=s.last_best_index)
            if method == 'cem_mppi':
                got['new_sigma_cem_mppi'], got['elite_indices'] = o[7], s.last_elite_indices
            assert_tick_parity(ref, got, cem=(method == 'cem_mppi'), noise=noise)
        for o in outs[1:]:
            assert np.array_equal(o[3], outs[0][3]) and np.array_equal(o[0], outs[0][0])


def test_fused_peer_exchange_production_mode_loopback(built):
    """Direct I/O + in-kernel Philox through the fused exchange equals the unsharded tick."""
    N, world = 10000, 2
    cfg = qp.SamplerConfig.for_robot('go2', num_parallel_computations=N, sampling_method='mppi')
    sc = synthetic.scenario('go2', 'trot', seed=2)
    key = np.array([42, 7], np.uint32)
    best0 = np.zeros(96, np.float32)
    whole_ctrl = qp.Sampling_MPC(cfg)
    whole = whole_ctrl.compute_control(sc['state'], sc['reference'], sc['contact_sequence'], best0, key, sc['phase_signal'], 1.4, 0)
    shards = [qp.Sampling_MPC(cfg, shard_rank=r, shard_world=world).peer_loopback() for r in range(world)]
    recs = [s.shard_partial(sc['state'], sc['reference'], sc['contact_sequence'], best0, key) for s in shards]
    for me, s in enumerate(shards):
        s.peer_inject(1 - me, recs[1 - me])
        out = s.compute_control(sc['state'], sc['reference'], sc['contact_sequence'], best0, key, sc['phase_signal'], 1.4, 0)
        assert s.last_best_index == whole_ctrl.last_best_index
        np.testing.assert_allclose(out[4], whole[4], rtol=1e-6)
        np.testing.assert_allclose(out[3], whole[3], rtol=1e-4, atol=1e-5)
        np.testing.assert_allclose(out[0], whole[0], rtol=1e-4, atol=1e-4)


def test_peer_attach_refuses_two_shards_on_one_gpu(built):
    """Two waiting kernels on one GPU are not guaranteed to run at the same time: attach must refuse."""
    from quadruped_pympc_b200 import _lib
    cfg = qp.SamplerConfig.for_robot('go2', num_parallel_computations=1000, sampling_method='mppi')
    a, b = (qp.Sampling_MPC(cfg, shard_rank=r, shard_world=2) for r in range(2))
    blobs = [a.peer_export(), b.peer_export()]
    with pytest.raises(_lib.QpmpcError, match='same GPU'):
        a.peer_attach(blobs)


@pytest.mark.parametrize('method', ['mppi', 'cem_mppi'])
def test_sharded_two_stage_shards_equal_unsharded(built, method):
    """Shards large enough for the two-stage launch (tick kernel + multi-block tail kernel): the shard records, and the
    tick finished from them, must match the unsharded controller (itself two-stage at this size)."""
    N, world = 120000, 3
    cfg = qp.SamplerConfig.for_robot('go2', num_parallel_computations=N, sampling_method=method)
    sc = synthetic.scenario('go2', 'trot', seed=6)
    key = np.array([42, 3], np.uint32)
    best0 = np.zeros(96, np.float32)
    sig = (np.full(96, 3.0, np.float32),) if method == 'cem_mppi' else ()
    ctrl = qp.Sampling_MPC(cfg)
    whole = ctrl.compute_control(sc['state'], sc['reference'], sc['contact_sequence'], best0, key, *sig,
                                 sc['phase_signal'], 1.4, 0)
    costs = np.asarray(whole[6])
    shards = [qp.Sampling_MPC(cfg, shard_rank=r, shard_world=world) for r in range(world)]
    recs = [s.shard_partial(sc['state'], sc['reference'], sc['contact_sequence'], best0, key, sigma=sig[0] if sig else None)
            for s in shards]
    out = shards[0].shard_finish(recs)
    assert shards[0].last_best_index == int(np.argmin(costs)) == ctrl.last_best_index
    np.testing.assert_allclose(out[4], whole[4], rtol=1e-6)
    np.testing.assert_allclose(out[3], whole[3], rtol=1e-4, atol=2e-5)
    np.testing.assert_allclose(out[0], whole[0], rtol=1e-4, atol=2e-4)
    if method == 'cem_mppi':
        np.testing.assert_array_equal(shards[0].last_elite_indices, np.argsort(costs, kind='stable')[:10])
        np.testing.assert_array_equal(ctrl.last_elite_indices, shards[0].last_elite_indices)
        np.testing.assert_allclose(out[7], whole[7], rtol=1e-4, atol=1e-6)


def test_peer_exchange_times_out_when_a_rank_never_ticks(built):
    """A shard whose peer never delivers its record (process gone) must not hang the controller: the in-kernel wait gives
    up after its time-out (2 s by default; 0.1 s here), the tick returns QPMPC_ERR_NCCL, and the handle keeps working once
    the peer is back."""
    import time
    cfg = qp.SamplerConfig.for_robot('go2', num_parallel_computations=2000, sampling_method='mppi')
    sc = synthetic.scenario('go2', 'trot', seed=2)
    best0 = np.zeros(96, np.float32)
    me = qp.Sampling_MPC(cfg, shard_rank=0, shard_world=2).peer_loopback().peer_timeout(0.1)
    peer = qp.Sampling_MPC(cfg, shard_rank=1, shard_world=2)
    t0 = time.perf_counter()
    with pytest.raises(_lib.QpmpcError) as e:
        me.compute_control(sc['state'], sc['reference'], sc['contact_sequence'], best0, np.array([42, 1], np.uint32),
                           sc['phase_signal'], 1.4, 0)
    dt = time.perf_counter() - t0
    assert e.value.code == _lib.ERR_NCCL and 0.09 < dt < 1.0, (e.value, dt)
    # the peer comes back: its record for the next tick is delivered, the tick completes and matches the unsharded one
    key = np.array([42, 2], np.uint32)
    rec = peer.shard_partial(sc['state'], sc['reference'], sc['contact_sequence'], best0, key)
    me.peer_inject(1, rec)
    out = me.compute_control(sc['state'], sc['reference'], sc['contact_sequence'], best0, key, sc['phase_signal'], 1.4, 0)
    whole = qp.Sampling_MPC(cfg).compute_control(sc['state'], sc['reference'], sc['contact_sequence'], best0, key,
                                                 sc['phase_signal'], 1.4, 0)
    np.testing.assert_allclose(out[3], whole[3], rtol=1e-4, atol=1e-5)
