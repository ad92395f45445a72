"""GPU: the tick loop of SRBDControllerInterface.compute_control (I:118-176) replayed verbatim against the B200 class
(dict inputs, prepare_state_and_reference, with_newkey / with_newsigma, several sampling iterations, several ticks with a
lift-off in between) and against the oracle driven the same way on the noise the device exports."""
import numpy as np
import pytest

import quadruped_pympc_b200 as qp
from quadruped_pympc_b200 import synthetic
from oracle.sampling_oracle import OracleConfig, SamplingOracle
from helpers import COST_RTOL, KNOT_RTOL, knot_atol

pytestmark = pytest.mark.gpu


def _dicts(sc):
    s, r = sc['state'], sc['reference']
    st = dict(position=s[0:3], linear_velocity=s[3:6], orientation=s[6:9], angular_velocity=s[9:12],
              foot_FL=s[12:15], foot_FR=s[15:18], foot_RL=s[18:21], foot_RR=s[21:24])
    rf = dict(ref_position=r[0:3], ref_linear_velocity=r[3:6], ref_orientation=r[6:9], ref_angular_velocity=r[9:12],
              ref_foot_FL=r[12:15].reshape(1, 3), ref_foot_FR=r[15:18].reshape(1, 3),
              ref_foot_RL=r[18:21].reshape(1, 3), ref_foot_RR=r[21:24].reshape(1, 3))
    return st, rf


@pytest.mark.parametrize('method', ['random_sampling', 'mppi', 'cem_mppi'])
def test_interface_tick_loop(built, method):
    kw = dict(sampling_method=method, control_parametrization='cubic_spline', num_parallel_computations=2000,
              num_sampling_iterations=3)
    controller = qp.Sampling_MPC(qp.SamplerConfig.for_robot('go2', **kw))
    orc = SamplingOracle(OracleConfig.for_robot('go2', **kw))
    best_ref = np.zeros(orc.P, np.float32)
    previous_contact_mpc = np.array([1, 1, 1, 1])
    sigma_cfg = 3                                              # cfg.mpc_params['sigma_cem_mppi']
    for tick in range(4):
        sc = synthetic.scenario('go2', 'trot', seed=10 + tick, sim_steps=60 * tick)
        state_current, ref_state = _dicts(sc)
        contact_sequence = np.concatenate([sc['contact_sequence'], np.ones((4, 12))], axis=1)   # (4, 2H) is allowed
        current_contact = contact_sequence[:, 0].copy()
        # ---- I:120-123
        state_jax, ref_jax = controller.prepare_state_and_reference(state_current, ref_state, current_contact,
                                                                    previous_contact_mpc)
        s_o, r_o, best_ref = orc.prepare_state_and_reference(state_current, ref_state, current_contact,
                                                            previous_contact_mpc, best_ref)
        np.testing.assert_array_equal(state_jax, s_o)
        np.testing.assert_array_equal(controller.best_control_parameters, best_ref)
        previous_contact_mpc = current_contact
        sigma_ref = None
        for iter_sampling in range(controller.num_sampling_iterations):          # I:125
            controller = controller.with_newkey()                                # I:126
            if controller.sampling_method == 'cem_mppi':
                if iter_sampling == 0:
                    controller = controller.with_newsigma(sigma_cfg)             # I:128-129
                    sigma_ref = np.full(orc.P, sigma_cfg, np.float32)
                noise = controller.generate_noise(controller.master_key, controller.sigma_cem_mppi)
                (nmpc_GRFs, nmpc_footholds, nmpc_predicted_state, controller.best_control_parameters, best_cost,
                 best_sample_freq, costs, sigma_cem_mppi) = controller.jitted_compute_control(
                    state_jax, ref_jax, contact_sequence, controller.best_control_parameters, controller.master_key,
                    controller.sigma_cem_mppi)                                   # I:131-147
                controller = controller.with_newsigma(sigma_cem_mppi)            # I:148
            else:
                noise = controller.generate_noise(controller.master_key)
                (nmpc_GRFs, nmpc_footholds, nmpc_predicted_state, controller.best_control_parameters, best_cost,
                 best_sample_freq, costs) = controller.jitted_compute_control(
                    state_jax, ref_jax, contact_sequence, controller.best_control_parameters, controller.master_key,
                    sc['phase_signal'], sc['nominal_step_freq'], 1)              # I:151-168
            ref = orc.compute_control(s_o, r_o, contact_sequence[:, :12], best_ref, noise, sigma=sigma_ref)
            rel = np.max(np.abs(np.asarray(costs) - ref['costs']) / np.abs(ref['costs']))
            assert rel < COST_RTOL
            assert controller.last_best_index == ref['best_index']
            atol = knot_atol(ref['costs'], np.asarray(costs), ref.get('weights'), noise)
            np.testing.assert_allclose(controller.best_control_parameters, ref['best_control_parameters'],
                                       rtol=KNOT_RTOL, atol=atol)
            np.testing.assert_allclose(np.array(nmpc_GRFs), ref['nmpc_GRFs'], rtol=KNOT_RTOL, atol=atol)
            # carry the ORACLE's state forward from the device's best so that rounding does not accumulate across ticks
            best_ref = np.asarray(controller.best_control_parameters, np.float32).copy()
            if method == 'cem_mppi':
                np.testing.assert_allclose(sigma_cem_mppi, ref['new_sigma_cem_mppi'], rtol=KNOT_RTOL, atol=1e-6)
                sigma_ref = np.asarray(sigma_cem_mppi, np.float32).copy()
        assert np.array(nmpc_GRFs).shape == (12,) and nmpc_footholds.shape == (12,)
        assert np.all(np.array(nmpc_GRFs).reshape(4, 3)[current_contact == 0] == 0)


@pytest.mark.parametrize('lanes', [1, 4, 14])
@pytest.mark.parametrize('method,adaptive', [('random_sampling', False), ('mppi', False), ('cem_mppi', False),
                                             ('mppi', True), ('cem_mppi', True)])
def test_device_resident_sampling_iterations_equal_the_host_loop(built, method, adaptive, lanes):
    """the loop I:125-168 (with_newkey / best / with_newsigma between iterations) in one call with the
    state carried on the device must equal the loop driven from the host, bit for bit."""
    kw = dict(sampling_method=method, control_parametrization='cubic_spline', num_parallel_computations=3000,
              num_sampling_iterations=4, optimize_step_freq=adaptive, lanes_per_sample=lanes)
    sc = synthetic.scenario('go2', 'trot', seed=21)
    cem = method == 'cem_mppi'
    tail = (sc['phase_signal'], sc['nominal_step_freq'], 1)

    host = qp.Sampling_MPC(qp.SamplerConfig.for_robot('go2', **kw))
    best = np.zeros(host.num_control_parameters, np.float32)
    host.with_newsigma(3)
    for _ in range(host.num_sampling_iterations):
        host.with_newkey()
        args = ((host.sigma_cem_mppi,) if cem else ()) + tail
        out_h = host.jitted_compute_control(sc['state'], sc['reference'], sc['contact_sequence'], best, host.master_key, *args)
        best = out_h[3]
        if cem:
            host.with_newsigma(out_h[7])

    dev = qp.Sampling_MPC(qp.SamplerConfig.for_robot('go2', **kw))
    dev.with_newsigma(3)
    dev.with_newkey()
    args = ((dev.sigma_cem_mppi,) if cem else ()) + tail
    out_d = dev.compute_control_iterations(sc['state'], sc['reference'], sc['contact_sequence'],
                                           np.zeros(dev.num_control_parameters, np.float32), dev.master_key, *args)
    for a, b in zip(out_h[:6], out_d[:6]):
        np.testing.assert_array_equal(np.asarray(a), np.asarray(b))
    np.testing.assert_array_equal(np.asarray(out_h[6]), np.asarray(out_d[6]))      # costs of the last iteration
    if cem:
        np.testing.assert_array_equal(out_h[7], out_d[7])
    np.testing.assert_array_equal(host.master_key, dev.master_key)
    assert host.last_best_index == dev.last_best_index


@pytest.mark.parametrize('lanes,N', [(1, 2500), (4, 2500), (14, 2500), (1, 26000)])
@pytest.mark.parametrize('method,adaptive,iters', [('random_sampling', False, 1), ('mppi', False, 3), ('cem_mppi', False, 2),
                                                   ('mppi', True, 2), ('cem_mppi', True, 1)])
def test_stateful_step_equals_the_interface_block(built, method, adaptive, iters, lanes, N):
    """`step()` (device-resident best / sigma / key counter / previous contact, host prep done in the
    kernel's staging) against the block I:118-168 driven from the host with the same class, over ticks that contain
    lift-offs and touch-downs.  Must agree bit for bit, including the carried state."""
    kw = dict(sampling_method=method, control_parametrization='cubic_spline', num_parallel_computations=N,
              num_sampling_iterations=iters, optimize_step_freq=adaptive, lanes_per_sample=lanes)   # N = 26 000: two-stage launches
    cem = method == 'cem_mppi'
    host = qp.Sampling_MPC(qp.SamplerConfig.for_robot('go2', **kw))
    dev = qp.Sampling_MPC(qp.SamplerConfig.for_robot('go2', **kw))
    previous_contact_mpc = np.array([1, 1, 1, 1])
    saw_liftoff = False
    for tick in range(6):
        sc = synthetic.scenario('go2', 'trot', seed=30 + tick, sim_steps=45 * tick)
        state_current, ref_state = _dicts(sc)
        for k in ('ref_foot_FL', 'ref_foot_FR', 'ref_foot_RL', 'ref_foot_RR'):      # footholds differ from the feet
            ref_state[k] = ref_state[k] + 0.01 * (tick + 1)
        contact_sequence = sc['contact_sequence']
        current_contact = contact_sequence[:, 0].copy()
        saw_liftoff |= bool(np.any((previous_contact_mpc == 1) & (current_contact == 0)))
        # ---- host-driven block, I:118-168
        state_jax, ref_jax = host.prepare_state_and_reference(state_current, ref_state, current_contact, previous_contact_mpc)
        previous_contact_mpc = current_contact
        for it in range(host.num_sampling_iterations):
            host.with_newkey()
            if cem and it == 0:
                host.with_newsigma(host.cfg.sigma_cem_mppi)
            tail = (sc['phase_signal'], sc['nominal_step_freq'], 1)
            args = ((host.sigma_cem_mppi,) if cem else ()) + tail
            out_h = host.jitted_compute_control(state_jax, ref_jax, contact_sequence, host.best_control_parameters,
                                                host.master_key, *args)
            host.best_control_parameters = out_h[3]
            if cem:
                host.with_newsigma(out_h[7])
        # ---- the same block in one call
        out_d = dev.step(state_current, ref_state, contact_sequence, sc['phase_signal'], sc['nominal_step_freq'], 1)
        for a, b in zip(out_h[:6], out_d[:6]):
            np.testing.assert_array_equal(np.asarray(a), np.asarray(b))
        if cem:
            np.testing.assert_array_equal(out_h[7], out_d[7])
        np.testing.assert_array_equal(host.master_key, dev.master_key)
        best, sigma, pc, key = dev.step_state()
        np.testing.assert_array_equal(best, out_h[3])
        np.testing.assert_array_equal(pc, current_contact.astype(int))
        assert key[1] == host.master_key[1] + 1
    assert saw_liftoff


@pytest.mark.gpu
@pytest.mark.parametrize('param', ['zero_order', 'linear_spline', 'cubic_spline'])
@pytest.mark.parametrize('method,lanes', [('mppi', 1), ('cem_mppi', 4), ('random_sampling', 1)])
def test_stateful_step_with_shift_solution(built, param, method, lanes):
    """mpc_params['shift_solution'] = True (config.py:194 has False): the shift of the carried best (S:513-561, called from
    prepare_state_and_reference S:571-573) done in the kernel's staging must equal the host-driven block bit for bit,
    single and batched, 2 sampling iterations (the shift applies once per tick, before iteration 0)."""
    from quadruped_pympc_b200.batched import BatchedSampling_MPC
    kw = dict(sampling_method=method, control_parametrization=param, num_parallel_computations=1500,
              num_sampling_iterations=2, lanes_per_sample=lanes, shift_solution=True)
    cem = method == 'cem_mppi'
    cfg = qp.SamplerConfig.for_robot('go2', **kw)
    host, dev, bat = qp.Sampling_MPC(cfg), qp.Sampling_MPC(cfg), BatchedSampling_MPC(cfg, 3)
    previous_contact_mpc = np.array([1, 1, 1, 1])
    for tick in range(5):
        sc = synthetic.scenario('go2', 'trot', seed=60 + tick, sim_steps=45 * tick)
        state_current, ref_state = _dicts(sc)
        contact_sequence = sc['contact_sequence']
        current_contact = contact_sequence[:, 0].copy()
        before = host.best_control_parameters.copy()
        state_jax, ref_jax = host.prepare_state_and_reference(state_current, ref_state, current_contact, previous_contact_mpc)
        if tick > 0:
            assert not np.array_equal(before, host.best_control_parameters)         # the shift did something
        previous_contact_mpc = current_contact
        for it in range(2):
            host.with_newkey()
            if cem and it == 0:
                host.with_newsigma(host.cfg.sigma_cem_mppi)
            args = ((host.sigma_cem_mppi,) if cem else ()) + (sc['phase_signal'], sc['nominal_step_freq'], 1)
            out_h = host.jitted_compute_control(state_jax, ref_jax, contact_sequence, host.best_control_parameters,
                                                host.master_key, *args)
            host.best_control_parameters = out_h[3]
            if cem:
                host.with_newsigma(out_h[7])
        out_d = dev.step(state_current, ref_state, contact_sequence, sc['phase_signal'], sc['nominal_step_freq'], 1)
        for a, b in zip(out_h[:6], out_d[:6]):
            np.testing.assert_array_equal(np.asarray(a), np.asarray(b))
        # batched: robot 0 of the batch has robot id 0 = the single controller's stream
        raw = np.concatenate([np.asarray(state_current[k], np.float64).reshape(-1) for k in qp.Sampling_MPC._STATE_KEYS])
        rf = np.concatenate([np.asarray(ref_state[k], np.float64).reshape(-1)[:3] for k in qp.Sampling_MPC._REF_KEYS])
        out_b = bat.step(np.tile(raw, (3, 1)), np.tile(rf, (3, 1)), np.tile(contact_sequence[None], (3, 1, 1)),
                         timing=np.tile(np.asarray(sc['phase_signal'], np.float32), (3, 1)),
                         nominal_step_frequency=sc['nominal_step_freq'], optimize_swing=1, iterations=2)
        np.testing.assert_array_equal(out_b[0][0], np.asarray(out_h[0]))
        np.testing.assert_array_equal(out_b[2][0], np.asarray(out_h[3]))


@pytest.mark.gpu
def test_shift_is_refused_for_the_gait_adaptive_controller(built):
    cfg = qp.SamplerConfig.for_robot('go2', sampling_method='mppi', num_parallel_computations=512, optimize_step_freq=True,
                                     shift_solution=True)
    ctrl = qp.Sampling_MPC(cfg)
    sc = synthetic.scenario('go2', 'trot', seed=1)
    state_current, ref_state = _dicts(sc)
    with pytest.raises(Exception):
        ctrl.step(state_current, ref_state, sc['contact_sequence'], sc['phase_signal'], sc['nominal_step_freq'], 1)


@pytest.mark.parametrize('case', ['mppi', 'random_sampling_adaptive', 'cem_mppi'])
def test_replay_of_the_reference_interface_trace_matches_the_oracle(built, case):
    """tests/golden/interface_trace.npz = the arguments that crossed the C ABI when the reference's own
    SRBDControllerInterface (INTEGRATION.md patch applied) drove this class (tools/make_interface_trace.py,
    tests/test_reference_interface.py).  Every recorded tick is replayed here on the GPU, in production mode, and must
    equal the oracle on the noise / step frequencies the device exports for the recorded key."""
    import os
    from helpers import assert_tick_parity
    from oracle.sampling_oracle import OracleConfig, SamplingOracle
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden', 'interface_trace.npz'))
    N, H, ns, par, method, adaptive, seed = [int(x) for x in z[f'{case}/config']]
    mass, mu, gmin, gmax = z[f'{case}/mass_mu_grf']
    robot = {15.019: 'go2', 24.637: 'aliengo'}[round(float(mass), 3)]
    kw = dict(sampling_method=['random_sampling', 'mppi', 'cem_mppi'][method],
              control_parametrization=['zero_order', 'linear_spline', 'cubic_spline'][par], num_parallel_computations=N,
              optimize_step_freq=bool(adaptive), num_splines=ns, horizon=H)
    ctrl = qp.Sampling_MPC(qp.SamplerConfig.for_robot(robot, seed=seed, **kw))
    orc = SamplingOracle(OracleConfig.for_robot(robot, **kw))
    cem = method == 2
    for t in range(z[f'{case}/key'].shape[0]):
        key = z[f'{case}/key'][t]
        st, rf, cs, best = z[f'{case}/state'][t], z[f'{case}/reference'][t], z[f'{case}/contact_sequence'][t], z[f'{case}/best'][t]
        sigma = z[f'{case}/sigma'][t] if cem else None
        ph, f, osw = z[f'{case}/phase_signal'][t], float(z[f'{case}/nominal_step_freq'][t]), int(z[f'{case}/optimize_swing'][t])
        noise = ctrl.generate_noise(key, sigma)
        freqs = ctrl.generate_step_freqs(key, f, osw) if adaptive else None
        ref = orc.compute_control(st, rf, cs, best, noise, sigma=sigma, timing=ph, nominal_step_frequency=f,
                                  optimize_swing=osw, step_frequencies=freqs)
        out = ctrl.jitted_compute_control(st, rf, cs, best, key, sigma) if cem else \
            ctrl.jitted_compute_control(st, rf, cs, best, key, ph, f, osw)
        got = dict(nmpc_GRFs=out[0], nmpc_predicted_state=out[2], best_control_parameters=out[3], best_cost=out[4],
                   best_freq=out[5], costs=np.asarray(out[6]), best_index=ctrl.last_best_index)
        if cem:
            got['new_sigma_cem_mppi'], got['elite_indices'] = out[7], ctrl.last_elite_indices
        assert_tick_parity(ref, got, cem=cem, noise=noise)


@pytest.mark.parametrize('method,adaptive', [('mppi', False), ('cem_mppi', False), ('random_sampling', True), ('cem_mppi', True)])
def test_cpython_binding_equals_ctypes_binding(built, method, adaptive):
    """The tick call has two bindings of the SAME C-ABI entry point: ctypes (quadruped_pympc_b200/_lib.py) and the CPython
    module csrc/qpfast.c (marshalling only, ~8 us less Python per tick).  Same inputs -- float64 arrays, a (4, 2H)
    contact table, a non-contiguous state, scalar sigma -- must give the same bits."""
    from quadruped_pympc_b200 import _lib
    assert _lib.load_fast() is not None, 'lib/_qpfast*.so has not been built'
    sc = synthetic.scenario('go2', 'trot', seed=5)
    outs = []
    for fast in (True, False):
        cfg = qp.SamplerConfig.for_robot('go2', sampling_method=method, control_parametrization='cubic_spline',
                                         num_parallel_computations=1500, optimize_step_freq=adaptive, fast_binding=fast)
        ctrl = qp.Sampling_MPC(cfg)
        state = np.asfortranarray(np.stack([sc['state'], sc['state']], axis=1))[:, 0]          # strided float64 view
        cs = np.concatenate([sc['contact_sequence'], np.ones((4, 12))], axis=1)                  # (4, 2H), I: FULL_STANCE shape
        best = np.linspace(-1, 1, 96).astype(np.float32)
        res = []
        for it in range(3):
            ctrl.with_newkey()
            args = ((3 if it == 0 else sigma),) if method == 'cem_mppi' else ()
            args = args + ((sc['phase_signal'], sc['nominal_step_freq'], 1) if (adaptive or method != 'cem_mppi') else ())
            out = ctrl.jitted_compute_control(state, sc['reference'], cs, best, ctrl.master_key, *args)
            best = out[3]
            if method == 'cem_mppi':
                sigma = out[7]
            res.append([np.asarray(x) for x in out] + [ctrl.last_best_index] +
                       ([ctrl.last_elite_indices] if method == 'cem_mppi' else []))
        assert (ctrl._fast is not None) == fast
        outs.append(res)
    for a, b in zip(*outs):
        for x, y in zip(a, b):
            np.testing.assert_array_equal(x, y)
