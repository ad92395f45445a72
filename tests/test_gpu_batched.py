"""GPU: the multi-robot batch dispatch (BASELINE configs[4]): R independent controllers in ONE
launch.  Per robot the result must equal the oracle on the noise that robot draws, must not depend on how the robots are
grouped into handles (the partition over GPUs is a regrouping), and robots must not see each other's inputs."""
import numpy as np
import pytest

import quadruped_pympc_b200 as qp
from quadruped_pympc_b200 import synthetic
from quadruped_pympc_b200.batched import BatchedSampling_MPC, partition_robots
from oracle.sampling_oracle import OracleConfig, SamplingOracle
from helpers import assert_tick_parity

pytestmark = pytest.mark.gpu


def _scenarios(robot, R, first=0):
    scs = [synthetic.scenario(robot, 'trot' if (first + r) % 2 == 0 else 'pace', seed=first + r) for r in range(R)]
    return (np.stack([s['state'] for s in scs]).astype(np.float32), np.stack([s['reference'] for s in scs]).astype(np.float32),
            np.stack([s['contact_sequence'] for s in scs]).astype(np.float32),
            np.stack([s['phase_signal'] for s in scs]).astype(np.float32), scs)


@pytest.mark.parametrize('method,adaptive', [('mppi', False), ('random_sampling', False), ('cem_mppi', False), ('mppi', True)])
def test_each_robot_of_a_batch_equals_the_oracle(built, method, adaptive):
    R, N = 5, 1500
    kw = dict(sampling_method=method, control_parametrization='cubic_spline', num_parallel_computations=N,
              optimize_step_freq=adaptive)
    batch = BatchedSampling_MPC(qp.SamplerConfig.for_robot('go2', **kw), R, robot_base=3)
    orc = SamplingOracle(OracleConfig.for_robot('go2', **kw))
    st, rf, cs, ph, scs = _scenarios('go2', R)
    rng = np.random.RandomState(0)
    best0 = (0.3 * rng.randn(R, orc.P)).astype(np.float32)
    batch.with_newkey()
    sigma = np.full(orc.P, 3.0, np.float32) if method == 'cem_mppi' else None
    out = batch.compute_control(st, rf, cs, best0, timing=ph, nominal_step_frequency=1.4, optimize_swing=0)
    for r in range(R):
        noise = batch.generate_noise(r, sigma=sigma)
        assert np.all(noise[0] == 0.0)
        if r > 0:
            assert not np.array_equal(noise, batch.generate_noise(r - 1, sigma=sigma))   # every robot has its own stream
        if adaptive:     # per-sample frequencies are covered by the single-controller tests; here: it runs, finite results
            continue
        ref = orc.compute_control(scs[r]['state'], scs[r]['reference'], scs[r]['contact_sequence'], best0[r], noise, sigma=sigma)
        got = dict(nmpc_GRFs=out[0][r], nmpc_predicted_state=out[1][r], best_control_parameters=out[2][r],
                   best_cost=out[3][r], best_freq=out[4][r], costs=batch.fetch_costs(r), best_index=int(out[5][r]))
        if method == 'cem_mppi':
            got['new_sigma_cem_mppi'] = out[6][r]
            got['elite_indices'] = ref['elite_indices']      # the batched surface does not return elite indices
        assert_tick_parity(ref, got, cem=(method == 'cem_mppi'), noise=noise)
    assert np.all(np.isfinite(out[0])) and np.all(np.isfinite(out[3]))


@pytest.mark.parametrize('R,N', [(8, 1000), (12, 25000)])
@pytest.mark.parametrize('method', ['mppi', 'cem_mppi'])
def test_result_does_not_depend_on_the_partition(built, method, R, N):
    """R robots in one handle == the same robots split 3 + (R - 3) (two 'ranks') == one handle per robot, bit for bit.
    (12, 25000): two-stage launches whose tail kernel runs with a different number of blocks per robot in every grouping.)"""
    cfg = qp.SamplerConfig.for_robot('go2', sampling_method=method, control_parametrization='linear_spline',
                                     num_parallel_computations=N, lanes_per_sample=1)
    st, rf, cs, ph, _ = _scenarios('go2', R)
    key = np.array([42, 9], np.uint32)

    def run(groups):
        outs = []
        for lo, hi in groups:
            b = BatchedSampling_MPC(cfg, hi - lo, robot_base=lo)
            outs.append(b.compute_control(st[lo:hi], rf[lo:hi], cs[lo:hi], np.zeros((hi - lo, 36), np.float32), key=key))
        return [np.concatenate([o[i] for o in outs]) for i in range(len(outs[0]))]

    whole = run([(0, R)])
    for groups in ([(0, 3), (3, R)], [(r, r + 1) for r in range(R)], [partition_robots(R, g, 4) for g in range(4)]):
        part = run(groups)
        for a, b in zip(whole, part):
            np.testing.assert_array_equal(a, b)


def test_robots_are_isolated(built):
    """Changing robot 2's state changes robot 2's result only."""
    R, N = 4, 800
    cfg = qp.SamplerConfig.for_robot('go2', sampling_method='mppi', num_parallel_computations=N)
    b = BatchedSampling_MPC(cfg, R)
    st, rf, cs, ph, _ = _scenarios('go2', R)
    key = np.array([42, 3], np.uint32)
    base = b.compute_control(st, rf, cs, np.zeros((R, 96), np.float32), key=key)
    st2 = st.copy()
    st2[2, 3:6] += 0.2
    pert = b.compute_control(st2, rf, cs, np.zeros((R, 96), np.float32), key=key)
    for r in range(R):
        same = np.array_equal(base[2][r], pert[2][r]) and base[3][r] == pert[3][r]
        assert same == (r != 2)


def test_batch_rejects_mixed_seeds(built):
    """Key word 0 is launch-uniform (kernel-parameter space): the struct entry point must refuse a batch that mixes seeds."""
    import ctypes as C
    from quadruped_pympc_b200 import _lib
    from quadruped_pympc_b200.sampling_mpc import make_c_config
    lib = _lib.load()
    cfg = qp.SamplerConfig.for_robot('go2', sampling_method='mppi', num_parallel_computations=256)
    c = make_c_config(cfg, 2)
    h = C.c_void_p()
    _lib.check(lib.qpmpc_create(C.byref(c), C.byref(h)))
    sc = synthetic.scenario('go2', 'trot', seed=0)
    keep = [sc['state'].astype(np.float32), sc['reference'].astype(np.float32),
            np.ascontiguousarray(sc['contact_sequence'], np.float32), np.zeros(96, np.float32)]
    fp = lambda a: a.ctypes.data_as(_lib._FP)
    ins, outs = (_lib.QpmpcTickIn * 2)(), (_lib.QpmpcTickOut * 2)()
    for r in range(2):
        i = ins[r]
        i.state, i.reference, i.contact_sequence, i.contact_ld, i.best = fp(keep[0]), fp(keep[1]), fp(keep[2]), 12, fp(keep[3])
        i.key[0], i.key[1] = 42 + r, 1
    rc = lib.qpmpc_tick_batched(h, ins, outs)
    assert rc == _lib.ERR_INVALID and b'share key[0]' in lib.qpmpc_last_error()
    lib.qpmpc_destroy(h)


@pytest.mark.parametrize('method,iters', [('mppi', 2), ('cem_mppi', 2), ('random_sampling', 1)])
def test_stateful_batch_step_equals_per_robot_host_loops(built, method, iters):
    """`BatchedSampling_MPC.step` (device-resident per-robot state, prep in the kernel) against one host-driven
    interface block per robot (single-robot batches with the same robot id), over ticks with contact changes."""
    R, N = 3, 1200
    cfg = qp.SamplerConfig.for_robot('go2', sampling_method=method, control_parametrization='cubic_spline',
                                     num_parallel_computations=N, lanes_per_sample=1)
    dev = BatchedSampling_MPC(cfg, R, robot_base=4)
    hosts = [BatchedSampling_MPC(cfg, 1, robot_base=4 + r) for r in range(R)]
    prev = np.ones((R, 4))
    P = dev.num_control_parameters
    cem = method == 'cem_mppi'
    for tick in range(4):
        scs = [synthetic.scenario('go2', 'trot', seed=50 + r, sim_steps=40 * tick + 13 * r) for r in range(R)]
        st = np.stack([s['state'] for s in scs]).astype(np.float32)
        rf = np.stack([s['reference'] for s in scs]).astype(np.float32)
        rf[:, 12:] += 0.02                                            # reference footholds differ from the feet
        cs = np.stack([s['contact_sequence'] for s in scs]).astype(np.float32)
        out_d = dev.step(st, rf, cs, iterations=iters)
        for r, hc in enumerate(hosts):
            cur = cs[r, :, 0]
            s_r = st[r].copy()
            for leg in range(4):                                       # S:588-595, S:612-625 on the host
                if cur[leg] == 0.0:
                    s_r[12 + 3 * leg:15 + 3 * leg] = rf[r, 12 + 3 * leg:15 + 3 * leg]
                if prev[r, leg] == 1 and cur[leg] == 0:
                    hc.best_control_parameters[0, leg * (P // 4):(leg + 1) * (P // 4)] = 0.0
            prev[r] = cur
            for it in range(iters):
                hc.with_newkey()
                sg = np.full((1, P), cfg.sigma_cem_mppi, np.float32) if (cem and it == 0) else hc.sigma_cem_mppi
                o = hc.compute_control(s_r[None], rf[r][None], cs[r][None], hc.best_control_parameters, sigma=sg)
                hc.best_control_parameters = o[2]
                if cem:
                    hc.sigma_cem_mppi = o[6]
            for k in range(6):
                np.testing.assert_array_equal(out_d[k][r], o[k][0])
            if cem:
                np.testing.assert_array_equal(out_d[6][r], o[6][0])


def test_stance_torque_epilogue(built):
    """GRFs * current_contact (I:225-230) and tau = -J^T GRF per leg (wb_interface.py:372-375) on the
    device-resident results, against the reference's numpy expressions in float64."""
    R, N = 6, 600
    cfg = qp.SamplerConfig.for_robot('go2', sampling_method='mppi', num_parallel_computations=N)
    b = BatchedSampling_MPC(cfg, R)
    st, rf, cs, ph, _ = _scenarios('go2', R)
    out = b.with_newkey().compute_control(st, rf, cs)
    rng = np.random.RandomState(4)
    jac = rng.randn(R, 4, 3, 3).astype(np.float32)             # feet_jac.leg[:, legs_qvel_idx.leg]
    contact = cs[:, :, 0]
    grf, tau = b.stance_torques(jac, contact)
    for r in range(R):
        for leg in range(4):
            F = out[0][r, 3 * leg:3 * leg + 3].astype(np.float64) * float(contact[r, leg])
            want = -np.matmul(jac[r, leg].astype(np.float64).T, F)
            np.testing.assert_allclose(grf[r, 3 * leg:3 * leg + 3], F, rtol=1e-7, atol=0)
            np.testing.assert_allclose(tau[r, 3 * leg:3 * leg + 3], want, rtol=1e-5, atol=1e-4)
    # single controller, direct I/O: the result packet is device-resident as well
    c = qp.Sampling_MPC(cfg)
    sc = synthetic.scenario('go2', 'trot', seed=0)
    o = c.with_newkey().jitted_compute_control(sc['state'], sc['reference'], sc['contact_sequence'], np.zeros(96, np.float32),
                                               c.master_key, sc['phase_signal'], 1.4, 0)
    g1, t1 = c.stance_torques(jac[0], sc['contact_sequence'][:, 0])
    for leg in range(4):
        F = o[0][3 * leg:3 * leg + 3].astype(np.float64) * float(sc['contact_sequence'][leg, 0])
        np.testing.assert_allclose(t1[3 * leg:3 * leg + 3], -np.matmul(jac[0, leg].astype(np.float64).T, F), rtol=1e-5, atol=1e-4)
        np.testing.assert_allclose(g1[3 * leg:3 * leg + 3], F, rtol=1e-7)


@pytest.mark.parametrize('method,adaptive', [('mppi', False), ('cem_mppi', False), ('random_sampling', True)])
@pytest.mark.parametrize('R', [64, 77, 300])
def test_chunk_pipelined_batch_equals_small_batches(built, method, adaptive, R):
    """R >= 64 robots with host arrays run as up to 8 chunks alternating between two streams (upload / kernels / download
    overlapped).  The result must be bit-identical to the same robots run in small un-pipelined handles (same robot ids),
    for ragged chunk sizes, stateless and stateful (2 iterations), and a failing robot must fail the call up front."""
    N = 400
    cfg = qp.SamplerConfig.for_robot('go2', sampling_method=method, control_parametrization='linear_spline',
                                     num_parallel_computations=N, optimize_step_freq=adaptive, lanes_per_sample=1)
    st, rf, cs, ph, _ = _scenarios('go2', R)
    key = np.array([42, 17], np.uint32)
    P = 36
    rng = np.random.RandomState(1)
    best0 = (0.2 * rng.randn(R, P)).astype(np.float32)
    kw = dict(timing=ph, nominal_step_frequency=1.4, optimize_swing=1)
    big = BatchedSampling_MPC(cfg, R)
    whole = big.compute_control(st, rf, cs, best0, key=key, **kw)
    parts = []
    for lo in range(0, R, 29):                                  # 29 < 64: the un-pipelined path
        hi = min(R, lo + 29)
        b = BatchedSampling_MPC(cfg, hi - lo, robot_base=lo)
        parts.append(b.compute_control(st[lo:hi], rf[lo:hi], cs[lo:hi], best0[lo:hi], key=key,
                                       timing=ph[lo:hi], nominal_step_frequency=1.4, optimize_swing=1))
    for i in range(len(whole)):
        np.testing.assert_array_equal(whole[i], np.concatenate([p[i] for p in parts]))
    np.testing.assert_array_equal(big.fetch_costs(R - 1), b.fetch_costs(R - 1 - lo))
    # stateful, two ticks of two iterations
    big2 = BatchedSampling_MPC(cfg, R)
    smalls = [(lo, min(R, lo + 29)) for lo in range(0, R, 29)]
    sm = [BatchedSampling_MPC(cfg, hi - lo, robot_base=lo) for lo, hi in smalls]
    for tick in range(2):
        st_t = st + 0.01 * tick
        w = big2.step(st_t, rf, cs, iterations=2, **kw)
        ps = [m.step(st_t[lo:hi], rf[lo:hi], cs[lo:hi], timing=ph[lo:hi], nominal_step_frequency=1.4, optimize_swing=1,
                     iterations=2) for m, (lo, hi) in zip(sm, smalls)]
        for i in range(len(w)):
            np.testing.assert_array_equal(w[i], np.concatenate([p[i] for p in ps]))
    # a bad contact table in the LAST chunk is refused before anything is launched: the carried state is untouched
    bad = cs.copy()
    bad[R - 1, 2, 3] = 0.5
    with pytest.raises(Exception):
        big2.step(st, rf, bad, iterations=1, **kw)
    again = big2.step(st + 0.02, rf, cs, iterations=2, **kw)
    ps = [m.step(st[lo:hi] + 0.02, rf[lo:hi], cs[lo:hi], timing=ph[lo:hi], nominal_step_frequency=1.4, optimize_swing=1,
                 iterations=2) for m, (lo, hi) in zip(sm, smalls)]
    np.testing.assert_array_equal(again[2], np.concatenate([p[2] for p in ps]))


@pytest.mark.parametrize('R,io_mode', [(2, 'direct'), (5, 'direct'), (1, 'memcpy')])
def test_graph_captured_tick_follows_a_changed_seed(built, R, io_mode):
    """Batches of 2 <= R < 64 and io_mode='memcpy' replay a CUDA graph captured on the first tick.  The Philox seed
    (key[0]) must not be frozen with it: a handle that ticks with key (42, 1) and then (43, 1) must return, for the
    second tick, what a fresh handle returns for (43, 1), and the tick must equal the oracle on generate_noise(key)."""
    N = 700
    kw = dict(sampling_method='mppi', control_parametrization='cubic_spline', num_parallel_computations=N, io_mode=io_mode)
    cfg = qp.SamplerConfig.for_robot('go2', **kw)
    st, rf, cs, ph, scs = _scenarios('go2', R)
    best0 = np.zeros((R, cfg.num_control_parameters), np.float32)
    used = BatchedSampling_MPC(cfg, R)
    first = used.compute_control(st, rf, cs, best0, key=np.array([42, 1], np.uint32))
    again = used.compute_control(st, rf, cs, best0, key=np.array([43, 1], np.uint32))
    fresh = BatchedSampling_MPC(cfg, R).compute_control(st, rf, cs, best0, key=np.array([43, 1], np.uint32))
    assert not np.array_equal(first[2], again[2])
    for a, b in zip(again, fresh):
        np.testing.assert_array_equal(a, b)
    orc = SamplingOracle(OracleConfig.for_robot('go2', **{k: v for k, v in kw.items() if k != 'io_mode'}))
    for r in range(R):
        noise = used.generate_noise(r, key=np.array([43, 1], np.uint32))
        ref = orc.compute_control(scs[r]['state'], scs[r]['reference'], scs[r]['contact_sequence'], best0[r], noise)
        assert again[5][r] == ref['best_index']
        np.testing.assert_allclose(again[3][r], ref['best_cost'], rtol=1e-5)
