"""Gait look-ahead (first half): contact tables from phase clocks.
CPU: the oracle restatement against tables produced by the reference's own PeriodicGaitGenerator (tools/make_golden_pgg.py).
GPU: the device kernel against the same fixtures (exact 0/1 tables), batched against the oracle, and a tick fed by the
device-resident tables against the same tick fed by host tables."""
import os

import numpy as np
import pytest

from oracle.sampling_oracle import gait_lookahead_contact_table as pgg_contact_sequence

GOLD = np.load(os.path.join(os.path.dirname(__file__), 'golden', 'gait_pgg.npz'))
H = 12


def _case(i):
    n = int(GOLD['n_dts'][i])
    return dict(phase=GOLD['phase'][i], init=GOLD['init'][i], freq=float(GOLD['freq'][i]), duty=float(GOLD['duty'][i]),
                offset=GOLD['offset'][i], dts=list(GOLD['dts'][i][:n]), lens=[int(x) for x in GOLD['lens'][i][:n]],
                table=GOLD['table'][i])


def test_oracle_reproduces_the_reference_gait_generator():
    assert len(GOLD['freq']) == 160
    for i in range(len(GOLD['freq'])):
        c = _case(i)
        got = pgg_contact_sequence(c['phase'], c['init'], c['freq'], c['duty'], c['offset'], H, c['dts'], c['lens'])
        np.testing.assert_array_equal(got, c['table'])
    assert 0.3 < GOLD['table'].mean() < 0.95 and GOLD['init'].any()


@pytest.mark.gpu
def test_device_tables_reproduce_the_reference_gait_generator(built):
    import quadruped_pympc_b200 as qp
    from quadruped_pympc_b200.gait import contact_sequences_from_phase
    ctrl = qp.Sampling_MPC(qp.SamplerConfig.for_robot('go2', sampling_method='mppi', num_parallel_computations=256))
    for i in range(len(GOLD['freq'])):
        c = _case(i)
        got = contact_sequences_from_phase(ctrl, c['phase'], c['freq'], c['duty'], c['offset'], init=c['init'],
                                           contact_sequence_dts=c['dts'], contact_sequence_lenghts=c['lens'])
        np.testing.assert_array_equal(got[0], c['table'].astype(np.float32))


@pytest.mark.gpu
@pytest.mark.parametrize('R', [37, 90])          # 90: the chunk-pipelined batched path
@pytest.mark.parametrize('method', ['mppi', 'cem_mppi'])
def test_batched_tables_feed_the_tick_from_the_device(built, method, R):
    import quadruped_pympc_b200 as qp
    from quadruped_pympc_b200 import synthetic
    from quadruped_pympc_b200.batched import BatchedSampling_MPC
    from quadruped_pympc_b200.gait import contact_sequences_from_phase, PHASE_OFFSETS
    cfg = qp.SamplerConfig.for_robot('go2', sampling_method=method, num_parallel_computations=500, lanes_per_sample=1)
    a, b = BatchedSampling_MPC(cfg, R), BatchedSampling_MPC(cfg, R)
    rng = np.random.RandomState(3)
    scs = [synthetic.scenario('go2', 'trot', seed=r) for r in range(R)]
    st = np.stack([s['state'] for s in scs]).astype(np.float32)
    rf = np.stack([s['reference'] for s in scs]).astype(np.float32)
    for tick in range(3):
        phase = rng.rand(R, 4)
        freq = rng.choice([1.3, 1.4, 2.0], R)
        tables = contact_sequences_from_phase(a, phase, freq, 0.65, 'trot')
        for r in range(R):
            want = pgg_contact_sequence(phase[r], None, float(freq[r]), 0.65, PHASE_OFFSETS['trot'], H, [cfg.dt], [H])
            np.testing.assert_array_equal(tables[r], want.astype(np.float32))
        tables[:, :, 0][tables[:, :, 0].sum(axis=1) == 0, 0] = 1.0   # (the comparison below needs no airborne start)
        if not np.array_equal(tables, contact_sequences_from_phase(a, phase, freq, 0.65, 'trot')):
            out_a = a.step(st, rf, tables)                             # tables were edited: host tables on both sides
        else:
            out_a = a.step(st, rf, None)                               # device-resident tables, nothing uploaded
        out_b = b.step(st, rf, tables)
        for x, y in zip(out_a, out_b):
            np.testing.assert_array_equal(x, y)
    # the armed tables are used once: a second call without tables is an error
    with pytest.raises(Exception, match='contact_sequence'):
        a.step(st, rf, None)
