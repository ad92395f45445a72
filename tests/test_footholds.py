"""Foothold reference (second half).
CPU: the oracle restatement against footholds produced by the reference's own FootholdReferenceGenerator
(tools/make_golden_footholds.py: 24 robots x 50 consecutive ticks, clips driven, non-zero manual CoM offsets).
GPU: the batched device kernel through the C ABI against the same fixtures (all 24 robots in one launch per tick, the
velocity-history rings carried on the device) and against the oracle after a history reset.
Tolerance 1e-12 absolute (float64 on both sides; sin / cos / sqrt and the 2 x 2 products may round differently)."""
import os

import numpy as np
import pytest

from oracle.sampling_oracle import FootholdReferenceOracle

GOLD = np.load(os.path.join(os.path.dirname(__file__), 'golden', 'gait_footholds.npz'))
TOL = 1e-12


def test_oracle_reproduces_the_reference_foothold_generator():
    R, T = GOLD['vel'].shape[:2]
    assert (R, T) == (24, 50)
    worst = 0.0
    for r in range(R):
        gen = FootholdReferenceOracle(float(GOLD['stance_time'][r]), GOLD['lift_off_z'][r], float(GOLD['hip_height'][r]),
                                      float(GOLD['hip_offset']), float(GOLD['gravity']))
        gen.com_pos_offset_b = GOLD['com_off_b'][r]
        for t in range(T):
            got = gen(GOLD['base_position'][r, t], GOLD['euler'][r, t], GOLD['vel'][r, t], GOLD['ref_vel'][r, t],
                      GOLD['hips'][r, t], float(GOLD['com_height'][r]))
            worst = max(worst, float(np.max(np.abs(got - GOLD['ref_feet'][r, t]))))
    assert worst < TOL, worst
    # the fixtures exercise both clips and the 20-entry window
    assert T > 20 and np.any(GOLD['com_off_b'] != 0)


def _by_offset(controller, rows, t, reset):
    """robots that share com_pos_offset_b go in one call (the offset is a per-generator attribute, one per call here)"""
    from quadruped_pympc_b200.gait import footholds_reference
    r = rows
    return footholds_reference(controller, GOLD['base_position'][r, t], GOLD['euler'][r, t], GOLD['vel'][r, t],
                               GOLD['ref_vel'][r, t], GOLD['hips'][r, t], GOLD['lift_off_z'][r], GOLD['stance_time'][r],
                               float(GOLD['com_height'][r[0]]), float(GOLD['hip_height'][r[0]]), float(GOLD['hip_offset']),
                               float(GOLD['gravity']), GOLD['com_off_b'][r[0]], reset_history=reset)


@pytest.mark.gpu
def test_device_footholds_reproduce_the_reference_generator(built):
    import quadruped_pympc_b200 as qp
    from quadruped_pympc_b200.batched import BatchedSampling_MPC
    R, T = GOLD['vel'].shape[:2]
    cfg = qp.SamplerConfig.for_robot('go2', sampling_method='mppi', num_parallel_computations=256)
    # com_height / hip_height / offset are scalars of a call: one single-robot controller per fixture robot
    worst = 0.0
    for r in range(0, R, 5):
        ctrl = qp.Sampling_MPC(cfg)
        for t in range(T):
            got = _by_offset(ctrl, np.array([r]), t, reset=(t == 0))
            worst = max(worst, float(np.max(np.abs(got[0] - GOLD['ref_feet'][r, t]))))
        del ctrl
    assert worst < TOL, worst
    # batched: R copies of one fixture robot with different inputs per slot = the R fixture rows of robots sharing scalars
    r0 = 1
    B = 9
    ctrl = BatchedSampling_MPC(cfg, B)
    oracles = []
    rng = np.random.RandomState(5)
    scale = rng.uniform(0.5, 1.5, B)
    for b in range(B):
        o = FootholdReferenceOracle(float(GOLD['stance_time'][r0]) * scale[b], GOLD['lift_off_z'][r0] + 0.01 * b,
                                    float(GOLD['hip_height'][r0]), 0.1, 9.81)
        o.com_pos_offset_b = GOLD['com_off_b'][3]
        oracles.append(o)
    from quadruped_pympc_b200.gait import footholds_reference
    for t in range(30):
        idx = (np.arange(B) * 2 + 1) % GOLD['vel'].shape[0]
        args = [GOLD[k][idx, t] for k in ('base_position', 'euler', 'vel', 'ref_vel', 'hips')]
        got = footholds_reference(ctrl, *args, np.stack([GOLD['lift_off_z'][r0] + 0.01 * b for b in range(B)]),
                                  float(GOLD['stance_time'][r0]) * scale, float(GOLD['com_height'][r0]),
                                  float(GOLD['hip_height'][r0]), 0.1, 9.81, GOLD['com_off_b'][3], reset_history=(t == 0))
        for b in range(B):
            want = oracles[b](args[0][b], args[1][b], args[2][b], args[3][b], args[4][b], float(GOLD['com_height'][r0]))
            worst = max(worst, float(np.max(np.abs(got[b] - want))))
    assert worst < TOL, worst
    # a reset empties the rings: the same inputs give the first-tick answer again
    a = footholds_reference(ctrl, *args, np.zeros((B, 4)), 0.3, 0.3, 0.3, reset_history=True)
    b2 = footholds_reference(ctrl, *args, np.zeros((B, 4)), 0.3, 0.3, 0.3, reset_history=True)
    np.testing.assert_array_equal(a, b2)
    c = footholds_reference(ctrl, *args, np.zeros((B, 4)), 0.3, 0.3, 0.3)
    np.testing.assert_array_equal(a, c)       # mean of two identical entries = the entry
