"""GPU: the two-samples-per-thread packed-FP32 kernel (csrc/qpmpc_x2.cuh, lanes_per_sample = 2).

The rollout arithmetic is one template (csrc/qpmpc_math.cuh) instantiated for float (tick_kernel) and float2
(tick_kernel_x2: FFMA2 / FMUL2 / FADD2), every operation an explicit round-to-nearest intrinsic, and both kernels draw the
same Philox rows -- so every sample must get the SAME BITS from both, whatever the tile boundaries."""
import numpy as np
import pytest

import quadruped_pympc_b200 as qp
from quadruped_pympc_b200 import synthetic
from quadruped_pympc_b200.batched import BatchedSampling_MPC
from helpers import METHODS, PARAMS, COST_RTOL

pytestmark = pytest.mark.gpu


def _tick(lanes, method, param, N, robot='go2', gait='trot', seed=11, ticks=2):
    cfg = qp.SamplerConfig.for_robot(robot, sampling_method=method, control_parametrization=param,
                                     num_parallel_computations=N, lanes_per_sample=lanes)
    ctrl = qp.Sampling_MPC(cfg)
    sc = synthetic.scenario(robot, gait, seed=seed)
    best = np.linspace(-1, 1, ctrl.num_control_parameters).astype(np.float32)
    outs = []
    for _ in range(ticks):
        ctrl.with_newkey()
        args = (np.full(ctrl.num_control_parameters, 2.0, np.float32),) if method == 'cem_mppi' else ()
        out = ctrl.jitted_compute_control(sc['state'], sc['reference'], sc['contact_sequence'], best, ctrl.master_key,
                                          *args, sc['phase_signal'], sc['nominal_step_freq'], 0)
        outs.append((np.asarray(out[6]).copy(), ctrl.last_best_index, out[3].copy(), out[0].copy(),
                     ctrl.last_elite_indices.copy() if method == 'cem_mppi' else None))
        best = out[3]
    return ctrl.kernel_name, outs


@pytest.mark.parametrize('N', [1, 2, 255, 257, 1000, 5001, 70000])
@pytest.mark.parametrize('method', METHODS)
@pytest.mark.parametrize('param', PARAMS)
def test_x2_costs_are_bit_identical_to_the_sample_per_thread_kernel(built, param, method, N):
    n1, o1 = _tick(1, method, param, N)
    n2, o2 = _tick(2, method, param, N)
    # zero-order (P = 144): two 256-sample tiles do not fit one SM, the request falls back to the plain kernel
    assert n1 == 'tick_kernel' and n2 == ('tick_kernel' if param == 'zero_order' else 'tick_kernel_x2')
    for (c1, i1, b1, g1, e1), (c2, i2, b2, g2, e2) in zip(o1, o2):
        np.testing.assert_array_equal(c2, c1)                       # every sample, bit for bit
        assert i2 == i1
        if method == 'random_sampling':
            np.testing.assert_array_equal(b2, b1)                   # argmin pick: no reduction order involved
            np.testing.assert_array_equal(g2, g1)
        else:                                                       # softmax sums grouped by 256- instead of 128-sample tiles
            np.testing.assert_allclose(b2, b1, rtol=1e-4, atol=2e-5)
            np.testing.assert_allclose(g2, g1, rtol=1e-4, atol=2e-4)
        if e1 is not None:
            np.testing.assert_array_equal(e2, e1)
        if method != 'random_sampling':
            break                                                   # the second tick starts from (slightly) different knots


def test_x2_is_opt_in_and_fixed_gait_only(built):
    def name(N, lanes=0, robots=None, **kw):
        cfg = qp.SamplerConfig.for_robot('go2', sampling_method='mppi', num_parallel_computations=N, lanes_per_sample=lanes, **kw)
        if robots is None:
            return qp.Sampling_MPC(cfg).kernel_name
        from quadruped_pympc_b200 import _lib
        return _lib.load().qpmpc_kernel_name(BatchedSampling_MPC(cfg, robots)._ensure()).decode()

    assert name(1000000) == 'tick_kernel' and name(10000, robots=8) == 'tick_kernel'   # never picked automatically (slower)
    assert name(5000, lanes=2) == 'tick_kernel_x2' and name(10000, lanes=2, robots=8) == 'tick_kernel_x2'
    assert name(5000, lanes=2, optimize_step_freq=True) == 'tick_kernel'                 # gait-adaptive: not packed
    assert name(100000, lanes=2, control_parametrization='zero_order') == 'tick_kernel'   # P = 144: two tiles do not fit one SM


def test_x2_batched_robots_equal_the_per_robot_kernel(built):
    """8 robots x 10k in one packed launch: per robot the costs must be the bits the sample-per-thread kernel produces."""
    R, N = 8, 10000
    out = {}
    for lanes in (1, 2):
        cfg = qp.SamplerConfig.for_robot('go2', sampling_method='mppi', control_parametrization='cubic_spline',
                                         num_parallel_computations=N, lanes_per_sample=lanes)
        b = BatchedSampling_MPC(cfg, R, robot_base=5)
        scs = [synthetic.scenario('go2', 'trot' if r % 2 == 0 else 'pace', seed=r) for r in range(R)]
        st = np.stack([s['state'] for s in scs]).astype(np.float32)
        rf = np.stack([s['reference'] for s in scs]).astype(np.float32)
        cs = np.stack([s['contact_sequence'] for s in scs]).astype(np.float32)
        b.with_newkey()
        res = b.compute_control(st, rf, cs, np.zeros((R, 96), np.float32))
        out[lanes] = (res, [b.fetch_costs(r) for r in range(R)])
    for r in range(R):
        np.testing.assert_array_equal(out[2][1][r], out[1][1][r])
    np.testing.assert_array_equal(out[2][0][5], out[1][0][5])            # best indices
    np.testing.assert_allclose(out[2][0][2], out[1][0][2], rtol=1e-4, atol=2e-5)
