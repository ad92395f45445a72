"""GPU: the CUDA tick (through the C ABI) against the golden vectors the reference's own source produced
(tests/golden/, see tools/make_golden.py), with the reference's raw draws injected as the noise tensor.

Tolerances are BASELINE.json's: sample costs rel 1e-5; returned GRFs / knots rel 1e-4; argmin exact."""
import numpy as np
import pytest

import quadruped_pympc_b200 as qp
from golden_util import Fixture, tick_fixtures, fixture_id
from helpers import COST_RTOL, KNOT_RTOL, knot_atol

pytestmark = pytest.mark.gpu
FIXTURES = tick_fixtures()


@pytest.mark.parametrize('lanes', [1, 4, 14])
@pytest.mark.parametrize('path', FIXTURES, ids=fixture_id)
def test_cuda_tick_reproduces_reference(built, path, lanes):
    fx = Fixture(path)
    orc = fx.oracle()
    ctrl = qp.Sampling_MPC(qp.SamplerConfig.for_robot(fx.robot, lanes_per_sample=lanes, **fx.config_kwargs()))
    sc, z = fx.scenario(), fx.z
    noise, freqs, sigma = fx.noise(orc), fx.step_freqs(orc), fx.sigma()
    tail = (sc['phase_signal'], sc['nominal_step_freq'], int(z['optimize_swing']))
    args = ((sigma,) + tail) if fx.method == 'cem_mppi' else tail
    out = ctrl.jitted_compute_control(sc['state'], sc['reference'], sc['contact_sequence'], z['best0'], ctrl.master_key,
                                      *args, noise=noise, step_freqs=freqs)
    exp = fx.expected()
    costs = np.asarray(out[6])
    rel = np.max(np.abs(costs - exp['costs']) / np.abs(exp['costs']))
    assert rel < COST_RTOL, f'cost rel err {rel:.3e}'
    assert ctrl.last_best_index == exp['best_index']
    np.testing.assert_allclose(out[4], exp['best_cost'], rtol=COST_RTOL)
    weights = None
    if fx.method != 'random_sampling':
        e = np.exp(-(exp['costs'].astype(np.float64) - float(exp['best_cost'])))
        weights = e / e.sum()
    atol = knot_atol(exp['costs'], costs, weights, noise)
    np.testing.assert_allclose(out[3], exp['best_control_parameters'], rtol=KNOT_RTOL, atol=atol)
    np.testing.assert_allclose(out[0], exp['nmpc_GRFs'], rtol=KNOT_RTOL, atol=atol)
    np.testing.assert_allclose(out[2], exp['nmpc_predicted_state'], rtol=KNOT_RTOL, atol=1e-6)
    if not np.isnan(exp['best_freq']):
        np.testing.assert_allclose(out[5], exp['best_freq'], rtol=1e-6)
    if fx.method == 'cem_mppi':
        np.testing.assert_allclose(out[7], exp['new_sigma_cem_mppi'], rtol=KNOT_RTOL, atol=1e-6)


@pytest.mark.parametrize('lanes', [1, 4, 14])
@pytest.mark.parametrize('path', FIXTURES, ids=fixture_id)
def test_cuda_rollout_seam_reproduces_reference(built, path, lanes):
    """jit_vectorized_rollout (S:176-177 / A:186-187): explicit parameters in, raw costs out."""
    fx = Fixture(path)
    z = fx.z
    params = z['seam_params']
    kw = fx.config_kwargs()
    kw['num_parallel_computations'] = params.shape[0]
    ctrl = qp.Sampling_MPC(qp.SamplerConfig.for_robot(fx.robot, lanes_per_sample=lanes, **kw))
    if fx.adaptive:
        got = ctrl.jit_vectorized_rollout(z['state'], z['reference'], z['phase_signal'], params, z['seam_freqs'])
    else:
        got = ctrl.jit_vectorized_rollout(z['state'], z['reference'], params, z['contact_sequence'])
    np.testing.assert_allclose(got, z['seam_costs'], rtol=COST_RTOL)
