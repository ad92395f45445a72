"""Pin the oracle (NumPy restatement + C twin) against golden vectors produced by the REFERENCE'S OWN SOURCE.

The fixtures under tests/golden/ were written by tools/make_golden.py, which executes the reference's unmodified
controllers/sampling/*.py and helpers/periodic_gait_generator_jax.py under a NumPy stand-in for jax (oracle/jaxshim;
jax itself is not installable in this image).  Here the oracle is fed the very same raw random draws and must
reproduce what the reference returned.  CPU only -- no GPU, no /root/reference at run time.
"""
import numpy as np
import pytest

from golden_util import Fixture, tick_fixtures, fixture_id, GOLDEN_DIR
from oracle import c_oracle
from oracle.sampling_oracle import OracleConfig, SamplingOracle

# The oracle and the shim-executed reference are both float32 NumPy, so they agree far inside the north_star
# tolerance (1e-5 on costs); the residual is the reference's dense e^T Q e through BLAS vs the diagonal sum.
COST_RTOL = 2e-6
VEC_RTOL = 1e-5

FIXTURES = tick_fixtures()


def test_fixtures_present():
    assert len(FIXTURES) >= 24, f'golden fixtures missing under {GOLDEN_DIR} (run tools/make_golden.py)'


def _check(fx, got, exp, cost_rtol=COST_RTOL, vec_rtol=VEC_RTOL):
    rel = np.max(np.abs(got['costs'] - exp['costs']) / np.maximum(np.abs(exp['costs']), 1e-12))
    assert rel < cost_rtol, f'cost rel err {rel:.3e}'
    assert got['best_index'] == exp['best_index']
    np.testing.assert_allclose(got['best_cost'], exp['best_cost'], rtol=cost_rtol)
    np.testing.assert_allclose(got['best_control_parameters'], exp['best_control_parameters'], rtol=vec_rtol, atol=2e-5)
    np.testing.assert_allclose(got['nmpc_GRFs'], exp['nmpc_GRFs'], rtol=vec_rtol, atol=2e-5)
    np.testing.assert_allclose(got['nmpc_predicted_state'], exp['nmpc_predicted_state'], rtol=vec_rtol, atol=1e-6)
    if not np.isnan(exp['best_freq']):
        np.testing.assert_allclose(got['best_freq'], exp['best_freq'], rtol=1e-6)
    if 'new_sigma_cem_mppi' in exp:
        np.testing.assert_allclose(got['new_sigma_cem_mppi'], exp['new_sigma_cem_mppi'], rtol=vec_rtol, atol=1e-6)


@pytest.mark.parametrize('path', FIXTURES, ids=fixture_id)
def test_numpy_oracle_reproduces_reference_tick(path):
    fx = Fixture(path)
    orc = fx.oracle()
    sc = fx.scenario()
    got = orc.compute_control(sc['state'], sc['reference'], sc['contact_sequence'], fx.z['best0'], fx.noise(orc),
                              sigma=fx.sigma(), timing=sc['phase_signal'],
                              nominal_step_frequency=sc['nominal_step_freq'],
                              optimize_swing=int(fx.z['optimize_swing']), step_frequencies=fx.step_freqs(orc))
    _check(fx, got, fx.expected())


@pytest.mark.parametrize('path', FIXTURES, ids=fixture_id)
def test_numpy_oracle_reproduces_reference_rollout_seam(path):
    """jit_vectorized_rollout (S:176-177 / A:186-187) on explicit parameters."""
    fx = Fixture(path)
    z = fx.z
    params = z['seam_params']
    kw = fx.config_kwargs()
    kw['num_parallel_computations'] = params.shape[0]
    orc = SamplingOracle(OracleConfig.for_robot(fx.robot, **kw))
    if fx.adaptive:
        got = orc.compute_rollout_adaptive(z['state'], z['reference'], z['phase_signal'], params, z['seam_freqs'])
    else:
        got = orc.compute_rollout(z['state'], z['reference'], params, z['contact_sequence'])
    np.testing.assert_allclose(got, z['seam_costs'], rtol=COST_RTOL)


@pytest.mark.parametrize('path', FIXTURES, ids=fixture_id)
def test_c_oracle_reproduces_reference_tick(path, built):
    fx = Fixture(path)
    orc = fx.oracle()
    corc = c_oracle.COracle(orc.cfg)
    sc = fx.scenario()
    got = corc.compute_control(sc['state'], sc['reference'], sc['contact_sequence'], fx.z['best0'], fx.noise(orc),
                               timing=sc['phase_signal'], freqs=fx.step_freqs(orc))
    _check(fx, got, fx.expected())


def test_float64_twin_bounds_reference_rounding():
    """The float64 twin of the oracle vs the float32 reference output: the float32 error budget of the path.

    Gait-adaptive rollouts contain a genuine float32 discontinuity: the in-rollout phase clock compares `t < duty` and
    `t >= 1` (G:73-87), and for some (phase, frequency) pairs the float32 and float64 clocks land on opposite sides.
    Those samples get a different contact table, hence a different cost; they are excluded here (and are exactly why
    the CUDA kernel keeps the reference's float32 operation order for the clock)."""
    worst = 0.0
    for path in FIXTURES:
        fx = Fixture(path)
        o32, o64 = fx.oracle(), fx.oracle(np.float64)
        sc = fx.scenario()
        noise = fx.noise(o32)
        freqs = fx.step_freqs(o32)
        got = o64.compute_control(sc['state'], sc['reference'], sc['contact_sequence'], fx.z['best0'], noise,
                                  sigma=fx.sigma(), timing=sc['phase_signal'],
                                  nominal_step_frequency=sc['nominal_step_freq'],
                                  optimize_swing=int(fx.z['optimize_swing']), step_frequencies=freqs)
        exp = fx.expected()
        rel = np.abs(got['costs'] - exp['costs']) / np.abs(exp['costs'])
        if fx.adaptive:
            same = np.all(o32.pgg_contact_sequence(sc['phase_signal'], freqs) ==
                          o64.pgg_contact_sequence(sc['phase_signal'], freqs), axis=(0, 1))
            assert same.mean() > 0.5
            rel = rel[same]
        worst = max(worst, rel.max())
    assert worst < 1e-5, worst       # the north_star's cost tolerance is wide enough for float32 round-off


def test_host_prepare_and_shift_match_reference():
    """prepare_state_and_reference (S:563-627) and shift_solution (S:513-561) of the product's host class."""
    import quadruped_pympc_b200 as qp
    z = np.load(f'{GOLDEN_DIR}/ref_host.npz')
    st = {k[3:]: z[k] for k in z.files if k.startswith('in_') and not k.startswith('in_ref_')}
    rf = {k[3:]: z[k] for k in z.files if k.startswith('in_ref_')}
    for param in ('zero_order', 'linear_spline', 'cubic_spline'):
        ctrl = qp.Sampling_MPC(qp.SamplerConfig.for_robot('go2', sampling_method='mppi', control_parametrization=param,
                                                          num_parallel_computations=8))
        ctrl.best_control_parameters = z[f'{param}_best_in'].copy()
        s24, r24 = ctrl.prepare_state_and_reference(dict(st), dict(rf), z['current_contact'], z['previous_contact'])
        np.testing.assert_array_equal(np.asarray(s24, np.float64), z[f'{param}_state24'])
        np.testing.assert_array_equal(np.asarray(r24, np.float64), z[f'{param}_ref24'])
        np.testing.assert_array_equal(ctrl.best_control_parameters, z[f'{param}_best_after_prepare'])
        shifted = ctrl.shift_solution(z[f'{param}_best_in'].copy(), 1.0 / 100)
        np.testing.assert_allclose(shifted, z[f'{param}_shifted'], rtol=1e-6, atol=1e-7)


def test_pgg_contact_sequence_matches_reference():
    """PeriodicGaitGeneratorJax.compute_contact_sequence (G:136-151) vs the oracle's in-rollout gait generator."""
    z = np.load(f'{GOLDEN_DIR}/ref_host.npz')
    orc = SamplingOracle(OracleConfig.for_robot('go2', optimize_step_freq=True))
    freqs = np.array([1.4, 2.0, 2.4], np.float32)
    cs = orc.pgg_contact_sequence(z['pgg_t0'], freqs)            # (4, H, 3)
    for i in range(3):
        np.testing.assert_array_equal(cs[:, :, i], z[f'pgg_contact_{i}'])
