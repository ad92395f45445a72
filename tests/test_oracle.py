"""CPU: properties of the oracle itself (beyond the golden vectors of test_golden.py) and of the host mirror of the
product's Philox noise generator."""
import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

from oracle import c_oracle
from oracle.sampling_oracle import OracleConfig, SamplingOracle, twin64
from quadruped_pympc_b200 import synthetic

PARAMS = ['zero_order', 'linear_spline', 'cubic_spline']


def test_philox4x32_10_known_answers(built):
    """Random123's published known-answer vectors for philox4x32-10."""
    kat = [([0] * 4, [0] * 2, [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]),
           ([0xffffffff] * 4, [0xffffffff] * 2, [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]),
           ([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0],
            [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1])]
    for ctr, key, want in kat:
        assert [int(x) for x in c_oracle.philox4x32_10(ctr, key)] == want


def test_noise_mirror_layout_and_moments(built):
    """Row a9: row 0 is zero; random_sampling blocks [1,1+k) and [1+k,1+2k) share their normals (same key);
    the uniform block stays inside [-s2, s2); the Gaussian rows have the right scale."""
    cfg = OracleConfig.for_robot('go2', sampling_method='random_sampling', num_parallel_computations=3000)
    c = c_oracle.COracle(cfg)
    nz = c.noise(42, 7)
    k = 3000 // 3
    assert np.all(nz[0] == 0)
    np.testing.assert_allclose(nz[1:1 + k] * (3.0 / 0.2), nz[1 + k:1 + 2 * k], rtol=2e-6)
    assert np.all(np.abs(nz[1 + 2 * k:]) <= 10.0) and nz[1 + 2 * k:].std() == pytest.approx(10 / np.sqrt(3), rel=0.02)
    assert nz[1 + k:1 + 2 * k].std() == pytest.approx(3.0, rel=0.02)
    cfg = OracleConfig.for_robot('go2', sampling_method='mppi', num_parallel_computations=2000)
    nz = c_oracle.COracle(cfg).noise(42, 1)
    assert np.all(nz[0] == 0) and nz[1:].std() == pytest.approx(3.0, rel=0.02) and abs(nz[1:].mean()) < 0.02
    assert not np.array_equal(nz, c_oracle.COracle(cfg).noise(42, 2))          # the tick counter changes the stream
    assert not np.array_equal(nz, c_oracle.COracle(cfg).noise(42, 1, robot=1))  # so does the robot id


def _normal_distribution_checks(z):
    """z: (rows, P) standard normals as the generator lays them out (8 per Philox block, two per 32-bit word)."""
    from scipy import stats
    flat = z.ravel()
    n = flat.size
    assert abs(flat.mean()) < 4.0 / np.sqrt(n) and abs(flat.std() - 1.0) < 4.0 / np.sqrt(2 * n)
    assert abs(np.mean(flat ** 4) - 3.0) < 0.03                                  # kurtosis of a normal
    d, _ = stats.kstest(flat[:2000000], 'norm')
    assert d < 1.63 / np.sqrt(min(n, 2000000))                                   # Kolmogorov-Smirnov, alpha = 0.01
    assert np.max(np.abs(flat)) <= 4.72                                          # 16-bit radius: |z| <= sqrt(32 ln 2)
    tail = np.mean(np.abs(flat) > 3.0)
    assert abs(tail - 0.0026998) < 4 * np.sqrt(0.0027 / n)                       # mass beyond 3 sigma
    # the eight normals of one block (and the two of one word) are uncorrelated, also in their squares
    oct = z[:, :8 * (z.shape[1] // 8)].reshape(-1, 8)
    c1 = np.corrcoef(oct, rowvar=False) - np.eye(8)
    c2 = np.corrcoef(oct ** 2, rowvar=False) - np.eye(8)
    assert np.max(np.abs(c1)) < 5.0 / np.sqrt(oct.shape[0]) and np.max(np.abs(c2)) < 5.0 / np.sqrt(oct.shape[0])
    # consecutive rows (counter word 1 differs by one) are uncorrelated
    assert abs(np.corrcoef(z[:-1].ravel(), z[1:].ravel())[0, 1]) < 5.0 / np.sqrt(n)


def test_noise_mirror_is_standard_normal(built):
    """The generator draws 8 normals per Philox4x32-10 block (16-bit radius, 16-bit angle per Box-Muller pair): moments,
    Kolmogorov-Smirnov distance, tail mass and independence inside a block of the host mirror (the device kernel is held
    to the mirror within 2e-4 in tests/test_gpu_parity.py, and runs the same checks on its own draw)."""
    cfg = OracleConfig.for_robot('go2', sampling_method='mppi', control_parametrization='cubic_spline',
                                 num_parallel_computations=30001)
    z = c_oracle.COracle(cfg).noise(42, 3)[1:] / np.float32(3.0)
    _normal_distribution_checks(z)


@pytest.mark.parametrize('param', PARAMS)
def test_zero_noise_row_reproduces_cost_of_best(param):
    """Row 0 of the noise is zero, so costs[0] is the cost of the previous best (S:647)."""
    orc = SamplingOracle(OracleConfig.for_robot('go2', sampling_method='mppi', control_parametrization=param,
                                                num_parallel_computations=16))
    sc = synthetic.scenario('go2', 'trot', seed=5)
    rng = np.random.RandomState(0)
    best = rng.randn(orc.P).astype(np.float32)
    noise = orc.assemble_noise_gaussian(rng.randn(15, orc.P), np.float32(3))
    out = orc.compute_control(sc['state'], sc['reference'], sc['contact_sequence'], best, noise)
    alone = orc.compute_rollout(sc['state'], sc['reference'], best[None, :], sc['contact_sequence'])
    assert out['costs'][0] == alone[0]


def test_hover_full_stance_analytic_step():
    """All legs in stance, zero knots, level body at rest: every leg carries m*g/4, the net wrench about the CoM of a
    symmetric stance is zero, so the state stays put and the cost is the constant tracking error."""
    orc = SamplingOracle(OracleConfig.for_robot('go2', control_parametrization='cubic_spline', num_parallel_computations=1))
    s = np.zeros(24); s[2] = 0.3
    feet = np.array([[0.2, 0.13, 0], [0.2, -0.13, 0], [-0.2, 0.13, 0], [-0.2, -0.13, 0]], float)
    s[12:] = feet.ravel()
    r = s.copy()
    out = orc.compute_control(s, r, np.ones((4, 12)), np.zeros(orc.P, np.float32), np.zeros((1, orc.P), np.float32))
    np.testing.assert_allclose(out['nmpc_GRFs'].reshape(4, 3)[:, 2], 15.019 * 9.81 / 4, rtol=1e-6)
    np.testing.assert_allclose(out['nmpc_GRFs'].reshape(4, 3)[:, :2], 0, atol=1e-7)
    np.testing.assert_allclose(out['nmpc_predicted_state'], s, atol=2e-6)
    assert out['costs'][0] < 1e-6


def test_clamp_invariants():
    """S:270-314: fz in [fmin, fmax], |fx|,|fy| <= mu*fz, NaN fz -> fmin."""
    orc = SamplingOracle(OracleConfig.for_robot('go2'))
    rng = np.random.RandomState(1)
    F = [[rng.randn(500).astype(np.float32) * 300 for _ in range(3)] for _ in range(4)]
    F[0][2][:5] = np.nan
    out = orc.enforce_force_constraints(F)
    for l in range(4):
        fx, fy, fz = out[l]
        assert np.all(fz >= orc.f_z_min) and np.all(fz <= orc.f_z_max)
        assert np.all(np.abs(fx) <= orc.mu * fz + 1e-6) and np.all(np.abs(fy) <= orc.mu * fz + 1e-6)
    assert np.all(out[0][2][:5] == orc.f_z_min)


def test_no_stance_leg_goes_through_the_nan_path():
    """mass*g/0 = inf (S:377-380); inf*0 = NaN; the clamp's `where` order turns NaN fz into fz_min -> finite cost."""
    orc = SamplingOracle(OracleConfig.for_robot('go2', sampling_method='mppi', num_parallel_computations=8))
    sc = synthetic.scenario('go2', 'trot', seed=1)
    cs = sc['contact_sequence'].copy(); cs[:, 3] = 0
    noise = orc.assemble_noise_gaussian(np.random.RandomState(0).randn(7, orc.P), np.float32(3))
    out = orc.compute_control(sc['state'], sc['reference'], cs, np.zeros(orc.P, np.float32), noise)
    assert np.all(np.isfinite(out['costs']))


def test_saturation_and_argmin_tie():
    """NaN / Inf costs -> 1e6 (S:686-687); argmin = first minimum (S:690)."""
    orc = SamplingOracle(OracleConfig.for_robot('go2'))
    c = orc._saturate(np.array([np.nan, np.inf, -np.inf, 3.0, 3.0], np.float32))
    assert list(c) == [1e6, 1e6, 1e6, 3.0, 3.0] and int(np.argmin(c)) == 3


@settings(max_examples=15, deadline=None)
@given(st.integers(0, 10_000), st.sampled_from(PARAMS), st.booleans())
def test_permuting_samples_permutes_costs(seed, param, adaptive):
    orc = SamplingOracle(OracleConfig.for_robot('aliengo', control_parametrization=param, optimize_step_freq=adaptive,
                                                num_parallel_computations=24))
    sc = synthetic.scenario('aliengo', 'pace', seed=seed % 50)
    rng = np.random.RandomState(seed)
    params = (3 * rng.randn(24, orc.P)).astype(np.float32)
    freqs = np.array([1.4, 2.0, 2.4], np.float32)[rng.randint(0, 3, 24)]
    perm = rng.permutation(24)
    if adaptive:
        a = orc.compute_rollout_adaptive(sc['state'], sc['reference'], sc['phase_signal'], params, freqs)
        b = orc.compute_rollout_adaptive(sc['state'], sc['reference'], sc['phase_signal'], params[perm], freqs[perm])
    else:
        a = orc.compute_rollout(sc['state'], sc['reference'], params, sc['contact_sequence'])
        b = orc.compute_rollout(sc['state'], sc['reference'], params[perm], sc['contact_sequence'])
    assert np.array_equal(a[perm], b)


def test_srbd_against_independent_statement():
    """The same SRBD equations as stated independently by the gradient controller's model
    (controllers/gradient/nominal/centroidal_model_nominal.py:201-272), here with numpy.linalg in float64."""
    cfg = OracleConfig.for_robot('go2')
    orc = twin64(cfg)
    rng = np.random.RandomState(3)
    x = rng.randn(24) * 0.3; x[6:8] *= 0.2
    F = rng.randn(12) * 20
    stance = np.array([1.0, 0.0, 1.0, 1.0])
    got = np.array([float(np.asarray(v).reshape(-1)[0]) for v in
                    orc.fd([np.array([v]) for v in x], [0.0] * 12 + [np.array([f]) for f in F], list(stance))])
    m, I = cfg.mass, cfg.inertia
    roll, pitch, yaw = x[6:9]; w = x[9:12]
    lin_acc = (F.reshape(4, 3) * stance[:, None]).sum(0) / m + np.array([0, 0, -9.81])
    tau = sum(np.cross(x[12 + 3 * l:15 + 3 * l] - x[0:3], F[3 * l:3 * l + 3]) * stance[l] for l in range(4))
    E = np.array([[1, 0, -np.sin(pitch)], [0, np.cos(roll), np.cos(pitch) * np.sin(roll)],
                  [0, -np.sin(roll), np.cos(pitch) * np.cos(roll)]])
    Rx = np.array([[1, 0, 0], [0, np.cos(roll), np.sin(roll)], [0, -np.sin(roll), np.cos(roll)]])
    Ry = np.array([[np.cos(pitch), 0, -np.sin(pitch)], [0, 1, 0], [np.sin(pitch), 0, np.cos(pitch)]])
    Rz = np.array([[np.cos(yaw), np.sin(yaw), 0], [-np.sin(yaw), np.cos(yaw), 0], [0, 0, 1]])
    b_R_w = Rx @ Ry @ Rz
    ang_acc = np.linalg.solve(I, b_R_w @ tau - np.cross(w, I @ w))
    want = np.concatenate([x[3:6], lin_acc, np.linalg.solve(E, w), ang_acc])
    np.testing.assert_allclose(got, want, rtol=1e-9, atol=1e-10)


def test_c_twin_matches_numpy_oracle_at_full_size(built):
    """BASELINE config 2 size (N = 10,000): C twin vs NumPy oracle on the same injected noise."""
    cfg = OracleConfig.for_robot('go2', sampling_method='mppi', control_parametrization='cubic_spline',
                                 num_parallel_computations=10000)
    orc, corc = SamplingOracle(cfg), c_oracle.COracle(cfg)
    sc = synthetic.scenario('go2', 'trot', seed=0)
    noise = corc.noise(42, 3)
    best = np.zeros(orc.P, np.float32)
    a = orc.compute_control(sc['state'], sc['reference'], sc['contact_sequence'], best, noise)
    b = corc.compute_control(sc['state'], sc['reference'], sc['contact_sequence'], best, noise)
    np.testing.assert_allclose(b['costs'], a['costs'], rtol=2e-6)
    assert a['best_index'] == b['best_index']
    np.testing.assert_allclose(b['best_control_parameters'], a['best_control_parameters'], rtol=1e-4, atol=1e-5)
