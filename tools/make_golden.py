"""Generate tests/golden/*.npz by executing the REFERENCE'S OWN SOURCE (read in place from /root/reference, never copied)
under the NumPy stand-in for jax in oracle/jaxshim (jax/jaxlib are not installable here).  Run in the build container:

    python tools/make_golden.py            # writes tests/golden/ref_*.npz  (a few minutes)

What is executed, unmodified: quadruped_pympc/config.py, controllers/sampling/centroidal_nmpc_jax.py,
controllers/sampling/centroidal_nmpc_jax_gait_adaptive.py, controllers/sampling/centroidal_model_jax.py,
helpers/periodic_gait_generator_jax.py.  Stubbed (not on the hot path): gym_quadruped.robot_cfgs, mujoco.viewer.

Every fixture stores the inputs, the RandomState seeds of the raw draws the (shim) jax.random calls produced, a checksum
of those draws, and everything the reference returned.  tests/test_golden.py regenerates the raw draws from the seeds
(numpy's legacy MT19937 streams are frozen), feeds them to the oracle and compares.  The GPU box never reads
/root/reference: only the committed .npz files travel.
"""
from __future__ import annotations

import argparse
import hashlib
import importlib
import os
import sys
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REFERENCE = os.environ.get('QPMPC_REFERENCE', '/root/reference')
GOLDEN = os.path.join(ROOT, 'tests', 'golden')


def install_reference_imports():
    sys.path.insert(0, os.path.join(ROOT, 'oracle', 'jaxshim'))
    sys.path.insert(0, REFERENCE)
    sys.path.insert(0, ROOT)
    # --- stubs for modules config.py drags in but the sampling path never uses -------------------------------
    mj = types.ModuleType('mujoco')
    mjv = types.ModuleType('mujoco.viewer')
    mjv.Handle = type('Handle', (), {})
    mj.viewer = mjv
    sys.modules['mujoco'], sys.modules['mujoco.viewer'] = mj, mjv
    gq = types.ModuleType('gym_quadruped')
    rc = types.ModuleType('gym_quadruped.robot_cfgs')

    class RobotConfig:                                    # only attribute access happens in config.py:11-16
        leg_joints = None
        feet_geom_names = None
        qpos0_js = None
        hip_height = 0.3

    rc.RobotConfig = RobotConfig
    rc.get_robot_config = lambda robot_name: RobotConfig()
    gq.robot_cfgs = rc
    sys.modules['gym_quadruped'], sys.modules['gym_quadruped.robot_cfgs'] = gq, rc
    import jax                                            # the shim
    assert jax.__version__.endswith('numpy-shim'), 'a real jax is installed: use it instead of the shim'
    import quadruped_pympc.config as config
    return jax, config


def draws_digest(log):
    h = hashlib.sha256()
    for kind, _key, seed, raw in log:
        h.update(kind.encode()); h.update(np.int64(seed).tobytes()); h.update(np.ascontiguousarray(raw).tobytes())
    return h.hexdigest()


def make_controller(config, robot, method, param, adaptive, N, **mp):
    from quadruped_pympc_b200.config import ROBOT_INERTIALS
    mass, inertia = ROBOT_INERTIALS[robot]
    config.robot, config.mass, config.inertia = robot, mass, np.array(inertia)
    p = config.mpc_params
    p.update(type='sampling', sampling_method=method, control_parametrization=param, optimize_step_freq=adaptive,
             num_parallel_computations=N, device='cpu', grf_max=mass * config.gravity_constant, grf_min=0,
             use_nonuniform_discretization=False, num_splines=2, horizon=12, dt=0.02, shift_solution=False)
    p.update(mp)
    name = 'centroidal_nmpc_jax_gait_adaptive' if adaptive else 'centroidal_nmpc_jax'
    mod = importlib.import_module(f'quadruped_pympc.controllers.sampling.{name}')
    return mod.Sampling_MPC()


def run_case(jax, config, tag, robot, gait, method, param, adaptive, N, seed, best_scale=0.5, optimize_swing=1, **mp):
    from quadruped_pympc_b200 import synthetic
    ctrl = make_controller(config, robot, method, param, adaptive, N, **mp)
    H = config.mpc_params['horizon']
    sc = synthetic.scenario(robot, gait, horizon=H, mpc_dt=config.mpc_params['dt'], seed=seed)
    rng = np.random.RandomState(seed)
    P = ctrl.num_control_parameters
    best0 = (best_scale * rng.randn(P)).astype(np.float32)
    key = jax.random.PRNGKey(1000 + seed)
    sigma = None
    jax.random.DRAW_LOG.clear()
    args = [sc['state'], sc['reference'], sc['contact_sequence'], best0, key]
    if method == 'cem_mppi':
        sigma = (np.linspace(1.0, 4.0, P) if seed % 2 else np.full(P, 3.0)).astype(np.float32)
        args += [sigma]
        if adaptive:
            args += [sc['phase_signal'], sc['nominal_step_freq'], optimize_swing]
    else:
        args += [sc['phase_signal'], sc['nominal_step_freq'], optimize_swing]
    out = ctrl.jitted_compute_control(*args)
    log = list(jax.random.DRAW_LOG)
    if adaptive and method == 'cem_mppi':
        # A:1123-1131 returns 7 values WITHOUT best_freq: (GRFs, footholds, pred, best, best_cost, costs, new_sigma)
        out = tuple(out[:5]) + (np.float32(np.nan), out[5], out[6])
    rec = dict(robot=robot, gait=gait, method=method, param=param, adaptive=int(adaptive), N=N, seed=seed,
               state=sc['state'], reference=sc['reference'], contact_sequence=sc['contact_sequence'],
               phase_signal=sc['phase_signal'], nominal_step_freq=sc['nominal_step_freq'],
               optimize_swing=optimize_swing, best0=best0,
               draw_kinds=np.array([k for k, *_ in log]), draw_seeds=np.array([s for _, _, s, _ in log], np.int64),
               draw_shapes=np.array([list(r.shape) + [0] * (2 - r.ndim) for *_, r in log], np.int64),
               draws_sha256=draws_digest(log),
               nmpc_GRFs=np.asarray(out[0]), nmpc_footholds=np.asarray(out[1]), nmpc_predicted_state=np.asarray(out[2]),
               best_control_parameters=np.asarray(out[3]), best_cost=np.asarray(out[4]), best_freq=np.asarray(out[5]),
               costs=np.asarray(out[6]))
    for k, v in mp.items():
        rec['mp_' + k] = v
    if sigma is not None:
        rec['sigma'] = sigma
    if method == 'cem_mppi' and len(out) > 7:
        rec['new_sigma_cem_mppi'] = np.asarray(out[7])
    # the RNG-free seam on its own explicit parameters (row a8)
    params = (4.0 * rng.randn(min(N, 64), P)).astype(np.float32)
    if adaptive:
        fr = np.asarray(config.mpc_params['step_freq_available'], np.float32)[rng.randint(0, 3, params.shape[0])]
        rec['seam_freqs'] = fr
        rec['seam_costs'] = np.asarray(ctrl.jit_vectorized_rollout(sc['state'], sc['reference'], sc['phase_signal'], params, fr))
    else:
        rec['seam_costs'] = np.asarray(ctrl.jit_vectorized_rollout(sc['state'], sc['reference'], params, sc['contact_sequence']))
    rec['seam_params'] = params
    os.makedirs(GOLDEN, exist_ok=True)
    path = os.path.join(GOLDEN, f'ref_{tag}.npz')
    np.savez_compressed(path, **rec)
    print(f'{tag}: best_cost {float(out[4]):.4f} best_freq {float(out[5]):.3f} cost range [{rec["costs"].min():.2f}, '
          f'{rec["costs"].max():.2f}] -> {os.path.relpath(path, ROOT)}', flush=True)


def host_cases(jax, config):
    """prepare_state_and_reference (S:563-627) incl. lift-off zeroing, and shift_solution (S:513-561)."""
    recs = {}
    for param in ('zero_order', 'linear_spline', 'cubic_spline'):
        ctrl = make_controller(config, 'go2', 'mppi', param, False, 8)
        rng = np.random.RandomState(77)
        P = ctrl.num_control_parameters
        best = rng.randn(P).astype(np.float32)
        ctrl.best_control_parameters = jax.numpy.array(best)
        rng = np.random.RandomState(78)          # the same state / reference dictionaries for every parametrisation
        st = dict(position=rng.randn(3), linear_velocity=rng.randn(3), orientation=rng.randn(3) * 0.1,
                  angular_velocity=rng.randn(3), foot_FL=rng.randn(3), foot_FR=rng.randn(3), foot_RL=rng.randn(3),
                  foot_RR=rng.randn(3))
        rf = dict(ref_position=rng.randn(3), ref_linear_velocity=rng.randn(3), ref_orientation=rng.randn(3) * 0.1,
                  ref_angular_velocity=rng.randn(3), ref_foot_FL=rng.randn(1, 3), ref_foot_FR=rng.randn(1, 3),
                  ref_foot_RL=rng.randn(1, 3), ref_foot_RR=rng.randn(1, 3))
        cur, prev = np.array([1, 0, 0, 1]), np.array([1, 1, 0, 1])
        s24, r24 = ctrl.prepare_state_and_reference(st, rf, cur, prev)
        recs[f'{param}_best_in'] = best
        recs[f'{param}_state24'] = np.asarray(s24, np.float64)
        recs[f'{param}_ref24'] = np.asarray(r24, np.float64)
        recs[f'{param}_best_after_prepare'] = np.asarray(ctrl.best_control_parameters)
        recs[f'{param}_shifted'] = np.asarray(ctrl.shift_solution(jax.numpy.array(best), 1.0 / 100))
        if param == 'cubic_spline':
            for k, v in {**st, **rf}.items():
                recs['in_' + k] = np.asarray(v)
            recs['current_contact'], recs['previous_contact'] = cur, prev
    # PeriodicGaitGeneratorJax.compute_contact_sequence (G:136-151) for the three available step frequencies
    from quadruped_pympc.helpers.periodic_gait_generator_jax import PeriodicGaitGeneratorJax
    pgg = PeriodicGaitGeneratorJax(duty_factor=0.65, step_freq=1.65, horizon=12, mpc_dt=0.02)
    t0 = np.array([0.5, 1.0, 1.0, 0.5], np.float32)
    for i, f in enumerate((1.4, 2.0, 2.4)):
        cs, t_new = pgg.compute_contact_sequence(simulation_dt=0.002, t=t0, step_freq=f)
        recs[f'pgg_contact_{i}'], recs[f'pgg_t_{i}'] = np.asarray(cs), np.asarray(t_new)
    recs['pgg_t0'] = t0
    path = os.path.join(GOLDEN, 'ref_host.npz')
    np.savez_compressed(path, **recs)
    print('host cases ->', os.path.relpath(path, ROOT), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--samples', type=int, default=160)
    ap.add_argument('--only', default='')
    ap.add_argument('--skip-existing', action='store_true')
    a = ap.parse_args()
    jax, config = install_reference_imports()
    N = a.samples
    cases = []
    seed = 0
    for adaptive in (False, True):
        for method in ('random_sampling', 'mppi', 'cem_mppi'):
            for param in ('zero_order', 'linear_spline', 'cubic_spline'):
                seed += 1
                robot = ('go2', 'aliengo', 'hyqreal1')[seed % 3]
                gait = ('trot', 'pace')[seed % 2]
                tag = f'{"adaptive" if adaptive else "fixed"}_{method}_{param}'
                n = max(N, 260) if (adaptive and method == 'cem_mppi') else N      # adaptive CEM needs > 200 elites
                cases.append(dict(tag=tag, robot=robot, gait=gait, method=method, param=param, adaptive=adaptive,
                                  N=n, seed=seed))
    # extra shapes: no leg in stance at some step (mass*g/0 -> NaN path), nonuniform dt, 3 splines, optimize_swing=0
    cases.append(dict(tag='fixed_mppi_cubic_nonuniform_dt', robot='go2', gait='trot', method='mppi', param='cubic_spline',
                      adaptive=False, N=N, seed=31, use_nonuniform_discretization=True))
    cases.append(dict(tag='fixed_mppi_linear_3splines', robot='go2', gait='pace', method='mppi', param='linear_spline',
                      adaptive=False, N=N, seed=32, num_splines=3))
    cases.append(dict(tag='adaptive_random_sampling_cubic_noswing', robot='aliengo', gait='trot', method='random_sampling',
                      param='cubic_spline', adaptive=True, N=N, seed=33, optimize_swing=0))
    # row a15 with optimize_swing = 0: mppi draws from step_freq_available regardless (A:835-838), cem_mppi collapses to the
    # nominal frequency (A:994-1000)
    cases.append(dict(tag='adaptive_mppi_linear_noswing', robot='go2', gait='pace', method='mppi', param='linear_spline',
                      adaptive=True, N=N, seed=35, optimize_swing=0))
    cases.append(dict(tag='adaptive_cem_mppi_cubic_noswing', robot='hyqreal1', gait='trot', method='cem_mppi', param='cubic_spline',
                      adaptive=True, N=max(N, 260), seed=36, optimize_swing=0))
    cases.append(dict(tag='fixed_random_sampling_cubic_bigbest', robot='hyqreal1', gait='trot', method='random_sampling',
                      param='cubic_spline', adaptive=False, N=N, seed=34, best_scale=40.0))
    for c in cases:
        if a.only and a.only not in c['tag']:
            continue
        if a.skip_existing and os.path.exists(os.path.join(GOLDEN, f"ref_{c['tag']}.npz")):
            continue
        run_case(jax, config, **c)
    if not a.only or a.only == 'host':
        host_cases(jax, config)


if __name__ == '__main__':
    main()
