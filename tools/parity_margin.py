"""How far inside the north_star tolerances the CUDA tick sits: cost / knot error statistics vs the oracle."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tests'))
import numpy as np
from quadruped_pympc_b200 import synthetic
from helpers import make_pair, seeded_noise, seeded_freqs, run_both

for method, param, adaptive in [('mppi', 'cubic_spline', False), ('random_sampling', 'zero_order', False),
                                ('mppi', 'linear_spline', False), ('cem_mppi', 'cubic_spline', True)]:
    for lanes in (1, 4):
        worst_c, med_c, worst_k, worst_g, exact = 0, [], 0, 0, 0
        for seed in range(6):
            oracle, ctrl, _ = make_pair('go2', 10000, control_parametrization=param, sampling_method=method,
                                        optimize_step_freq=adaptive, lanes_per_sample=lanes)
            sc = synthetic.scenario('go2', 'trot', seed=seed)
            rng = np.random.RandomState(seed)
            best0 = (0.5 * rng.randn(oracle.P)).astype(np.float32)
            sigma = np.full(oracle.P, 3.0, np.float32) if method == 'cem_mppi' else None
            noise = seeded_noise(oracle, seed, sigma)
            freqs = seeded_freqs(oracle, seed) if adaptive else None
            ref, got = run_both(oracle, ctrl, sc, best0, noise, sigma=sigma, freqs=freqs)
            rel = np.abs(got['costs'] - ref['costs']) / np.abs(ref['costs'])
            worst_c = max(worst_c, rel.max()); med_c.append(np.median(rel)); exact += np.mean(rel == 0)
            worst_k = max(worst_k, np.abs(got['best_control_parameters'] - ref['best_control_parameters']).max())
            worst_g = max(worst_g, np.abs(got['nmpc_GRFs'] - ref['nmpc_GRFs']).max())
            assert got['best_index'] == ref['best_index']
        print(f'{method:16s} {param:13s} adaptive={int(adaptive)} lanes={lanes}: cost rel err max {worst_c:.2e} median '
              f'{np.mean(med_c):.2e} bit-equal {exact / 6:.1%} | knots abs max {worst_k:.2e} | GRF abs max {worst_g:.2e}', flush=True)
