"""Soak of the direct-I/O tick through the public API: N ticks back to back, latency percentiles and a consistency check
(the carried best must stay finite; every result must belong to its own tick: best_cost changes with the key).

    python tools/soak_e2e.py [ticks=200000]
"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import quadruped_pympc_b200 as qp
from quadruped_pympc_b200 import synthetic


def main():
    ticks = int(sys.argv[1]) if len(sys.argv) > 1 else 200000
    qp.build()
    cfg = qp.SamplerConfig.for_robot('go2', sampling_method='mppi', control_parametrization='cubic_spline',
                                     num_parallel_computations=10000)
    c = qp.Sampling_MPC(cfg)
    scs = [synthetic.scenario('go2', 'trot', seed=s, sim_steps=5 * s) for s in range(64)]
    best = np.zeros(c.num_control_parameters, np.float32)
    lat = np.zeros(ticks)
    costs = np.zeros(ticks, np.float32)
    for i in range(ticks):
        sc = scs[i & 63]
        c.with_newkey()
        t0 = time.perf_counter()
        out = c.compute_control(sc['state'], sc['reference'], sc['contact_sequence'], best, c.master_key,
                                sc['phase_signal'], sc['nominal_step_freq'], 0)
        lat[i] = time.perf_counter() - t0
        best = out[3]
        costs[i] = out[4]
    lat = lat[100:] * 1e6                 # the first call creates the handle (CUDA context, allocations)
    assert np.all(np.isfinite(best)) and np.all(np.isfinite(costs))
    # the same 64 scenarios recur with a drifting best: consecutive results must differ (no stale packet delivered twice)
    assert np.count_nonzero(np.diff(costs) == 0.0) == 0
    print(f'{ticks} ticks, first 100 dropped ({c.kernel_name}): p50 {np.percentile(lat, 50):.1f} us  p99 {np.percentile(lat, 99):.1f}  '
          f'p99.9 {np.percentile(lat, 99.9):.1f}  max {lat.max():.1f}  mean {lat.mean():.1f}  total {lat.sum() * 1e-6:.1f} s')


if __name__ == '__main__':
    main()
