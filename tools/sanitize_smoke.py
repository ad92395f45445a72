"""Small workload for compute-sanitizer (memcheck / racecheck / synccheck): every kernel variant and the last-block paths
(merge, CEM two-level top-k, elite refit, stateful staging, fused exchange in loopback) at sizes that stay fast under the
tool.      compute-sanitizer --tool racecheck python tools/sanitize_smoke.py"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import quadruped_pympc_b200 as qp
from quadruped_pympc_b200 import synthetic
from quadruped_pympc_b200.batched import BatchedSampling_MPC

sc = synthetic.scenario('go2', 'trot', seed=1)
for lanes in (1, 4):
    for method, param, adaptive, N in (('mppi', 'cubic_spline', False, 700), ('random_sampling', 'zero_order', False, 300),
                                       ('cem_mppi', 'linear_spline', True, 5000), ('cem_mppi', 'cubic_spline', False, 400)):
        cfg = qp.SamplerConfig.for_robot('go2', sampling_method=method, control_parametrization=param,
                                         num_parallel_computations=N, optimize_step_freq=adaptive, lanes_per_sample=lanes)
        c = qp.Sampling_MPC(cfg)
        P = c.num_control_parameters
        best = np.zeros(P, np.float32)
        for _ in range(2):
            c.with_newkey()
            args = ((np.full(P, 3.0, np.float32),) if method == 'cem_mppi' else ()) + (sc['phase_signal'], 1.4, 1)
            out = c.jitted_compute_control(sc['state'], sc['reference'], sc['contact_sequence'], best, c.master_key, *args)
            best = out[3]
        st = dict(position=sc['state'][0:3], linear_velocity=sc['state'][3:6], orientation=sc['state'][6:9],
                  angular_velocity=sc['state'][9:12], foot_FL=sc['state'][12:15], foot_FR=sc['state'][15:18],
                  foot_RL=sc['state'][18:21], foot_RR=sc['state'][21:24])
        c.step(sc['state'], sc['reference'], sc['contact_sequence'], sc['phase_signal'], 1.4, 1, iterations=2)
        print('ok', lanes, method, param, adaptive, float(out[4]), flush=True)
    # sharded, fused exchange in loopback (peers played by the host)
    cfg = qp.SamplerConfig.for_robot('go2', sampling_method='cem_mppi', num_parallel_computations=900, lanes_per_sample=lanes)
    shards = [qp.Sampling_MPC(cfg, shard_rank=r, shard_world=3).peer_loopback() for r in range(3)]
    key = np.array([42, 1], np.uint32)
    sig = np.full(96, 3.0, np.float32)
    recs = [s.shard_partial(sc['state'], sc['reference'], sc['contact_sequence'], np.zeros(96, np.float32), key, sigma=sig) for s in shards]
    for me, s in enumerate(shards):
        for w in range(3):
            if w != me:
                s.peer_inject(w, recs[w])
        s.jitted_compute_control(sc['state'], sc['reference'], sc['contact_sequence'], np.zeros(96, np.float32), key, sig,
                                 sc['phase_signal'], 1.4, 1)
    print('ok sharded', lanes, flush=True)
b = BatchedSampling_MPC(qp.SamplerConfig.for_robot('go2', sampling_method='mppi', num_parallel_computations=300), 3)
scs = [synthetic.scenario('go2', 'trot', seed=r) for r in range(3)]
o = b.with_newkey().compute_control(np.stack([s['state'] for s in scs]), np.stack([s['reference'] for s in scs]),
                                    np.stack([s['contact_sequence'] for s in scs]))
print('ok batched', o[3], flush=True)
