#!/usr/bin/env python
"""bench.py -- the sampling-MPC tick on B200 (and the CPU reference arm beside it).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--robots R]

Workload (BASELINE.json configs[1]): Go2, trot, MPPI, cubic-spline parametrisation (P = 96 knots), N = 10,000 samples,
H = 12, dt = 0.02; synthetic seeded states (quadruped_pympc_b200/synthetic.py).  One "step" = one MPC tick of every
controller a rank owns (default: 1 controller per GPU = configs[1] itself; --robots R batches R independent controllers
per GPU in one launch = configs[4]'s per-GPU share).  N > 1 GPUs: robots are partitioned over ranks with no data-path
collective, weak scaling.

  value   rollout-steps/s = ranks * R * N * H / (max over ranks of the summed per-tick device times), inputs resident in
          HBM, CUDA events on the handle's stream around each tick kernel, L2 flushed (256 MiB memset) before every
          timed tick, outside the events.
  e2e     the same metric through the public API (`Sampling_MPC.compute_control` -> C ABI `qpmpc_tick`) with HOST
          buffers: pinned H2D of the tick packet, kernel, D2H of the result inside the timed region, every tick.
  multi_robot   a second measurement in the same run: 512 robots per GPU in one batched launch (ticks/s, rollout-steps/s,
          and its own roofline fraction) -- the "multi-robot ticks/sec" part of the metric.
  roofline      FP32-FMA-pipe roofline of the tick kernel (north_star: no tensor cores, not HBM-bound): algorithmic flop
          per launch / mean launch duration, against the FP32 FMA peak measured in this process.
  cpu_baseline  the C/OpenMP port of the reference tick (oracle/sampling_oracle.c) on the host cores, bounded sample.
  --impl reference   times that CPU port alone (the reference's JAX sampler cannot be installed here: no jax wheel).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

ROBOT, GAIT, METHOD, PARAM, N_SAMPLES, HORIZON = 'go2', 'trot', 'mppi', 'cubic_spline', 10000, 12
P_KNOTS = 96
# algorithmic flop per tick per controller, SURVEY.md 8(a)/(d): N * (H * 573 + P [perturb] + (2P + 1) [MPPI])
FLOP_PER_ROLLOUT_STEP = 573
FLOP_PER_TICK = N_SAMPLES * (HORIZON * FLOP_PER_ROLLOUT_STEP + P_KNOTS + 2 * P_KNOTS + 1)
# algorithmic HBM bytes per tick per controller: input packet + costs (N floats) + output packet
MULTI_ROBOTS_PER_GPU = 512


def workload_name(robots):
    return (f'{ROBOT} {GAIT}, {METHOD}, {PARAM} (P={P_KNOTS}), N={N_SAMPLES} samples, H={HORIZON}; '
            f'{robots} controller(s) per GPU')


# ----------------------------------------------------------------------------------------------------------------
# clocks sampler (nvidia-smi during the timed region)
# ----------------------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ('clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
         'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, index=0):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(self.index), f'--query-gpu={self.Q}',
                                          '--format=csv,noheader,nounits', '-lms', '100'],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), line.strip()))

    def stop(self, t0=None, t1=None):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        time.sleep(0.15)
        self.proc.terminate()
        sm, smax, reasons = [], None, set()
        for t, line in self.rows:
            if t0 is not None and not (t0 - 0.05 <= t <= t1 + 0.15):
                continue
            f = [x.strip() for x in line.split(',')]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); smax = float(f[1])
            except ValueError:
                continue
            for name, v in zip(('hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap'), f[2:6]):
                if v.lower().startswith('active'):
                    reasons.add(name)
        return {'sm_mhz': float(np.median(sm)) if sm else None, 'sm_max_mhz': smax, 'reasons': sorted(reasons),
                'samples': len(sm)}
