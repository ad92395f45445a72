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

HORIZON = 12
# BASELINE.json configs[0..2] as concrete synthetic workloads; flop_step = algorithmic FP32 flop per
# rollout-step (378 zero-order, 417 linear, 573 cubic, 640 adaptive cubic)
WORKLOADS = {
    'config0': dict(robot='aliengo', gait='trot', method='random_sampling', param='zero_order', N=10000, adaptive=False,
                    P=144, flop_step=378, baseline='configs[0]: Aliengo trot, random_sampling, zero-order, 10k samples'),
    'config1': dict(robot='go2', gait='trot', method='mppi', param='cubic_spline', N=10000, adaptive=False,
                    P=96, flop_step=573, baseline='configs[1]: Go2 trot, MPPI, cubic spline, 10k samples'),
    'config2': dict(robot='hyqreal1', gait='pace', method='cem_mppi', param='cubic_spline', N=50000, adaptive=True,
                    P=96, flop_step=640, baseline='configs[2]: HyQReal pace, CEM-MPPI + step-frequency adaptation, 50k samples'),
}
HEAD = WORKLOADS['config1']          # the configuration BASELINE.json's metric is quoted on
ROBOT, GAIT, METHOD, PARAM, N_SAMPLES = HEAD['robot'], HEAD['gait'], HEAD['method'], HEAD['param'], HEAD['N']
P_KNOTS = HEAD['P']
MULTI_ROBOTS_PER_GPU = 512


def flop_per_tick(w, n=None):
    """algorithmic flop per tick per controller, SURVEY.md 8(a)/(d): N * (H * F_step + P [perturb] + (2P + 1) [MPPI / CEM])"""
    n = w['N'] if n is None else n
    return n * (HORIZON * w['flop_step'] + w['P'] + (2 * w['P'] + 1 if w['method'] != 'random_sampling' else 0))


def packet_floats(w):
    """(tick packet, result packet) floats of one controller: include/qpmpc.h / csrc/qpmpc_kernels.cuh layout"""
    elite = (200 if w['adaptive'] else 10) if w['method'] == 'cem_mppi' else 0
    r4 = lambda x: (x + 3) // 4 * 4
    return r4(60 + 4 * HORIZON + 2 * w['P']), r4(40 + 2 * w['P'] + elite)


FLOP_PER_TICK = flop_per_tick(HEAD)


def bench_config(robots, world):
    return {'workload': workload_name(robots),
            'parallelism': f'robots partitioned over {world} GPU(s), no collective',
            'l2': 'flushed (256 MiB memset) before every timed tick, outside the timed span',
            'timing': 'value = host wall clock per tick (launch call -> cudaStreamSynchronize returns, L2 flush done before '
                      'the clock starts), ranks released by a barrier, max over ranks; device = a separate pass with CUDA '
                      'events around the kernel',
            'noise': 'in-kernel Philox4x32-10 (GPU arm) / host Philox mirror (CPU arm)'}


def workload_name(robots, w=None):
    w = HEAD if w is None else w
    return (f"{w['robot']} {w['gait']}, {w['method']}{' + step-frequency adaptation' if w['adaptive'] else ''}, "
            f"{w['param']} (P={w['P']}), N={w['N']} samples, H={HORIZON}; {robots} controller(s) per GPU")


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


# ----------------------------------------------------------------------------------------------------------------
# CPU arm: the C/OpenMP port of the reference tick (oracle/sampling_oracle.c), noise generation included
# ----------------------------------------------------------------------------------------------------------------
def host_threads():
    """threads the CPU arm may use: every core this process is allowed on (torchrun exports OMP_NUM_THREADS=1 to its
    workers -- that is a default for N ranks sharing a box, not a limit for the one rank that runs the CPU arm)"""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_tick_factory(w=None, threads=None):
    """Returns (tick(counter, controller=0) -> best_cost, threads).  One call = one full reference tick on the host: draw
    the (N,P) noise [and the per-sample step frequencies], perturb, N rollouts with cost, the control-law update, GRF
    read-out (what jitted_compute_control does)."""
    from oracle import c_oracle
    from oracle.sampling_oracle import OracleConfig, SamplingOracle
    from quadruped_pympc_b200 import synthetic
    w = HEAD if w is None else w
    c_oracle.build()
    c_oracle.set_num_threads(host_threads() if threads is None else threads)
    ocfg = OracleConfig.for_robot(w['robot'], sampling_method=w['method'], control_parametrization=w['param'],
                                  num_parallel_computations=w['N'], optimize_step_freq=w['adaptive'])
    orc = c_oracle.COracle(ocfg)
    np_orc = SamplingOracle(ocfg)
    scs, bests = {}, {}
    sigma = np.full(orc.P, 3.0, np.float32) if w['method'] == 'cem_mppi' else None

    def tick(counter, controller=0):
        if controller not in scs:
            scs[controller] = synthetic.scenario(w['robot'], w['gait'], seed=controller)
            bests[controller] = np.zeros(orc.P, np.float32)
        sc = scs[controller]
        noise = orc.noise(42, counter, sigma=sigma, robot=controller)
        freqs = None
        if w['adaptive']:
            freqs = np_orc.assemble_step_frequencies(orc.freq_choice(42, counter, robot=controller), sc['nominal_step_freq'], 1)
        out = orc.compute_control(sc['state'], sc['reference'], sc['contact_sequence'], bests[controller], noise,
                                  timing=sc['phase_signal'], freqs=freqs)
        bests[controller] = out['best_control_parameters']
        return float(out['best_cost'])

    return tick, c_oracle.num_threads()


def time_cpu(seconds_budget, w=None, min_ticks=3, max_ticks=100000, warmup=2, single_core=True):
    w = HEAD if w is None else w
    tick, threads = cpu_tick_factory(w)
    for i in range(warmup):
        tick(i)
    t0 = time.perf_counter()
    n = 0
    while n < max_ticks and (n < min_ticks or time.perf_counter() - t0 < seconds_budget):
        tick(warmup + n)
        n += 1
    dt = time.perf_counter() - t0
    res = {'value': n * w['N'] * HORIZON / dt, 'unit': 'rollout-steps/s', 'cores': threads, 'kind': 'port',
           'sample': f"{n} full ticks of the same workload (noise draw + {w['N']} rollouts x H={HORIZON} + {w['method']} update) "
                     f'in {dt:.2f} s, C/OpenMP port of the reference tick (oracle/sampling_oracle.c), {threads} threads; the '
                     f"reference's own JAX sampler is not installable on this box (no jax wheel, no network)",
           'ms_per_tick': 1e3 * dt / n}
    if single_core:
        # the same port on ONE core (SURVEY 8d ii): a few ticks are enough, a tick takes ~threads x longer
        from oracle import c_oracle
        c_oracle.set_num_threads(1)
        t1 = time.perf_counter()
        n1 = 0
        while n1 < 3 or (time.perf_counter() - t1 < 0.15 * seconds_budget and n1 < 50):
            tick(warmup + n + n1)
            n1 += 1
        res['single_core_value'] = n1 * w['N'] * HORIZON / (time.perf_counter() - t1)
        c_oracle.set_num_threads(threads)
    return res


def run_reference_arm(args):
    """--impl reference: the reference's CPU implementation of the path.  The reference's own JAX sampler cannot be
    installed here (no jax/jaxlib wheel, no network), so this times the C/OpenMP port of its tick on all host cores.
    The workload is the GPU arm's: `--gpus N` means N controllers (one per GPU there), so one step here ticks N
    controllers on the box's host cores; rank 0 alone runs it (the other ranks exit)."""
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    world = max(1, int(args.gpus))
    tick, threads = cpu_tick_factory(HEAD)
    for i in range(args.warmup):
        for c in range(world):
            tick(i, c)
    t0 = time.perf_counter()
    for i in range(args.steps):
        for c in range(world):
            tick(args.warmup + i, c)
    dt = time.perf_counter() - t0
    value = world * args.steps * N_SAMPLES * HORIZON / dt
    sample = (f'{args.steps} steps x {world} controller(s) = {args.steps * world} full ticks, C/OpenMP port of the reference '
              f'tick (jax is not installable here), {threads} threads')
    line = {'metric': 'rollout-steps/s', 'value': value, 'unit': 'rollout-steps/s', 'impl': 'reference',
            'n_gpus': args.gpus, 'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': 1e3 * dt / args.steps,
            'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
            'config': bench_config(1, args.gpus),
            'cpu_baseline': {'value': value, 'unit': 'rollout-steps/s', 'cores': threads, 'kind': 'port', 'sample': sample},
            'e2e': {'value': value, 'unit': 'rollout-steps/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
            'tick_latency_us': {'p50': 1e6 * dt / (args.steps * world)}}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------------------------
# GPU arm
# ----------------------------------------------------------------------------------------------------------------
def _sampler_cfg(device, w=None, n_samples=None):
    import quadruped_pympc_b200 as qp
    w = HEAD if w is None else w
    return qp.SamplerConfig.for_robot(w['robot'], sampling_method=w['method'], control_parametrization=w['param'],
                                      num_parallel_computations=int(w['N'] if n_samples is None else n_samples),
                                      optimize_step_freq=w['adaptive'], device=device)


def make_resident_handle(device, robots, robot_base, n_samples=None, lanes=None, w=None):
    """A handle with `robots` controllers whose tick packets are uploaded once and stay in HBM."""
    from quadruped_pympc_b200 import _lib, synthetic
    from quadruped_pympc_b200.sampling_mpc import make_c_config
    lib = _lib.load()
    w = HEAD if w is None else w
    scfg = _sampler_cfg(device, w, n_samples)
    if lanes is not None:
        scfg.lanes_per_sample = int(lanes)
    c = make_c_config(scfg, robots)
    c.robot_base = robot_base
    h = C.c_void_p()
    _lib.check(lib.qpmpc_create(C.byref(c), C.byref(h)))
    ins = (_lib.QpmpcTickIn * robots)()
    keep = []
    fp = lambda a: a.ctypes.data_as(_lib._FP)
    for r in range(robots):
        sc = synthetic.scenario(w['robot'], w['gait'], seed=robot_base + r)
        st, rf = sc['state'].astype(np.float32), sc['reference'].astype(np.float32)
        cs = np.ascontiguousarray(sc['contact_sequence'], np.float32)
        best = np.zeros(w['P'], np.float32)
        sigma = np.full(w['P'], 3.0, np.float32)
        ph = sc['phase_signal'].astype(np.float32)
        keep += [st, rf, cs, best, ph, sigma]
        i = ins[r]
        i.state, i.reference, i.contact_sequence, i.contact_ld, i.best = fp(st), fp(rf), fp(cs), cs.shape[1], fp(best)
        if w['method'] == 'cem_mppi':
            i.sigma = fp(sigma)
        i.phase_signal, i.nominal_step_freq, i.optimize_swing = fp(ph), float(sc['nominal_step_freq']), int(w['adaptive'])
        i.key[0], i.key[1] = 42, 1
    _lib.check(lib.qpmpc_upload_resident(h, ins))
    return h, keep


def run_resident(h, steps, warmup, flush=1, wall=False):
    """-> per-tick device times (ms) of `steps` resident ticks after `warmup` untimed ones, the launch count [and, when
    wall=True, the per-tick host wall times (ms) of a SECOND pass of `steps` ticks timed without events]."""
    from quadruped_pympc_b200 import _lib
    lib = _lib.load()
    if warmup > 0:
        wm = np.zeros(warmup, np.float32)
        _lib.check(lib.qpmpc_bench_resident(h, warmup, flush, wm.ctypes.data_as(_lib._FP)))
    n0, n1 = C.c_uint64(0), C.c_uint64(0)
    _lib.check(lib.qpmpc_kernel_launch_count(h, C.byref(n0)))
    ms, wl = np.zeros(steps, np.float32), np.zeros(steps, np.float32)
    _lib.check(lib.qpmpc_bench_resident(h, steps, flush, ms.ctypes.data_as(_lib._FP)))            # device: CUDA events
    if wall:                                                                                          # host wall clock, no events
        scratch = np.zeros(steps, np.float32)
        _lib.check(lib.qpmpc_bench_resident_wall(h, steps, flush, scratch.ctypes.data_as(_lib._FP), wl.ctypes.data_as(_lib._FP)))
    _lib.check(lib.qpmpc_kernel_launch_count(h, C.byref(n1)))
    if wall:
        return ms.astype(np.float64), int(n1.value - n0.value), wl.astype(np.float64)
    return ms.astype(np.float64), int(n1.value - n0.value)


def run_e2e(device, steps, warmup, w=None):
    """The user-facing call with host buffers: Sampling_MPC.compute_control -> qpmpc_tick (H2D + kernel + D2H per tick).
    -> (per-tick wall times in s, h2d bytes, d2h bytes, launches, kernel name)"""
    import quadruped_pympc_b200 as qp
    from quadruped_pympc_b200 import _lib, synthetic
    w = HEAD if w is None else w
    ctrl = qp.Sampling_MPC(_sampler_cfg(device, w))
    sc = synthetic.scenario(w['robot'], w['gait'], seed=0)
    best = np.zeros(w['P'], np.float32)
    sigma = np.full(w['P'], 3.0, np.float32)
    lat = np.zeros(steps)
    h = None
    n0 = C.c_uint64(0)
    opt = int(w['adaptive'])
    for i in range(warmup + steps):
        if i == warmup:
            h = ctrl._ensure()
            _lib.check(_lib.load().qpmpc_kernel_launch_count(h, C.byref(n0)))
        ctrl.with_newkey()
        t0 = time.perf_counter()
        if w['method'] == 'cem_mppi':
            out = ctrl.compute_control(sc['state'], sc['reference'], sc['contact_sequence'], best, ctrl.master_key, sigma,
                                       sc['phase_signal'], sc['nominal_step_freq'], opt)
        else:
            out = ctrl.compute_control(sc['state'], sc['reference'], sc['contact_sequence'], best, ctrl.master_key,
                                       sc['phase_signal'], sc['nominal_step_freq'], opt)
        t1 = time.perf_counter()
        best = out[3]
        if i >= warmup:
            lat[i - warmup] = t1 - t0
    n1 = C.c_uint64(0)
    _lib.check(_lib.load().qpmpc_kernel_launch_count(h, C.byref(n1)))
    in_floats, out_floats = packet_floats(w)
    assert np.isfinite(float(out[4]))
    name = ctrl.kernel_name
    ctrl.close()
    return lat, 4 * in_floats, 4 * out_floats, int(n1.value - n0.value), name


def run_config(device, w, steps, warmup, fp32_peak, with_cpu, cpu_seconds):
    """One BASELINE config measured like the headline: resident device ticks (events), e2e through the public API with host
    buffers, FP32 roofline fraction from the algorithmic flop count, optionally its own CPU baseline."""
    from quadruped_pympc_b200 import _lib
    lib = _lib.load()
    h, keep = make_resident_handle(device, 1, 0, w=w)
    kernel = lib.qpmpc_kernel_name(h).decode()
    ms, launches, wl = run_resident(h, steps, warmup, wall=True)
    lib.qpmpc_destroy(h)
    lat, h2d, d2h, e2e_launches, _ = run_e2e(device, steps, warmup, w)
    units = w['N'] * HORIZON
    tf = flop_per_tick(w) / (float(ms.mean()) * 1e-3) / 1e12
    out = {'baseline_config': w['baseline'], 'workload': workload_name(1, w), 'kernel': f'qp::{kernel}', 'steps': steps,
           'value': units * steps / (wl.sum() * 1e-3), 'unit': 'rollout-steps/s',
           'device_value': units * steps / (ms.sum() * 1e-3),
           'tick_latency_us': {'device_p50': float(np.percentile(ms, 50) * 1e3), 'device_p99': float(np.percentile(ms, 99) * 1e3),
                               'e2e_p50': float(np.percentile(lat, 50) * 1e6), 'e2e_p99': float(np.percentile(lat, 99) * 1e6)},
           'e2e': {'value': units * steps / lat.sum(), 'unit': 'rollout-steps/s', 'h2d_bytes_per_step': h2d,
                   'd2h_bytes_per_step': d2h},
           'roofline': {'bound': 'fp32', 'achieved': tf, 'peak': fp32_peak, 'unit': 'TFLOP/s', 'frac': tf / fp32_peak,
                        'algorithmic_flop_per_launch': flop_per_tick(w), 'flop_per_rollout_step': w['flop_step']},
           'gpu_launches': launches + e2e_launches}
    if with_cpu:
        out['cpu_baseline'] = time_cpu(cpu_seconds, w, single_core=False)
    return out


def run_sharded(device, world, n_samples, ticks, barrier, max_over_ranks):
    """BASELINE configs[3]: ONE controller of `n_samples` samples split over `world` GPUs.  Every rank rolls out its slice
    and the shard records are exchanged inside the tick kernel (stores into the peers' mailboxes over NVLink, one launch
    per rank per tick); all ranks end with the same control.  Timed through the public API with host buffers.  Before the
    timed ticks the sharded result is compared ONCE with an unsharded controller of the same size on the same key
    (`parity_ok`: same argmin, best cost rel 1e-5, knots / GRFs within 1e-4 + the softmax-merge rounding)."""
    import quadruped_pympc_b200 as qp
    from quadruped_pympc_b200 import synthetic
    from quadruped_pympc_b200.distributed import make_sharded_controller
    cfg = _sampler_cfg(device, HEAD, n_samples)
    ctrl = make_sharded_controller(cfg, exchange='peer')
    sc = synthetic.scenario(ROBOT, GAIT, seed=0)
    best = np.zeros(P_KNOTS, np.float32)
    args = (sc['phase_signal'], sc['nominal_step_freq'], 0)
    # ---- parity: sharded vs unsharded, same key ----
    ctrl.with_newkey()
    key = ctrl.master_key.copy()
    out_s = ctrl.compute_control(sc['state'], sc['reference'], sc['contact_sequence'], best, key, *args)
    idx_s = ctrl.last_best_index
    whole = qp.Sampling_MPC(_sampler_cfg(device, HEAD, n_samples))
    out_u = whole.compute_control(sc['state'], sc['reference'], sc['contact_sequence'], best, key, *args)
    ok = (idx_s == whole.last_best_index and abs(float(out_s[4]) - float(out_u[4])) <= 1e-5 * abs(float(out_u[4]))
          and np.allclose(out_s[3], out_u[3], rtol=1e-4, atol=1e-4) and np.allclose(out_s[0], out_u[0], rtol=1e-4, atol=1e-3))
    whole.close()
    parity_ok = max_over_ranks(0.0 if ok else 1.0) == 0.0
    lat = np.zeros(ticks)
    for i in range(10 + ticks):
        if i == 10:
            barrier()
        ctrl.with_newkey()
        t0 = time.perf_counter()
        out = ctrl.compute_control(sc['state'], sc['reference'], sc['contact_sequence'], best, ctrl.master_key, *args)
        t1 = time.perf_counter()
        best = out[3]
        if i >= 10:
            lat[i - 10] = t1 - t0
    barrier()
    total = max_over_ranks(lat.sum())
    kernel = ctrl.kernel_name
    ctrl.close()
    return {'samples': n_samples, 'shards': world, 'kernel': f'qp::{kernel}',
            'exchange': 'fused in the tick kernel (NVLink peer stores), 1 launch/tick/rank',
            'ticks': ticks, 'e2e_p50_us': float(np.percentile(lat, 50) * 1e6), 'e2e_p99_us': float(np.percentile(lat, 99) * 1e6),
            'rollout_steps_per_s': ticks * n_samples * HORIZON / total, 'record_bytes_per_rank': 4 * (4 + P_KNOTS),
            'parity_ok': bool(parity_ok), 'parity_check': 'first tick vs an unsharded controller of the same size on the same key, every rank'}


def run_multi_e2e(device, robots, robot_base, ticks, barrier, max_over_ranks, world):
    """The batched dispatch through its public API with HOST arrays (BatchedSampling_MPC.compute_control ->
    qpmpc_tick_batched_flat): per tick R packets are packed, copied H2D, one launch, R result packets D2H."""
    import quadruped_pympc_b200 as qp
    from quadruped_pympc_b200 import synthetic
    from quadruped_pympc_b200.batched import BatchedSampling_MPC
    ctrl = BatchedSampling_MPC(_sampler_cfg(device), robots, robot_base=robot_base)
    scs = [synthetic.scenario(ROBOT, GAIT, seed=robot_base + r) for r in range(robots)]
    st = np.stack([s['state'] for s in scs]).astype(np.float32)
    rf = np.stack([s['reference'] for s in scs]).astype(np.float32)
    cs = np.stack([s['contact_sequence'] for s in scs]).astype(np.float32)
    best = np.zeros((robots, P_KNOTS), np.float32)
    lat = np.zeros(ticks)
    for i in range(3 + ticks):
        if i == 3:
            barrier()
        ctrl.with_newkey()
        t0 = time.perf_counter()
        out = ctrl.compute_control(st, rf, cs, best)
        t1 = time.perf_counter()
        best = out[2]
        if i >= 3:
            lat[i - 3] = t1 - t0
    barrier()
    total = max_over_ranks(lat.sum())
    ctrl.close()
    in_floats, out_floats = packet_floats(HEAD)
    return {'e2e_ms_per_batched_tick': float(np.median(lat) * 1e3),
            'e2e_robot_ticks_per_s': world * robots * ticks / total,
            'e2e_rollout_steps_per_s': world * robots * N_SAMPLES * HORIZON * ticks / total,
            'e2e_h2d_bytes_per_tick': 4 * in_floats * robots, 'e2e_d2h_bytes_per_tick': 4 * out_floats * robots,
            'e2e_api': 'BatchedSampling_MPC.compute_control -> qpmpc_tick_batched_flat (host arrays)'}


def measure_fp32_peak(device):
    from quadruped_pympc_b200 import _lib
    tf = C.c_float(0)
    _lib.check(_lib.load().qpmpc_measure_fp32_peak(device, C.byref(tf)))
    return float(tf.value)


def ncu_traffic_bytes():
    """dram bytes per launch of the tick kernel from the committed `ncu --set full` capture (profiles/), or None."""
    p = os.path.join(ROOT, 'profiles', 'tick_kernel_ncu_summary.json')
    try:
        with open(p) as f:
            return json.load(f).get('dram_bytes_per_launch')
    except (OSError, ValueError):
        return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=1000)   # p50 / p99 over >= 1000 ticks
    ap.add_argument('--warmup', type=int, default=50)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--robots', type=int, default=1, help='controllers per GPU in the headline measurement')
    ap.add_argument('--multi-robots', type=int, default=MULTI_ROBOTS_PER_GPU)
    ap.add_argument('--sharded-samples', type=int, default=1000000,
                    help='N > 1 GPUs: samples of the one large controller sharded over the ranks (0 = skip)')
    ap.add_argument('--no-sweep', dest='sweep', action='store_false', help='skip the 1k..1M single-controller sample sweeps')
    ap.add_argument('--no-configs', dest='configs', action='store_false', help='skip BASELINE configs[0] and configs[2]')
    ap.add_argument('--cpu-seconds', type=float, default=10.0)
    ap.add_argument('--no-cpu-baseline', action='store_true')
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == 'reference':
        return run_reference_arm(args)

    # stdout carries exactly ONE line (the JSON): everything libraries print meanwhile (NCCL's version banner, ...) goes
    # to stderr
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)

    import torch
    import torch.distributed as dist
    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    if not torch.cuda.is_available():
        raise SystemExit('bench.py needs a CUDA device: the sampling path has no CPU fallback '
                         '(use --impl reference for the CPU arm)')
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return float(x)
        t = torch.tensor([x], dtype=torch.float64, device='cuda')
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    import quadruped_pympc_b200 as qp
    from quadruped_pympc_b200 import _lib
    qp.build()
    lib = _lib.load()
    R = args.robots
    fp32_peak = measure_fp32_peak(local)

    # ---- headline: R controller(s) per GPU, resident in HBM ------------------------------------------------
    h, keep = make_resident_handle(local, R, rank * R)
    sampler = ClockSampler(local)
    sampler.start()
    run_resident(h, 1, args.warmup)                  # warm-up ticks (+1), untimed
    barrier()
    t0 = time.perf_counter()
    ms, launches, wall = run_resident(h, args.steps, 0, wall=True)
    barrier()
    t1 = time.perf_counter()
    clocks = sampler.stop(t0, t1)
    units = world * R * N_SAMPLES * HORIZON * args.steps
    # value: host wall clock per tick (launch call -> completion seen by the host), all ranks released by the barrier
    # above, max over ranks -- N ranks on one box share PCIe, the host and the driver, and this clock sees it.
    # device_*: CUDA events around the kernel alone (what the roofline is computed from).
    total_wall_ms = max_over_ranks(wall.sum())
    total_ms = max_over_ranks(ms.sum())
    value = units / (total_wall_ms * 1e-3)
    device_value = units / (total_ms * 1e-3)
    kernel_ms = float(ms.mean())
    achieved_tf = R * FLOP_PER_TICK / (kernel_ms * 1e-3) / 1e12
    in_floats, out_floats = packet_floats(HEAD)
    hbm_bytes = R * (4 * in_floats + 4 * N_SAMPLES + 4 * out_floats)
    kernel_name = lib.qpmpc_kernel_name(h).decode()      # the variant qpmpc_create picked for this launch shape
    lib.qpmpc_destroy(h)

    # ---- e2e through the public API with host buffers (rank-local controller; aggregate = world x) -----------
    barrier()
    lat, h2d, d2h, e2e_launches, _ = run_e2e(local, args.steps, args.warmup)
    barrier()
    e2e_total = max_over_ranks(lat.sum())
    e2e_value = world * N_SAMPLES * HORIZON * args.steps / e2e_total

    # ---- multi-robot: `multi_robots` controllers per GPU in one batched launch ---------------------------------
    MR = args.multi_robots
    multi = None
    if MR > 0:
        hm, keepm = make_resident_handle(local, MR, rank * MR)
        m_kernel = lib.qpmpc_kernel_name(hm).decode()
        msteps = max(3, min(args.steps, 20))
        run_resident(hm, 1, 3)
        barrier()
        mms, _, mwall = run_resident(hm, msteps, 0, wall=True)
        barrier()
        m_total = max_over_ranks(mwall.sum())
        m_tf = MR * FLOP_PER_TICK / (float(mms.mean()) * 1e-3) / 1e12
        multi = {'robots_per_gpu': MR, 'robots': MR * world, 'steps': msteps, 'kernel': f'qp::{m_kernel}<CUBIC,MPPI>',
                 'ms_per_batched_tick': float(mwall.mean()), 'device_ms_per_batched_tick': float(mms.mean()),
                 'robot_ticks_per_s': world * MR * msteps / (m_total * 1e-3),
                 'rollout_steps_per_s': world * MR * N_SAMPLES * HORIZON * msteps / (m_total * 1e-3),
                 'roofline': {'bound': 'fp32', 'achieved': m_tf, 'peak': fp32_peak, 'unit': 'TFLOP/s',
                              'frac': m_tf / fp32_peak, 'kernel_ms': float(mms.mean()),
                              'algorithmic_flop_per_launch': MR * FLOP_PER_TICK}}
        lib.qpmpc_destroy(hm)
        multi.update(run_multi_e2e(local, MR, rank * MR, msteps, barrier, max_over_ranks, world))

    # ---- single-controller sample sweep on one GPU (BASELINE configs[3] at 1 GPU): kernel time, inputs resident ----
    sweep = None
    if world == 1 and args.sweep:
        sweep = []
        for n in (1000, 10000, 100000, 1000000):
            hs, keeps = make_resident_handle(local, 1, 0, n_samples=n)
            s_kernel = lib.qpmpc_kernel_name(hs).decode()
            run_resident(hs, 1, 3)
            sms, _ = run_resident(hs, 20, 0)
            lib.qpmpc_destroy(hs)
            t = float(np.median(sms)) * 1e-3
            sweep.append({'samples': n, 'kernel': f'qp::{s_kernel}', 'kernel_us': t * 1e6, 'rollout_steps_per_s': n * HORIZON / t,
                          'fp32_frac': flop_per_tick(HEAD, n) / t / 1e12 / fp32_peak})

    # ---- one large controller, samples sharded over the ranks (BASELINE configs[3]); fused peer exchange --------
    sharded, sharded_sweep = None, None
    if world > 1 and args.sharded_samples > 0:
        sticks = min(args.steps, 100)
        sharded = run_sharded(local, world, args.sharded_samples, sticks, barrier, max_over_ranks)
        if args.sweep:
            sharded_sweep = [run_sharded(local, world, n, sticks, barrier, max_over_ranks)
                             for n in (1000, 10000, 100000) if n != args.sharded_samples] + [sharded]

    # ---- BASELINE configs[0] and configs[2], measured like the headline (rank 0's GPU) ---------------------------
    configs = None
    if rank == 0 and args.configs:
        csteps = max(20, min(args.steps, 200))
        with_cpu = (world == 1 and not args.no_cpu_baseline)
        configs = [run_config(local, WORKLOADS['config0'], csteps, min(args.warmup, 20), fp32_peak, with_cpu,
                              0.4 * args.cpu_seconds),
                   run_config(local, WORKLOADS['config2'], csteps, min(args.warmup, 20), fp32_peak, False, 0.0)]
    barrier()

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = time_cpu(args.cpu_seconds)

    if rank == 0:
        line = {
            'metric': 'rollout-steps/s', 'value': value, 'unit': 'rollout-steps/s', 'n_gpus': world,
            'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': total_wall_ms / args.steps,
            'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
            'config': bench_config(R, world),
            'device': {'value': device_value, 'unit': 'rollout-steps/s', 'ms_per_step': total_ms / args.steps,
                       'timing': 'CUDA events around the tick kernel only, summed per rank, max over ranks'},
            'tick_latency_us': {'device_p50': float(np.percentile(ms, 50) * 1e3), 'device_p99': float(np.percentile(ms, 99) * 1e3),
                                'wall_p50': float(np.percentile(wall, 50) * 1e3), 'wall_p99': float(np.percentile(wall, 99) * 1e3),
                                'e2e_p50': float(np.percentile(lat, 50) * 1e6), 'e2e_p99': float(np.percentile(lat, 99) * 1e6)},
            'e2e': {'value': e2e_value, 'unit': 'rollout-steps/s', 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': d2h,
                    'api': 'Sampling_MPC.compute_control -> qpmpc_tick (pinned H2D + kernel + D2H per tick)'},
            'gpu_launches': launches + e2e_launches,
            'roofline': {'bound': 'fp32', 'achieved': achieved_tf, 'peak': fp32_peak, 'unit': 'TFLOP/s',
                         'frac': achieved_tf / fp32_peak, 'traffic': ncu_traffic_bytes(),
                         'kernel': f'qp::{kernel_name}<CUBIC,false,MPPI>', 'kernel_ms': kernel_ms,
                         'algorithmic_flop_per_launch': R * FLOP_PER_TICK,
                         'peak_source': 'FP32 FMA microbenchmark measured in this process (MEASURED_PEAKS.json has no '
                                        'FP32 CUDA-core figure); latency-bound at 10k samples, see multi_robot.roofline',
                         'hbm': {'algorithmic_bytes_per_launch': hbm_bytes,
                                 'achieved_gbs': hbm_bytes / (kernel_ms * 1e-3) / 1e9}},
            'clocks': clocks,
        }
        if multi:
            line['multi_robot'] = multi
        if sharded:
            line['sharded_controller'] = sharded
        if sharded_sweep:
            line['sharded_sweep'] = sharded_sweep
        if sweep:
            line['sample_sweep'] = sweep
        if configs:
            line['configs'] = configs
        if cpu:
            line['cpu_baseline'] = cpu
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(line) + '\n').encode())
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
