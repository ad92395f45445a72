"""Latency / throughput probe (not the bench): kernel variants x sample counts x robots, with phase breakdown.

    python tools/probe_latency.py [quick|sweep|phases|e2e]
"""
import ctypes as C
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import quadruped_pympc_b200 as qp
from quadruped_pympc_b200 import synthetic, _lib
from quadruped_pympc_b200.sampling_mpc import make_c_config


def resident(method='mppi', param='cubic_spline', N=10000, R=1, lanes=0, adaptive=False, robot='go2', steps=50, flush=1):
    """-> (median kernel ms, phases dict)"""
    lib = _lib.load()
    cfg = qp.SamplerConfig.for_robot(robot, sampling_method=method, control_parametrization=param,
                                     num_parallel_computations=N, optimize_step_freq=adaptive, lanes_per_sample=lanes)
    c = make_c_config(cfg, R)
    h = C.c_void_p()
    _lib.check(lib.qpmpc_create(C.byref(c), C.byref(h)))
    P = cfg.num_control_parameters
    ins = (_lib.QpmpcTickIn * R)()
    keep = []
    fp = lambda a: a.ctypes.data_as(_lib._FP)
    for r in range(R):
        sc = synthetic.scenario(robot, 'trot', seed=r)
        arrs = [sc['state'].astype(np.float32), sc['reference'].astype(np.float32),
                np.ascontiguousarray(sc['contact_sequence'], np.float32), np.zeros(P, np.float32),
                sc['phase_signal'].astype(np.float32), np.full(P, 3.0, np.float32)]
        keep += arrs
        i = ins[r]
        i.state, i.reference, i.contact_sequence, i.contact_ld, i.best = fp(arrs[0]), fp(arrs[1]), fp(arrs[2]), 12, fp(arrs[3])
        i.phase_signal, i.sigma, i.nominal_step_freq, i.optimize_swing = fp(arrs[4]), fp(arrs[5]), 1.4, 1
        i.key[0], i.key[1] = 42, 1
    _lib.check(lib.qpmpc_upload_resident(h, ins))
    ms = np.zeros(steps, np.float32)
    _lib.check(lib.qpmpc_bench_resident(h, 5, flush, ms.ctypes.data_as(_lib._FP)))
    _lib.check(lib.qpmpc_bench_resident(h, steps, flush, ms.ctypes.data_as(_lib._FP)))
    ph = (C.c_uint64 * 16)()
    _lib.check(lib.qpmpc_debug_phase_times(h, ph))
    ph = np.array(ph[:], np.int64)
    lib.qpmpc_destroy(h)
    return float(np.median(ms)), ph


def fmt_phases(ph):
    b0 = ph[0]
    names = {1: 'staged', 2: 'noise', 3: 'rollout', 4: 'reduced', 5: 'published', 8: 'last:ticket', 9: 'last:merged',
             6: 'f:tail', 7: 'f:elites', 10: 'last:best', 11: 'last:done', 12: 'm:argmin', 13: 'm:chunk0sum', 14: 'm:vdone', 15: 'last:publish'}
    return ' '.join(f'{names[k]}@{(ph[k] - b0) / 1e3:.1f}' for k in sorted(names) if ph[k] > 0)


def e2e(method='mppi', param='cubic_spline', N=10000, lanes=0, io_mode='direct', use_graph=True, ticks=2000):
    cfg = qp.SamplerConfig.for_robot('go2', sampling_method=method, control_parametrization=param,
                                     num_parallel_computations=N, io_mode=io_mode, use_graph=use_graph, lanes_per_sample=lanes)
    c = qp.Sampling_MPC(cfg)
    sc = synthetic.scenario('go2', 'trot', seed=0)
    best = np.zeros(c.num_control_parameters, np.float32)
    lat = []
    for i in range(ticks + 50):
        c.with_newkey()
        t0 = time.perf_counter_ns()
        out = c.jitted_compute_control(sc['state'], sc['reference'], sc['contact_sequence'], best, c.master_key,
                                       sc['phase_signal'], 1.4, 1)
        t1 = time.perf_counter_ns()
        best = out[3]
        if i >= 50:
            lat.append((t1 - t0) * 1e-3)
    lat = np.array(lat)
    # the same tick through the bare C entry point (structs already filled): what the Python wrapper adds
    lib, h, b = _lib.load(), c._ensure(), c._buf
    clat = []
    for i in range(ticks):
        t0 = time.perf_counter_ns()
        lib.qpmpc_tick(h, C.byref(b.cin), C.byref(b.cout))
        clat.append((time.perf_counter_ns() - t0) * 1e-3)
    print(f'    bare qpmpc_tick: p50 {np.percentile(clat, 50):.1f} us  p99 {np.percentile(clat, 99):.1f} us')
    if io_mode == 'direct':                # host-side timeline of the direct-I/O tick
        ns = (C.c_uint64 * 4)()
        lib.qpmpc_debug_host_times(h, 1, None)
        seg = []
        for i in range(ticks):
            t0 = time.perf_counter_ns()
            lib.qpmpc_tick(h, C.byref(b.cin), C.byref(b.cout))
            t1 = time.perf_counter_ns()
            lib.qpmpc_debug_host_times(h, 1, ns)
            seg.append([ns[1] - ns[0], ns[2] - ns[1], ns[3] - ns[2], (t1 - t0) - (ns[3] - ns[0])])
        lib.qpmpc_debug_host_times(h, 0, None)
        m = np.median(np.array(seg, dtype=np.float64), axis=0) * 1e-3
        print(f'    host timeline (median us): packet + cudaLaunchKernel {m[0]:.1f} | launch -> stamp seen {m[1]:.1f} | '
              f'unpack {m[2]:.1f} | pack_one + ctypes {m[3]:.1f}')
    print(f'e2e N={N} lanes={lanes} io={io_mode} graph={use_graph}: p50 {np.percentile(lat, 50):.1f} us  p99 '
          f'{np.percentile(lat, 99):.1f} us  best_cost {float(out[4]):.3f}', flush=True)


def batched_e2e(R=512, N=10000, ticks=30):
    """host arrays in / out through BatchedSampling_MPC.compute_control, with the host-side timeline of the C call"""
    from quadruped_pympc_b200.batched import BatchedSampling_MPC
    cfg = qp.SamplerConfig.for_robot('go2', sampling_method='mppi', control_parametrization='cubic_spline', num_parallel_computations=N)
    ctrl = BatchedSampling_MPC(cfg, R)
    scs = [synthetic.scenario('go2', 'trot', seed=r) for r in range(R)]
    st = np.stack([s['state'] for s in scs]).astype(np.float32)
    rf = np.stack([s['reference'] for s in scs]).astype(np.float32)
    cs = np.stack([s['contact_sequence'] for s in scs]).astype(np.float32)
    best = np.zeros((R, ctrl.num_control_parameters), np.float32)
    lib, h = _lib.load(), ctrl._ensure()
    lib.qpmpc_debug_host_times(h, 1, None)
    ns = (C.c_uint64 * 4)()
    rows = []
    for i in range(ticks + 3):
        ctrl.with_newkey()
        t0 = time.perf_counter_ns()
        out = ctrl.compute_control(st, rf, cs, best)
        t1 = time.perf_counter_ns()
        best = out[2]
        lib.qpmpc_debug_host_times(h, 1, ns)
        if i >= 3:
            rows.append([t1 - t0, ns[1] - ns[0], ns[2] - ns[1], ns[3] - ns[2], (t1 - t0) - (ns[3] - ns[0])])
    m = np.median(np.array(rows, dtype=np.float64), axis=0) * 1e-3
    print(f'batched e2e R={R} N={N}: call {m[0]:.0f} us = validate + pack + issue {m[1]:.0f} | wait for the last chunk {m[2]:.0f} | '
          f'unpack of the last chunk {m[3]:.0f} | flat wrapper + python {m[4]:.0f}', flush=True)


if __name__ == '__main__':
    what = sys.argv[1] if len(sys.argv) > 1 else 'quick'
    tf = C.c_float(0)
    _lib.check(_lib.load().qpmpc_measure_fp32_peak(0, C.byref(tf)))
    print('fp32 peak TFLOP/s', tf.value, flush=True)
    if what in ('one4', 'one1'):
        ms, ph = resident(lanes=int(what[3]), steps=10)
        print(f'config2 lanes={what[3]}: kernel {ms * 1e3:.1f} us | {fmt_phases(ph)}', flush=True)
    if what in ('quick', 'phases'):
        for lanes in (1, 14, 4):
            ms, ph = resident(lanes=lanes)
            print(f'config2 lanes={lanes}: kernel {ms * 1e3:.1f} us | {fmt_phases(ph)}', flush=True)
        for lanes in (1, 14, 4):
            ms, ph = resident(method='random_sampling', param='zero_order', robot='aliengo', lanes=lanes)
            print(f'config1 lanes={lanes}: kernel {ms * 1e3:.1f} us | {fmt_phases(ph)}', flush=True)
        for lanes in (1, 4):
            ms, ph = resident(method='cem_mppi', N=50000, adaptive=True, robot='hyqreal1', lanes=lanes, steps=20)
            print(f'config3 lanes={lanes}: kernel {ms * 1e3:.1f} us | {fmt_phases(ph)}', flush=True)
    if what in ('quick', 'sweep'):
        for N in (1000, 2000, 4000, 6000, 8000, 10000, 14000, 18000, 20000, 40000, 100000, 1000000):
            row = []
            for lanes in (1, 4, 14):            # 14 = sample per thread + noise helpers (falls back to 1 above 148 tiles)
                ms, _ = resident(N=N, lanes=lanes, steps=30)
                row.append(ms * 1e3)
            print(f'N={N:8d}  R=1   lanes1 {row[0]:9.1f} us   lanes4 {row[1]:9.1f} us   lanes14 {row[2]:9.1f} us', flush=True)
        for R in (2, 8, 32, 128, 512):
            row = []
            for lanes in (1, 4):
                ms, _ = resident(N=10000, R=R, lanes=lanes, steps=10)
                row.append(ms * 1e3)
            print(f'N=   10000  R={R:<4d} lanes1 {row[0]:9.1f} us   lanes4 {row[1]:9.1f} us', flush=True)
    if what in ('quick', 'batched'):
        batched_e2e()
    if what in ('quick', 'e2e'):
        for lanes in (1, 14, 4):
            e2e(lanes=lanes)
        e2e(lanes=1, io_mode='memcpy')
        e2e(lanes=1, io_mode='memcpy', use_graph=False)
