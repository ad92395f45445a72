"""Quick latency probe (not the bench): e2e tick p50/p99 for a few transport modes + device-only kernel time."""
import sys, os, time, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import quadruped_pympc_b200 as qp
from quadruped_pympc_b200 import synthetic, _lib

def run(method='mppi', param='cubic_spline', N=10000, io_mode='memcpy', use_graph=True, ticks=2000, adaptive=False, robot='go2'):
    cfg = qp.SamplerConfig.for_robot(robot, sampling_method=method, control_parametrization=param,
                                     num_parallel_computations=N, io_mode=io_mode, use_graph=use_graph, optimize_step_freq=adaptive)
    c = qp.Sampling_MPC(cfg)
    sc = synthetic.scenario(robot, 'trot', seed=0)
    best = np.zeros(c.num_control_parameters, np.float32)
    sig = (3.0,) if method == 'cem_mppi' else ()
    lat = []
    for i in range(ticks + 50):
        c.with_newkey()
        t0 = time.perf_counter_ns()
        out = c.jitted_compute_control(sc['state'], sc['reference'], sc['contact_sequence'], best, c.master_key, *sig, sc['phase_signal'], 1.4, 1)
        t1 = time.perf_counter_ns()
        best = out[3]
        if i >= 50: lat.append((t1 - t0) * 1e-3)
    lat = np.array(lat)
    # device-only
    lib = _lib.load(); h = c._ensure(); b = c._buf
    _lib.check(lib.qpmpc_upload_resident(h, C.byref(b.cin)))
    ms = np.zeros(200, np.float32)
    _lib.check(lib.qpmpc_bench_resident(h, 200, 0, ms.ctypes.data_as(_lib._FP)))
    msf = np.zeros(50, np.float32)
    _lib.check(lib.qpmpc_bench_resident(h, 50, 1, msf.ctypes.data_as(_lib._FP)))
    print(f'{robot} {method} {param} N={N} adaptive={adaptive} io={io_mode} graph={use_graph}: e2e p50 {np.percentile(lat,50):.1f} us p99 {np.percentile(lat,99):.1f} us | '
          f'kernel p50 {np.percentile(ms,50)*1e3:.1f} us (flushed {np.percentile(msf,50)*1e3:.1f} us) best_cost {float(out[4]):.3f}', flush=True)

if __name__ == '__main__':
    tf = C.c_float(0)
    _lib.check(_lib.load().qpmpc_measure_fp32_peak(0, C.byref(tf)))
    print('fp32 peak TFLOP/s', tf.value, flush=True)
    run(io_mode='memcpy', use_graph=True)
    run(io_mode='memcpy', use_graph=False)
    run(io_mode='mapped', use_graph=False)
    run(method='random_sampling', param='zero_order', robot='aliengo')
    run(method='cem_mppi', N=50000, adaptive=True, robot='hyqreal1', ticks=300)
    for N in (1000, 100000, 1000000):
        run(N=N, ticks=200)
