"""Kernel-variant comparison on resident inputs (CUDA events, L2 flushed): sample per thread (lanes 1) vs two samples per
thread in packed FP32 (lanes 2) for the batched launch and the single-controller sample sweep.
    python tools/probe_throughput.py > gpurun_out/probe_throughput.log"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
from quadruped_pympc_b200 import _lib

lib = _lib.load()
peak = bench.measure_fp32_peak(0)
print(f'fp32 peak (FFMA microbenchmark) {peak:.2f} TFLOP/s')
flop_per_sample = bench.FLOP_PER_TICK / bench.N_SAMPLES
cases = [(512, 10000), (64, 10000), (8, 10000), (1, 50000), (1, 100000), (1, 300000), (1, 1000000)]
for R, N in cases:
    for lanes in (1, 2):
        h, keep = bench.make_resident_handle(0, R, 0, n_samples=N, lanes=lanes)
        name = lib.qpmpc_kernel_name(h).decode()
        bench.run_resident(h, 1, 3)
        ms, _ = bench.run_resident(h, 10 if R * N > 2000000 else 30, 0)
        lib.qpmpc_destroy(h)
        t = float(np.median(ms)) * 1e-3
        print(f'R={R:4d} N={N:8d} lanes={lanes} {name:15s}: {t * 1e6:9.1f} us  {R * N * bench.HORIZON / t:.3e} rollout-steps/s  '
              f'{R * N * flop_per_sample / t / 1e12:6.2f} TFLOP/s = {R * N * flop_per_sample / t / 1e12 / peak:.3f} of peak', flush=True)
