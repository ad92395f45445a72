"""Resident-tick kernel times of the current tree for the launch shapes the A/B probe of an older tree also runs."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
from quadruped_pympc_b200 import _lib
lib = _lib.load()
for R, N, lanes in [(512, 10000, 1), (512, 10000, 2), (1, 10000, 0), (1, 10000, 1), (1, 1000, 0), (1, 100000, 1), (1, 1000000, 1)]:
    h, keep = bench.make_resident_handle(0, R, 0, n_samples=N, lanes=lanes)
    name = lib.qpmpc_kernel_name(h).decode()
    bench.run_resident(h, 1, 3)
    ms, _ = bench.run_resident(h, 10 if R * N > 2000000 else 200, 0)
    lib.qpmpc_destroy(h)
    print(f'NEW R={R} N={N} lanes={lanes} {name}: median {np.median(ms)*1e3:.1f} us mean {ms.mean()*1e3:.1f}', flush=True)
