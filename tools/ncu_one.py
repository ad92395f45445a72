"""One resident launch shape for an ncu capture: python tools/ncu_one.py R N lanes [ticks] [config0|config1|config2]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from quadruped_pympc_b200 import _lib
R, N, lanes = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
ticks = int(sys.argv[4]) if len(sys.argv) > 4 else 4
w = bench.WORKLOADS[sys.argv[5]] if len(sys.argv) > 5 else bench.HEAD
h, keep = bench.make_resident_handle(0, R, 0, n_samples=N, lanes=lanes, w=w)
print(_lib.load().qpmpc_kernel_name(h).decode())
ms, _ = bench.run_resident(h, ticks, 0, flush=0)
print(ms)
