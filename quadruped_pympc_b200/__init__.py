"""quadruped_pympc_b200 -- B200-native sampling-based centroidal MPC (drop-in for Quadruped-PyMPC's sampler).

Only what the hot path needs lives here: `csrc/` (CUDA kernels + C ABI), the ctypes binding, and the host-side
mirror of the reference's `Sampling_MPC` surface.  (The directory name uses an underscore because
`quadruped-pympc_b200` is not an importable Python identifier.)
"""
from .config import SamplerConfig, ROBOT_INERTIALS  # noqa: F401
from .sampling_mpc import Sampling_MPC, LazyCosts  # noqa: F401
from ._lib import QpmpcError, build, load, device_count  # noqa: F401

__all__ = ['SamplerConfig', 'Sampling_MPC', 'LazyCosts', 'QpmpcError', 'build', 'load', 'device_count',
           'ROBOT_INERTIALS']
