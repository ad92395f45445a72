"""Configuration of the B200 sampling controller.

`SamplerConfig` carries exactly the keys `Sampling_MPC.__init__` / `Centroidal_Model_JAX.__init__` read from the
reference's `quadruped_pympc/config.py` (config.py:84-109,181-196; mass/inertia config.py:19-72), so the reference's
config module can be passed straight in (`SamplerConfig.from_reference(config)`), and the backend is selected by one
extra key there: ``mpc_params['sampling_backend'] = 'b200'``.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Optional, Sequence

import numpy as np

# (mass, inertia) per robot name, config.py:19-72
_I_GO = ((1.58460467e-01, 1.21660000e-04, -1.55444692e-02),
         (1.21660000e-04, 4.68645637e-01, -3.12000000e-05),
         (-1.55444692e-02, -3.12000000e-05, 5.24474661e-01))
_I_ALIENGO = ((0.2310941359705289, -0.0014987128245817424, -0.021400468992761768),
              (-0.0014987128245817424, 1.4485084687476608, 0.0004641447134275615),
              (-0.021400468992761768, 0.0004641447134275615, 1.503217877350808))
_I_HYQREAL = ((4.55031444e+00, 2.75249434e-03, -5.11957307e-01),
              (2.75249434e-03, 2.02411774e+01, -7.38560592e-04),
              (-5.11957307e-01, -7.38560592e-04, 2.14269772e+01))
ROBOT_INERTIALS = {
    'go1': (12.019, _I_GO), 'go2': (15.019, _I_GO), 'aliengo': (24.637, _I_ALIENGO), 'b2': (83.49, _I_ALIENGO),
    'hyqreal1': (108.40, _I_HYQREAL), 'hyqreal2': (126.69, _I_HYQREAL), 'mini_cheetah': (12.5, _I_GO),
    'spot': (50.34, _I_ALIENGO), 'pegasus': (87.43, _I_ALIENGO),
}

# diag(Q), S/centroidal_nmpc_jax.py:118-130
Q_DIAG_DEFAULT = (0.0, 0.0, 1500.0, 200.0, 200.0, 200.0, 500.0, 500.0, 0.0, 20.0, 20.0, 50.0)


@dataclass
class SamplerConfig:
    mass: float = 15.019
    inertia: Sequence = field(default_factory=lambda: _I_GO)
    horizon: int = 12
    dt: float = 0.02
    grf_max: Optional[float] = None            # default mass * 9.81 (config.py:96)
    grf_min: float = 0.0
    mu: float = 0.5
    use_nonuniform_discretization: bool = False
    horizon_fine_grained: int = 2
    dt_fine_grained: float = 0.01
    optimize_step_freq: bool = False
    step_freq_available: Sequence[float] = (1.4, 2.0, 2.4)
    sampling_method: str = 'random_sampling'
    control_parametrization: str = 'cubic_spline'
    num_splines: int = 2
    num_parallel_computations: int = 10000
    num_sampling_iterations: int = 1
    sigma_cem_mppi: float = 3.0
    sigma_mppi: float = 3.0
    sigma_random_sampling: Sequence[float] = (0.2, 3.0, 10.0)
    shift_solution: bool = False
    # --- keys that only exist for this backend ---
    device: int = 0                  # CUDA ordinal (the reference's 'device': 'gpu'|'cpu' has no CPU meaning here)
    seed: int = 42                   # jax.random.PRNGKey(42), S/centroidal_nmpc_jax.py:167
    io_mode: str = 'direct'          # 'direct' (packet in kernel params, result stamped into mapped host memory) | 'memcpy' | 'mapped'
    use_graph: bool = True
    elite_k: int = 0                 # 0 = reference default (10 / 200)
    lanes_per_sample: int = 0        # kernel variant: 0 = auto, 1 = sample per thread, 2 = two samples per thread (packed FP32), 4 = four lanes per sample, 14 = 1 + noise helpers
    fast_binding: bool = True        # tick through the CPython binding (csrc/qpfast.c) when built; False: ctypes (same C ABI)
    q_diag: Sequence[float] = Q_DIAG_DEFAULT

    def __post_init__(self):
        if self.grf_max is None:
            self.grf_max = self.mass * 9.81
        self.inertia = np.asarray(self.inertia, dtype=np.float64).reshape(3, 3)

    @classmethod
    def for_robot(cls, robot: str, **kw) -> "SamplerConfig":
        mass, inertia = ROBOT_INERTIALS[robot]
        return cls(mass=mass, inertia=inertia, **kw)

    @classmethod
    def from_reference(cls, config_module, **overrides) -> "SamplerConfig":
        """Build from the reference's `quadruped_pympc.config` module (reads the same keys as S:27-33,55,98-105,160-164
        and M:38-53)."""
        mp = config_module.mpc_params
        kw = dict(
            mass=float(config_module.mass), inertia=np.asarray(config_module.inertia),
            horizon=int(mp['horizon']), dt=float(mp['dt']), grf_max=float(mp['grf_max']), grf_min=float(mp['grf_min']),
            mu=float(mp['mu']), use_nonuniform_discretization=bool(mp['use_nonuniform_discretization']),
            horizon_fine_grained=int(mp['horizon_fine_grained']), dt_fine_grained=float(mp['dt_fine_grained']),
            optimize_step_freq=bool(mp['optimize_step_freq']), step_freq_available=tuple(mp['step_freq_available']),
            sampling_method=mp['sampling_method'], control_parametrization=mp['control_parametrization'],
            num_splines=int(mp['num_splines']), num_parallel_computations=int(mp['num_parallel_computations']),
            num_sampling_iterations=int(mp['num_sampling_iterations']), sigma_cem_mppi=float(mp['sigma_cem_mppi']),
            sigma_mppi=float(mp['sigma_mppi']), sigma_random_sampling=tuple(mp['sigma_random_sampling']),
            shift_solution=bool(mp['shift_solution']),
        )
        for k in ('b200_device', 'b200_seed', 'b200_io_mode'):
            if k in mp:
                kw[k[5:]] = mp[k]
        kw.update(overrides)
        return cls(**kw)

    # S/centroidal_nmpc_jax.py:52-93
    @property
    def params_per_leg(self) -> int:
        if self.control_parametrization == 'linear_spline':
            return (self.num_splines + 1) * 3
        if self.control_parametrization == 'cubic_spline':
            return 4 * 3 * self.num_splines
        return self.horizon * 3

    @property
    def num_control_parameters(self) -> int:
        return 4 * self.params_per_leg

    def dts(self):
        """M/centroidal_model_jax.py:42-53"""
        if self.use_nonuniform_discretization:
            return [self.dt_fine_grained] * self.horizon_fine_grained + \
                   [self.dt] * (self.horizon - self.horizon_fine_grained)
        return [self.dt] * self.horizon
