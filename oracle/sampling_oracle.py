"""CPU oracle for the sampling-based centroidal MPC hot path of Quadruped-PyMPC.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is product code: only
``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import it, and only as the checker or the timed CPU
baseline.  The product path (``quadruped_pympc_b200``) never imports this module.

PARITY PINNED by the reference's own source: ``tools/make_golden.py`` executes the
reference's UNMODIFIED controllers/sampling/*.py, centroidal_model_jax.py,
helpers/periodic_gait_generator_jax.py and config.py from /root/reference under a
NumPy stand-in for the ~40 jax entry points they use (``oracle/jaxshim``; jax/jaxlib
0.9.2, gym_quadruped and mujoco are not installable in this image -- no network) and
freezes 22 tick fixtures + host-side fixtures under ``tests/golden/``.  This file and
its C twin reproduce every fixture (costs rel 2e-6, argmin exact, knots/GRFs/sigma
1e-5; tests/test_golden.py).  What the fixtures pin: formulas, indexing, op order and
control flow of the reference as written.  What they cannot pin: XLA's own fusion /
re-association (covered by the north_star's float32 tolerance) and threefry bit
streams (noise is injected, as the north_star prescribes).
This file is an op-for-op NumPy restatement of the reference's Python, vectorised
over the sample axis, with every function citing the reference file:line it follows
(paths relative to /root/reference/quadruped_pympc/).  Further cross-checks: (i) a
float64 twin of itself, (ii) an independent statement of the same SRBD equations
(controllers/gradient/nominal/centroidal_model_nominal.py:201-272, restated in
tests/test_oracle.py with numpy.linalg), (iii) analytic properties and (iv) the C
twin in oracle/sampling_oracle.c.

Conventions
-----------
* ``S/`` = controllers/sampling/ ; ``A/`` = the gait-adaptive twin
  controllers/sampling/centroidal_nmpc_jax_gait_adaptive.py.
* All arithmetic is done in ``cfg.dtype`` (np.float32 = the reference's
  ``dtype_general``; np.float64 = the twin used to bound rounding error).
* Every per-sample scalar of the reference is an array of shape (N,) here; the
  reference's ``vmap`` (S/centroidal_nmpc_jax.py:176) is the leading axis.
* Noise is always *injected*: the oracle never draws random numbers for the
  controller itself (north_star protocol: compare on identical noise tensors).
"""
from __future__ import annotations

import dataclasses
from dataclasses import dataclass, field
from typing import Optional, Sequence

import numpy as np

# --------------------------------------------------------------------------------------
# Robot constants, config.py:19-72 (mass, inertia per robot name)
# --------------------------------------------------------------------------------------
_I_GO = np.array([[1.58460467e-01, 1.21660000e-04, -1.55444692e-02],
                  [1.21660000e-04, 4.68645637e-01, -3.12000000e-05],
                  [-1.55444692e-02, -3.12000000e-05, 5.24474661e-01]])
_I_ALIENGO = np.array([[0.2310941359705289, -0.0014987128245817424, -0.021400468992761768],
                       [-0.0014987128245817424, 1.4485084687476608, 0.0004641447134275615],
                       [-0.021400468992761768, 0.0004641447134275615, 1.503217877350808]])
_I_HYQREAL = np.array([[4.55031444e+00, 2.75249434e-03, -5.11957307e-01],
                       [2.75249434e-03, 2.02411774e+01, -7.38560592e-04],
                       [-5.11957307e-01, -7.38560592e-04, 2.14269772e+01]])
ROBOTS = {
    'go1': (12.019, _I_GO), 'go2': (15.019, _I_GO), 'aliengo': (24.637, _I_ALIENGO),
    'b2': (83.49, _I_ALIENGO), 'hyqreal1': (108.40, _I_HYQREAL), 'hyqreal2': (126.69, _I_HYQREAL),
    'mini_cheetah': (12.5, _I_GO), 'spot': (50.34, _I_ALIENGO), 'pegasus': (87.43, _I_ALIENGO),
}


@dataclass
class OracleConfig:
    """The subset of config.py the sampler reads (config.py:84-109,181-196)."""
    mass: float = 15.019
    inertia: np.ndarray = field(default_factory=lambda: _I_GO.copy())
    horizon: int = 12                       # config.py:92
    dt: float = 0.02                        # config.py:93
    grf_max: Optional[float] = None         # config.py:96  (mass * 9.81)
    grf_min: float = 0.0                    # config.py:97
    mu: float = 0.5                         # config.py:98
    use_nonuniform_discretization: bool = False   # config.py:101
    horizon_fine_grained: int = 2
    dt_fine_grained: float = 0.01
    optimize_step_freq: bool = False        # config.py:108  -> adaptive twin
    step_freq_available: Sequence[float] = (1.4, 2.0, 2.4)  # config.py:109
    sampling_method: str = 'random_sampling'       # config.py:184
    control_parametrization: str = 'cubic_spline'  # config.py:185
    num_splines: int = 2                    # config.py:186
    num_parallel_computations: int = 10000  # config.py:187
    num_sampling_iterations: int = 1
    sigma_cem_mppi: float = 3.0
    sigma_mppi: float = 3.0
    sigma_random_sampling: Sequence[float] = (0.2, 3.0, 10.0)
    shift_solution: bool = False
    dtype: type = np.float32                # S/centroidal_nmpc_jax.py:17

    def __post_init__(self):
        if self.grf_max is None:
            self.grf_max = self.mass * 9.81
        self.inertia = np.asarray(self.inertia, dtype=np.float64).reshape(3, 3)

    @classmethod
    def for_robot(cls, robot: str, **kw) -> "OracleConfig":
        m, inertia = ROBOTS[robot]
        return cls(mass=m, inertia=inertia.copy(), **kw)

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
        return self.params_per_leg * 4


# Q diagonal, S/centroidal_nmpc_jax.py:118-130 (the remaining 12 entries are 0)
Q_DIAG = np.array([0, 0, 1500, 200, 200, 200, 500, 500, 0, 20, 20, 50] + [0] * 12, dtype=np.float64)

MAX_SAMPLING_FORCES_X = 10   # S/centroidal_nmpc_jax.py:39-41
MAX_SAMPLING_FORCES_Y = 10
MAX_SAMPLING_FORCES_Z = 30
PGG_DUTY_ADAPTIVE = 0.65     # A:179 (hard-coded duty factor of the in-rollout gait generator)


class SamplingOracle:
    """NumPy restatement of ``Sampling_MPC`` + ``Centroidal_Model_JAX`` (+ adaptive twin)."""

    def __init__(self, cfg: OracleConfig):
        self.cfg = cfg
        t = self.t = cfg.dtype
        self.N = int(cfg.num_parallel_computations)
        self.H = int(cfg.horizon)
        self.ns = int(cfg.num_splines)
        self.P_leg = cfg.params_per_leg
        self.P = cfg.num_control_parameters
        self.param = cfg.control_parametrization
        self.adaptive = bool(cfg.optimize_step_freq)
        # model constants, S/centroidal_model_jax.py:38-56
        self.mass = cfg.mass                                  # python float (weakly typed)
        self.inertia = cfg.inertia.astype(t)
        self.inertia_inv = self.calculate_inverse(self.inertia)
        if cfg.use_nonuniform_discretization:                  # S/centroidal_model_jax.py:42-51
            fine = np.tile(t(cfg.dt_fine_grained), cfg.horizon_fine_grained)
            self.dts = np.concatenate((fine, np.tile(t(cfg.dt), self.H - cfg.horizon_fine_grained))).astype(t)
        else:                                                  # :53
            self.dts = np.tile(t(cfg.dt), self.H).astype(t)
        self.Q = np.diag(Q_DIAG).astype(t)                     # S/centroidal_nmpc_jax.py:118-130
        self.mu = t(cfg.mu)
        self.f_z_max = t(cfg.grf_max)
        self.f_z_min = t(cfg.grf_min)
        self.mg = t(self.mass * 9.81)   # folded in float64 on the host, then cast (Appendix A.1)
        self.scale_xy = t(MAX_SAMPLING_FORCES_Z / MAX_SAMPLING_FORCES_X)   # = 3.0
        self.step_freq_delta = np.asarray(cfg.step_freq_available, dtype=t)  # A:182

    # ---------------------------------------------------------------- model
    def calculate_inverse(self, A):
        """3x3 inverse by cofactors; S/centroidal_model_jax.py:67-91. A[...,3,3] or nested lists of (N,)."""
        a11, a12, a13 = A[0][0], A[0][1], A[0][2]
        a21, a22, a23 = A[1][0], A[1][1], A[1][2]
        a31, a32, a33 = A[2][0], A[2][1], A[2][2]
        DET = a11 * (a33 * a22 - a32 * a23) - a21 * (a33 * a12 - a32 * a13) + a31 * (a23 * a12 - a22 * a13)
        rows = [
            [(a33 * a22 - a32 * a23), -(a33 * a12 - a32 * a13), (a23 * a12 - a22 * a13)],
            [-(a33 * a21 - a31 * a23), (a33 * a11 - a31 * a13), -(a23 * a11 - a21 * a13)],
            [(a32 * a21 - a31 * a22), -(a32 * a11 - a31 * a12), (a22 * a11 - a21 * a12)],
        ]
        out = [[rows[i][j] / DET for j in range(3)] for i in range(3)]
        if isinstance(A, np.ndarray) and A.ndim == 2:
            return np.array(out, dtype=A.dtype)
        return out

    @staticmethod
    def _dot3(row, v):
        """3-term dot product in index order (jnp.dot of a length-3 row)."""
        return (row[0] * v[0] + row[1] * v[1]) + row[2] * v[2]

    def fd(self, states, inputs, contact_status):
        """State derivative, S/centroidal_model_jax.py:93-162.

        states: list of 24 arrays (N,) (or scalars), inputs: list of 24, contact_status: list of 4.
        Returns list of 12."""
        t = self.t
        foot = [states[12 + 3 * l:15 + 3 * l] for l in range(4)]        # :102
        force = [inputs[12 + 3 * l:15 + 3 * l] for l in range(4)]       # :103
        com = states[0:3]
        stance = contact_status
        lin_vel = states[3:6]                                            # :108
        # :111-116  temp = F_fl*s_fl + F_fr*s_fr + F_rl*s_rl + F_rr*s_rr
        temp = [((force[0][k] * stance[0] + force[1][k] * stance[1]) + force[2][k] * stance[2])
                + force[3][k] * stance[3] for k in range(3)]
        gravity = [t(0), t(0), t(-9.81)]                                 # :117
        inv_m = t(1) / t(self.mass)                                      # :118
        lin_acc = [inv_m * temp[k] + gravity[k] for k in range(3)]
        w = states[9:12]                                                 # :121
        roll, pitch, yaw = states[6], states[7], states[8]               # :122
        sr, cr = np.sin(roll), np.cos(roll)
        sp, cp = np.sin(pitch), np.cos(pitch)
        sy, cy = np.sin(yaw), np.cos(yaw)
        zero = np.zeros_like(np.asarray(roll, dtype=t))
        one = zero + t(1)
        E = [[one, zero, -sp], [zero, cr, cp * sr], [zero, -sr, cp * cr]]   # :124-130
        # :132-135   skew(r) @ F * stance, r = foot - com
        temp2 = None
        for l in range(4):
            r = [foot[l][k] - com[k] for k in range(3)]
            F = force[l]
            sk = [(zero * F[0] + (-r[2]) * F[1]) + r[1] * F[2],
                  (r[2] * F[0] + zero * F[1]) + (-r[0]) * F[2],
                  ((-r[1]) * F[0] + r[0] * F[1]) + zero * F[2]]
            term = [sk[k] * stance[l] for k in range(3)]
            temp2 = term if temp2 is None else [temp2[k] + term[k] for k in range(3)]
        Einv = self.calculate_inverse(E)                                 # :137
        euler_rates = [self._dot3(Einv[i], w) for i in range(3)]
        b_R_w = [                                                        # :141-155
            [cp * cy, cp * sy, -sp],
            [sr * sp * cy - cr * sy, sr * sp * sy + cr * cy, sr * cp],
            [cr * sp * cy + sr * sy, cr * sp * sy - sr * cy, cr * cp],
        ]
        I, Iinv = self.inertia, self.inertia_inv
        Iw = [self._dot3(I[i], w) for i in range(3)]                     # :157
        wxIw = [(zero * Iw[0] + (-w[2]) * Iw[1]) + w[1] * Iw[2],
                (w[2] * Iw[0] + zero * Iw[1]) + (-w[0]) * Iw[2],
                ((-w[1]) * Iw[0] + w[0] * Iw[1]) + zero * Iw[2]]
        t1 = [self._dot3(Iinv[i], wxIw) for i in range(3)]
        Rt = [self._dot3(b_R_w[i], temp2) for i in range(3)]             # :158
        t2 = [self._dot3(Iinv[i], Rt) for i in range(3)]
        ang_acc = [-t1[i] + t2[i] for i in range(3)]
        return list(lin_vel) + lin_acc + euler_rates + ang_acc           # :162

    def integrate(self, state, inputs, contact_status, n):
        """Explicit Euler, S/centroidal_model_jax.py:164-174."""
        d = self.fd(state, inputs, contact_status)
        dt = self.dts[n]                                                 # :171
        new = [state[i] + d[i] * dt for i in range(12)]                  # :172
        return new + list(state[12:])                                    # :174

    # ---------------------------------------------------------------- splines
    def _chunk_index(self, step):
        """index = max(where(step >= linspace(0,H,ns+1), arange(ns+1), 0)); S/centroidal_nmpc_jax.py:187-189."""
        t = self.t
        bounds = np.linspace(0, self.H, self.ns + 1).astype(t)
        stepf = np.asarray(step).astype(t)
        idx = np.zeros(np.shape(stepf), dtype=np.int32)
        for i in range(self.ns + 1):
            idx = np.maximum(idx, np.where(stepf >= bounds[i], np.int32(i), np.int32(0)))
        return idx

    def _tau(self, step, horizon_leg, index):
        """tau = step/(horizon_leg/ns) - index ; S/centroidal_nmpc_jax.py:191-194 (=214-217)."""
        t = self.t
        if isinstance(horizon_leg, (int, float)):
            denom = t(horizon_leg / self.ns)           # python-level division, then weak cast
        else:
            denom = np.asarray(horizon_leg, dtype=t) / t(self.ns)
        tau = np.asarray(step).astype(t) / denom
        return tau - index.astype(t)

    @staticmethod
    def _take(parameters, idx):
        """parameters: (N, P_leg); idx: int array (N,) or scalar -> (N,) gather (numpy wrap for negatives)."""
        idx = np.asarray(idx)
        if idx.ndim == 0:
            return parameters[:, int(idx)]
        return np.take_along_axis(parameters, idx[:, None].astype(np.int64) % parameters.shape[1], axis=1)[:, 0]

    def compute_linear_spline(self, parameters, step, horizon_leg):
        """S/centroidal_nmpc_jax.py:181-201."""
        t = self.t
        index = self._chunk_index(step)
        q = self._tau(step, horizon_leg, index)
        shift = self.ns + 1
        g = lambda o: self._take(parameters, index + o)
        f_x = (t(1) - q) * g(0) + q * g(1)
        f_y = (t(1) - q) * g(shift) + q * g(shift + 1)
        f_z = (t(1) - q) * g(shift * 2) + q * g(shift * 2 + 1)
        return f_x, f_y, f_z

    def compute_cubic_spline(self, parameters, step, horizon_leg):
        """S/centroidal_nmpc_jax.py:204-257 (start_index = 10*index, 12 knots read)."""
        t = self.t
        index = self._chunk_index(step)
        q = self._tau(step, horizon_leg, index)
        start = 10 * index                                             # :219
        a = t(2) * q * q * q - t(3) * q * q + t(1)                     # :222
        b = (q * q * q - t(2) * q * q + q) * t(1.0)                    # :223
        c = t(-2) * q * q * q + t(3) * q * q                           # :224
        d = (q * q * q - q * q) * t(1.0)                               # :225
        g = lambda o: self._take(parameters, start + o)
        out = []
        for o in (0, 4, 8):                                            # :227-255
            p0, p1, p2, p3 = g(o), g(o + 1), g(o + 2), g(o + 3)
            phi = t(0.5) * (((p2 - p1) / t(1.0)) + ((p1 - p0) / t(1.0)))
            phi_next = t(0.5) * (((p3 - p2) / t(1.0)) + ((p2 - p1) / t(1.0)))
            out.append(a * p1 + b * phi + c * p2 + d * phi_next)
        return out[0], out[1], out[2]

    def compute_zero_order_spline(self, parameters, step, horizon_leg):
        """S/centroidal_nmpc_jax.py:259-268 (index = int16(step))."""
        index = np.asarray(step).astype(np.int16).astype(np.int32)
        return (self._take(parameters, index), self._take(parameters, index + self.H),
                self._take(parameters, index + self.H * 2))

    def spline(self, parameters, step, horizon_leg):
        if self.param == 'linear_spline':
            return self.compute_linear_spline(parameters, step, horizon_leg)
        if self.param == 'cubic_spline':
            return self.compute_cubic_spline(parameters, step, horizon_leg)
        return self.compute_zero_order_spline(parameters, step, horizon_leg)

    # ---------------------------------------------------------------- constraints
    def enforce_force_constraints(self, F):
        """S/centroidal_nmpc_jax.py:270-314.  F: list of 4 [fx, fy, fz]; same where-order as the reference."""
        mu, fmin, fmax = self.mu, self.f_z_min, self.f_z_max
        out = []
        fz_all = [np.where(F[l][2] > fmin, F[l][2], fmin) for l in range(4)]      # :278-281
        fz_all = [np.where(fz_all[l] < fmax, fz_all[l], fmax) for l in range(4)]  # :284-287
        for l in range(4):                                                        # :294-312
            fz = fz_all[l]
            fx = np.where(F[l][0] > -mu * fz, F[l][0], -mu * fz)
            fx = np.where(fx < mu * fz, fx, mu * fz)
            fy = np.where(F[l][1] > -mu * fz, F[l][1], -mu * fz)
            fy = np.where(fy < mu * fz, fy, mu * fz)
            out.append([fx, fy, fz])
        return out

    def _forces_at(self, legs_params, steps, horizon_legs, contact_n):
        """Spline -> gravity compensation -> contact mask/scale -> clamp.
        S/centroidal_nmpc_jax.py:353-409 (and the identical tails :707-745)."""
        with np.errstate(divide='ignore', invalid='ignore', over='ignore'):
            F = [list(self.spline(legs_params[l], steps[l], horizon_legs[l])) for l in range(4)]
            nst = ((contact_n[0] + contact_n[1]) + contact_n[2]) + contact_n[3]      # :377-379
            fref = self.mg / nst                                                   # :380
            for l in range(4):
                F[l][2] = fref + F[l][2]                                           # :382-385
            for l in range(4):                                                     # :388-402
                F[l][0] = F[l][0] * contact_n[l] / self.scale_xy
                F[l][1] = F[l][1] * contact_n[l] / self.scale_xy
                F[l][2] = F[l][2] * contact_n[l]
            return self.enforce_force_constraints(F)

    # ---------------------------------------------------------------- gait generator (adaptive)
    def pgg_contact_sequence(self, timing, step_freq):
        """PeriodicGaitGeneratorJax.compute_contact_sequence; helpers/periodic_gait_generator_jax.py:68-89,136-151.
        timing: (4,), step_freq: (N,) -> contact (4, H, N) in cfg.dtype."""
        t = self.t
        f = np.asarray(step_freq, dtype=t)
        tt = [np.full(f.shape, t(timing[l]), dtype=t) for l in range(4)]
        seq = np.zeros((4, self.H) + f.shape, dtype=t)
        mpc_dt, duty = t(self.cfg.dt), t(PGG_DUTY_ADAPTIVE)
        for n in range(self.H):
            for l in range(4):
                tt[l] = np.where(tt[l] >= t(1.0), t(0), tt[l])        # :73-76
                tt[l] = tt[l] + mpc_dt * f                            # :79-82
                seq[l, n] = np.where(tt[l] < duty, t(1.0), t(0.0))   # :84-87
        return seq

    # ---------------------------------------------------------------- rollouts
    def compute_rollout(self, initial_state, reference, control_parameters, contact_sequence):
        """Fixed-gait rollout cost for every row of control_parameters (N,P) -> (N,).
        S/centroidal_nmpc_jax.py:316-496 under vmap (:176)."""
        t = self.t
        cp = np.ascontiguousarray(control_parameters, dtype=t)
        N = cp.shape[0]
        x0 = np.asarray(initial_state, dtype=t)
        ref = np.asarray(reference, dtype=t)
        cs = np.asarray(contact_sequence, dtype=t)
        state = [np.full(N, x0[i], dtype=t) for i in range(24)]
        cost = np.zeros(N, dtype=t)
        legs = [cp[:, l * self.P_leg:(l + 1) * self.P_leg] for l in range(4)]
        with np.errstate(divide='ignore', invalid='ignore', over='ignore'):
            for n in range(self.H):
                steps = [np.int32(n)] * 4                                  # :348-351
                contact_n = [cs[l, n] for l in range(4)]
                F = self._forces_at(legs, steps, [self.H] * 4, contact_n)
                inputs = [t(0)] * 12 + [F[l][k] for l in range(4) for k in range(3)]   # :411-439
                state = self.integrate(state, inputs, contact_n, n)        # :446
                cost = cost + self._error_cost(state, ref)                 # :450-451,491
        return cost

    def _error_cost(self, state, ref):
        """e^T Q e with the dense diagonal Q, accumulated in index order; S/centroidal_nmpc_jax.py:450-451."""
        acc = None
        for j in range(24):
            e = state[j] - ref[j]
            term = (e * self.Q[j, j]) * e
            acc = term if acc is None else acc + term
        return acc

    def compute_rollout_adaptive(self, initial_state, reference, timing, control_parameters, step_frequency):
        """Gait-adaptive rollout; A:326-501 under vmap (A:186)."""
        t = self.t
        cp = np.ascontiguousarray(control_parameters, dtype=t)
        N = cp.shape[0]
        x0 = np.asarray(initial_state, dtype=t)
        ref = np.asarray(reference, dtype=t)
        f = np.asarray(step_frequency, dtype=t)
        cs = self.pgg_contact_sequence(np.asarray(timing, dtype=t), f)      # A:341-343
        hl = [cs[l].sum(axis=0, dtype=t) + t(1) for l in range(4)]           # A:345-348
        state = [np.full(N, x0[i], dtype=t) for i in range(24)]
        cost = np.zeros(N, dtype=t)
        n_ = [np.full(N, -1, dtype=np.int32) for _ in range(4)]              # A:339
        legs = [cp[:, l * self.P_leg:(l + 1) * self.P_leg] for l in range(4)]
        with np.errstate(divide='ignore', invalid='ignore', over='ignore'):
            for n in range(self.H):
                contact_n = [cs[l, n] for l in range(4)]
                n_ = [(n_[l].astype(t) + t(1) * contact_n[l]).astype(np.int32) for l in range(4)]  # A:353-356
                F = self._forces_at(legs, n_, hl, contact_n)
                inputs = [t(0)] * 12 + [F[l][k] for l in range(4) for k in range(3)]
                state = self.integrate(state, inputs, contact_n, n)
                cost = cost + self._error_cost(state, ref)
            cost = cost + (f - t(1.3)) * t(100) * (f - t(1.3))              # A:500
        return cost

    # ---------------------------------------------------------------- noise assembly (row a9)
    def assemble_noise_random_sampling(self, normal_rows, uniform01_rows):
        """Build ``additional_random_parameters`` for random_sampling from raw draws.
        S/centroidal_nmpc_jax.py:647-677.  normal_rows: (k,P) standard normals used for BOTH Gaussian
        blocks (the reference passes the same key to both draws, so both blocks see the same normals);
        uniform01_rows: (N-1-2k, P) in [0,1) mapped to [-s2, s2) as jax.random.uniform does."""
        t, N, P = self.t, self.N, self.P
        k = int(N / 3)
        s = [t(x) for x in self.cfg.sigma_random_sampling]
        noise = np.zeros((N, P), dtype=t)
        noise[1:1 + k] = s[0] * np.asarray(normal_rows, dtype=t)[:k]
        noise[1 + k:1 + 2 * k] = s[1] * np.asarray(normal_rows, dtype=t)[:k]
        u = np.asarray(uniform01_rows, dtype=t)
        noise[1 + 2 * k:N] = np.maximum(-s[2], u * (s[2] - (-s[2])) + (-s[2]))
        return noise

    def assemble_noise_gaussian(self, normal_rows, sigma):
        """mppi (S/centroidal_nmpc_jax.py:806-812: sigma*normal) / cem_mppi (:951-958: normal*sigma)."""
        t = self.t
        noise = np.zeros((self.N, self.P), dtype=t)
        noise[1:] = np.asarray(normal_rows, dtype=t)[:self.N - 1] * np.asarray(sigma, dtype=t)
        return noise

    # ---------------------------------------------------------------- control laws
    def _rollout_any(self, state, reference, contact_sequence, params, timing, step_frequencies):
        if self.adaptive:
            return self.compute_rollout_adaptive(state, reference, timing, params, step_frequencies)
        return self.compute_rollout(state, reference, params, contact_sequence)

    def _saturate(self, costs):
        """NaN -> 1e6, +-Inf -> 1e6 ; S/centroidal_nmpc_jax.py:686-687."""
        t = self.t
        costs = np.where(np.isnan(costs), t(1000000), costs)
        return np.where(np.isinf(costs), t(1000000), costs)

    def _tail(self, state, contact_sequence, best):
        """GRF read-out at step 0 + predicted state; S/centroidal_nmpc_jax.py:695-784 (=:839-928,:985-1073)."""
        t = self.t
        cs = np.asarray(contact_sequence, dtype=t)
        b = np.asarray(best, dtype=t)[None, :]
        legs = [b[:, l * self.P_leg:(l + 1) * self.P_leg] for l in range(4)]
        contact0 = [cs[l, 0] for l in range(4)]
        # spline(best_leg, 0.0, 1) : python-float step, horizon_leg = 1            (:707-710)
        F = self._forces_at(legs, [0.0] * 4, [1] * 4, contact0)
        grfs = np.array([F[l][k][0] for l in range(4) for k in range(3)], dtype=t)  # :747
        x0 = np.asarray(state, dtype=t)
        st = [np.array([x0[i]], dtype=t) for i in range(24)]
        inputs = [t(0)] * 12 + [F[l][k] for l in range(4) for k in range(3)]
        with np.errstate(divide='ignore', invalid='ignore', over='ignore'):
            pred = self.integrate(st, inputs, contact0, 0)                          # :784
        pred = np.array([np.asarray(p).reshape(-1)[0] for p in pred], dtype=t)
        return grfs, pred

    def compute_control(self, state, reference, contact_sequence, best_control_parameters, noise,
                        sigma=None, timing=None, nominal_step_frequency=None, optimize_swing=None,
                        step_frequencies=None):
        """One tick with injected noise (N,P) [and injected per-sample step frequencies (N,) when adaptive].

        Follows compute_control_random_sampling / _mppi / _cem_mppi, S/centroidal_nmpc_jax.py:629-1094 and the
        adaptive twins A:629-1131.  Returns a dict with the reference's tuple members plus the indices."""
        t = self.t
        method = self.cfg.sampling_method
        best_old = np.asarray(best_control_parameters, dtype=t)
        noise = np.asarray(noise, dtype=t)
        params = best_old[None, :] + noise                                  # :680 | :814 | :960
        costs = self._rollout_any(state, reference, contact_sequence, params, timing, step_frequencies)
        costs = self._saturate(costs)
        best_index = int(np.argmin(costs))                                  # :690 (first minimum)
        best_cost = costs[best_index]
        out = {}
        if method == 'random_sampling':
            best = params[best_index].copy()                                # :692
        else:
            with np.errstate(under='ignore'):
                exp_costs = np.exp(t(-1.0 / 1.0) * (costs - best_cost))     # :828-830
            denom = np.sum(exp_costs, dtype=t)                              # :831
            weights = exp_costs / denom                                     # :832
            weighted = weights[:, None] * noise                             # :833-835
            best = best_old + np.sum(weighted, axis=0, dtype=t)             # :836
            out['weights'] = weights
        grfs, pred = self._tail(state, contact_sequence, best)
        out.update(nmpc_GRFs=grfs, nmpc_footholds=np.zeros(12, dtype=t), nmpc_predicted_state=pred,
                   best_control_parameters=best, best_cost=best_cost, costs=costs, best_index=best_index)
        if self.adaptive:
            out['best_freq'] = np.asarray(step_frequencies, dtype=t)[best_index]   # A:705,861,1022
        else:
            out['best_freq'] = t(1.65) if method == 'cem_mppi' else t(1.4)  # :786,930,1083
        if method == 'cem_mppi':
            k = 200 if self.adaptive else 10                                # A:1115 | S:1075
            idx = np.argsort(costs, kind='stable')[:k]
            elite = noise[idx]
            mean = np.sum(elite, axis=0, dtype=t) / t(elite.shape[0])
            X = elite - mean[None, :]
            var = np.sum(X * X, axis=0, dtype=t) / t(elite.shape[0] - 1) + t(1e-8)   # cov diag, ddof=1 (:1077-1078)
            if self.adaptive:                                               # A:1118-1121 (variance, no sqrt)
                s_new = np.where(var > t(3), t(3), var)
                s_new = np.where(s_new < t(0.5), t(0.5), s_new)
            else:                                                           # S:1079-1081
                s_new = np.sqrt(var)
                s_new = np.where(s_new > t(5), t(5), s_new)
                s_new = np.where(s_new < t(0.2), t(0.2), s_new)
            out['new_sigma_cem_mppi'] = s_new.astype(t)
            out['elite_indices'] = idx.astype(np.int64)
        return out

    # ---------------------------------------------------------------- step-frequency assembly (row a15)
    def assemble_step_frequencies(self, choice_idx, nominal_step_frequency, optimize_swing):
        """Per-sample step frequency from injected choice indices in {0,1,2}; A:688-692 | A:835-838 | A:994-1000."""
        t = self.t
        ci = np.asarray(choice_idx, dtype=np.int64)
        method = self.cfg.sampling_method
        if method == 'random_sampling':
            avail = np.where(bool(optimize_swing), self.step_freq_delta, t(nominal_step_frequency)).astype(t)
            return avail[ci]
        if method == 'mppi':
            return self.step_freq_delta[ci]
        avail = np.array([0.0, 0.2, 0.4], dtype=t)
        return (avail[ci] * t(optimize_swing) + t(nominal_step_frequency)).astype(t)

    # ---------------------------------------------------------------- host-side prep (row a16)
    def prepare_state_and_reference(self, state_current, reference_state, current_contact, previous_contact,
                                    best_control_parameters, mpc_frequency=100):
        """S/centroidal_nmpc_jax.py:563-627.  Returns (state24, ref24, best) as float64/np arrays like the
        reference (the jit boundary casts to float32)."""
        best = np.array(best_control_parameters, dtype=self.t)
        if self.cfg.shift_solution:                                        # :571-573
            best = self.shift_solution(best, 1.0 / mpc_frequency)
        s = np.concatenate((state_current['position'], state_current['linear_velocity'],
                            state_current['orientation'], state_current['angular_velocity'],
                            state_current['foot_FL'], state_current['foot_FR'],
                            state_current['foot_RL'], state_current['foot_RR'])).reshape((24,)).astype(np.float64)
        names = ['ref_foot_FL', 'ref_foot_FR', 'ref_foot_RL', 'ref_foot_RR']
        for l in range(4):                                                 # :588-595
            if current_contact[l] == 0.0:
                s[12 + 3 * l:15 + 3 * l] = np.asarray(reference_state[names[l]]).reshape((3,))
        r = np.concatenate((reference_state['ref_position'], reference_state['ref_linear_velocity'],
                            reference_state['ref_orientation'], reference_state['ref_angular_velocity'],
                            *[np.asarray(reference_state[nm]).reshape((3,)) for nm in names])).reshape((24,))
        for l in range(4):                                                 # :612-625
            if previous_contact[l] == 1 and current_contact[l] == 0:
                best[l * self.P_leg:(l + 1) * self.P_leg] = 0.0
        return s, r.astype(np.float64), best

    def shift_solution(self, best, step):
        """S/centroidal_nmpc_jax.py:513-561: writes the spline value at `step` into knots [0],[2],[4] of each leg."""
        best = np.array(best, dtype=self.t)
        for l in range(4):
            leg = best[l * self.P_leg:(l + 1) * self.P_leg].copy()
            fx, fy, fz = self.spline(leg[None, :], step, self.H)
            leg[0], leg[2], leg[4] = fx[0], fy[0], fz[0]
            best[l * self.P_leg:(l + 1) * self.P_leg] = leg
        return best


def twin64(cfg: OracleConfig) -> SamplingOracle:
    """The float64 twin of an oracle configuration (bounds float32 rounding error in tests)."""
    return SamplingOracle(dataclasses.replace(cfg, dtype=np.float64))


# ======================================================================================================================
# Gait look-ahead of the callers: helpers/periodic_gait_generator.py (the numpy generator, float64)
# ======================================================================================================================
def gait_lookahead_contact_table(phase_signal, init, step_freq, duty_factor, phase_offset, horizon, contact_sequence_dts,
                                 contact_sequence_lenghts):
    """`PeriodicGaitGenerator.compute_contact_sequence` (helpers/periodic_gait_generator.py:93-118, non-FULL_STANCE
    branch) with `run()` (:48-76) restated on copies of the generator's state: phase_signal (4,) float64, init (4,) bool.
    Returns the (4, horizon) table; the inputs are not modified (the reference restores them, :117)."""
    t = np.array(phase_signal, dtype=np.float64)
    ini = [bool(x) for x in (init if init is not None else [False] * 4)]
    table = np.zeros((len(t), horizon))

    def advance(dt):                                         # run(), :48-76
        contact = np.zeros(len(t))
        for leg in range(len(t)):
            t[leg] += dt * step_freq                         # :53
            t[leg] = t[leg] % 1.0                            # :56
            if ini[leg]:                                     # :61-67
                contact[leg] = 1
                if not t[leg] <= phase_offset[leg]:
                    ini[leg] = False
                    t[leg] = 0
            else:                                            # :71-74
                contact[leg] = 1 if t[leg] < duty_factor else 0
        return contact

    table[:, 0] = advance(0.0)                               # :106
    j = 0
    for i in range(1, horizon):                              # :110-115
        if i >= contact_sequence_lenghts[j]:
            j += 1
        table[:, i] = advance(contact_sequence_dts[j])
    return table


class FootholdReferenceOracle:
    """TEST INFRASTRUCTURE.  numpy float64 restatement of FootholdReferenceGenerator.compute_footholds_reference
    (helpers/foothold_reference_generator.py:53-157) for one robot; state = the deque of the last
    `vel_moving_average_length` horizontal-frame base velocities (:16,27).  Pinned against the reference class itself by
    tests/golden/gait_footholds.npz (tools/make_golden_footholds.py)."""

    def __init__(self, stance_time, lift_off_z, hip_height, hip_offset=0.1, gravity=9.81, vel_moving_average_length=20):
        self.stance_time, self.hip_height, self.hip_offset, self.gravity = stance_time, hip_height, hip_offset, gravity
        self.lift_off_z = np.asarray(lift_off_z, np.float64).copy()      # (4,) FL FR RL RR
        self.length = vel_moving_average_length
        self.hist = []
        self.com_pos_offset_b = np.zeros(3)                              # :32

    def __call__(self, base_position, base_ori_euler_xyz, base_xy_lin_vel, ref_base_xy_lin_vel, hips_position,
                 com_height_nominal):
        base_position, hips = np.asarray(base_position, np.float64), np.asarray(hips_position, np.float64).reshape(4, 3)
        roll, pitch, yaw = (float(a) for a in base_ori_euler_xyz)
        c, s = np.cos(yaw), np.sin(yaw)
        w2h = np.array([[c, s], [-s, c]])                                # :87-90
        v_h = w2h @ np.asarray(base_xy_lin_vel, np.float64)              # :93
        rv_h = w2h @ np.asarray(ref_base_xy_lin_vel, np.float64)         # :94
        self.hist.append(v_h)                                            # deque(maxlen), :97
        self.hist = self.hist[-self.length:]
        acc = np.zeros(2)
        for v in self.hist:                                              # np.mean(axis=0): row after row, :98
            acc = acc + v
        mvg = acc / len(self.hist)
        lim = self.hip_height * 1.5
        delta = np.minimum(np.maximum((self.stance_time / 2.0) * rv_h, -lim), lim)          # :104-105
        err = np.sqrt(com_height_nominal / self.gravity) * (mvg - rv_h)                     # :109
        err = np.minimum(np.maximum(err, -0.05), 0.05)                                      # :110-111
        cr, sr, cp, sp = np.cos(roll), np.sin(roll), np.cos(pitch), np.sin(pitch)
        b2w = np.array([[c * cp, c * sp * sr - s * cr, c * sp * cr + s * sr],               # extrinsic xyz = Rz Ry Rx, :140
                        [s * cp, s * sp * sr + c * cr, s * sp * cr - c * sr],
                        [-sp, cp * sr, cp * cr]])
        off_w = b2w @ self.com_pos_offset_b                                                 # :141
        out = np.zeros((4, 3))
        for leg in range(4):
            f = w2h @ (hips[leg, :2] - base_position[:2])                                   # :118-121
            f[1] += self.hip_offset if leg in (0, 2) else -self.hip_offset                  # :126-129
            f = f + (delta + err)                                                           # :131
            out[leg, :2] = w2h.T @ f + base_position[:2] + off_w[:2]                        # :134-146
            out[leg, 2] = self.lift_off_z[leg]                                              # :150-151
        return out
