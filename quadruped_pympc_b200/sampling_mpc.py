"""`Sampling_MPC`: the reference's sampling controller surface, computed by libqpmpc_b200 on a B200.

Drop-in for `quadruped_pympc.controllers.sampling.centroidal_nmpc_jax[_gait_adaptive].Sampling_MPC` as used by
`SRBDControllerInterface` (interfaces/srbd_controller_interface.py:77-83,118-168): same attribute names
(`num_sampling_iterations`, `sampling_method`, `best_control_parameters`, `master_key`, `sigma_cem_mppi`), same method
names and argument meaning (`prepare_state_and_reference`, `with_newkey`, `with_newsigma`, `jitted_compute_control`,
`compute_control`, `shift_solution`, `reset`), same return tuples.  Host-side bookkeeping that the reference does in
numpy (S:513-627) stays numpy here; everything the reference jits (S:629-1094, A:629-1131) is one CUDA launch.

Differences a maintainer should know (all documented in DESIGN.md):
  * random bits come from a counter-based Philox4x32-10 keyed by `master_key = (seed, counter)` instead of JAX threefry;
    `with_newkey()` increments the counter.  `noise=` / `step_freqs=` keyword arguments inject the reference's tensors.
  * `costs` is returned as a lazy array (fetched from the GPU only if somebody reads it; the interface does not).
  * no CPU fallback: constructing works anywhere, ticking requires the CUDA library and a device.
"""
from __future__ import annotations

import ctypes as C
import os
import sys
from typing import Optional

import numpy as np

from . import _lib
from .config import SamplerConfig


def _fp(a: Optional[np.ndarray]):
    return a.ctypes.data_as(_lib._FP) if a is not None else None


def _f32(a, shape=None) -> np.ndarray:
    a = np.ascontiguousarray(np.asarray(a), dtype=np.float32)
    return a if shape is None else a.reshape(shape)


def make_c_config(cfg: SamplerConfig, num_robots=1, shard_rank=0, shard_world=1) -> _lib.QpmpcConfig:
    lib = _lib.load()
    c = _lib.QpmpcConfig()
    lib.qpmpc_config_defaults(C.byref(c))
    if cfg.sampling_method not in _lib.METHOD:
        # the reference prints and sys.exit(1)s here (S:107-110)
        print("Error: sampling method not recognized")
        sys.exit(1)
    c.num_samples = int(cfg.num_parallel_computations)
    c.horizon = int(cfg.horizon)
    c.num_splines = int(cfg.num_splines)
    c.parametrization = _lib.PARAMETRIZATION.get(cfg.control_parametrization, 0)   # S:82: anything else = zero order
    c.method = _lib.METHOD[cfg.sampling_method]
    c.adaptive = int(bool(cfg.optimize_step_freq))
    c.num_robots = int(num_robots)
    c.device = int(cfg.device)
    c.io_mode = _lib.IO_MODE[cfg.io_mode]
    c.use_graph = int(bool(cfg.use_graph))
    c.shard_rank, c.shard_world = int(shard_rank), int(shard_world)
    c.elite_k = int(cfg.elite_k)
    c.seed = int(cfg.seed) & 0xFFFFFFFF
    c.mass = float(cfg.mass)
    c.inertia[:] = [float(x) for x in np.asarray(cfg.inertia, dtype=np.float32).ravel()]
    if cfg.horizon > _lib.MAX_HORIZON:
        raise _lib.QpmpcError(_lib.ERR_INVALID, f'horizon {cfg.horizon} > {_lib.MAX_HORIZON}')
    for i, d in enumerate(cfg.dts()):
        c.dts[i] = float(d)
    c.mu, c.grf_min, c.grf_max = float(cfg.mu), float(cfg.grf_min), float(cfg.grf_max)
    c.q_diag[:] = [float(x) for x in cfg.q_diag]
    c.sigma_random_sampling[:] = [float(x) for x in cfg.sigma_random_sampling]
    c.sigma_mppi = float(cfg.sigma_mppi)
    c.step_freq_available[:] = [float(x) for x in cfg.step_freq_available]
    c.pgg_duty_factor = 0.65                                                        # A:179
    c.lanes_per_sample = int(cfg.lanes_per_sample)
    return c


class LazyCosts:
    """`costs` of the last tick; copied from the GPU on first use (np.asarray(costs))."""

    def __init__(self, owner, robot=0):
        self._owner, self._robot, self._tick, self._value = owner, robot, owner._tick_serial, None

    def __array__(self, dtype=None, copy=None):
        if self._value is None:
            if self._owner._tick_serial != self._tick:
                raise RuntimeError('costs of an earlier tick were overwritten by a later tick before being read')
            self._value = self._owner._fetch_costs(self._robot)
        return self._value if dtype is None else self._value.astype(dtype)

    def __len__(self):
        return self._owner._n_local

    @property
    def shape(self):
        return (self._owner._n_local,)


_ZERO_FOOTHOLDS = np.zeros(12, dtype=np.int32)
_ZERO_FOOTHOLDS.setflags(write=False)


class _TickBuffers:
    """ctypes in/out structs bound ONCE to preallocated float32 staging arrays: a tick only copies values into them
    (no allocation, no pointer re-binding on the tick path).  Views onto the fixed-size members of the out struct let
    `_unpack` copy results without going through Python lists."""

    def __init__(self, P, elite_k, H):
        self.cin, self.cout = _lib.QpmpcTickIn(), _lib.QpmpcTickOut()
        self.H = H
        self.state, self.reference = np.zeros(24, np.float32), np.zeros(24, np.float32)
        self.contact = np.zeros((4, H), np.float32)
        self.best_in, self.sigma_in = np.zeros(P, np.float32), np.zeros(P, np.float32)
        self.timing = np.zeros(4, np.float32)
        self.best_out, self.sigma_out = np.zeros(P, np.float32), np.zeros(P, np.float32)
        self.elite = np.full(max(elite_k, 1), -1, np.int32)
        i, o = self.cin, self.cout
        i.state, i.reference, i.contact_sequence, i.contact_ld = _fp(self.state), _fp(self.reference), _fp(self.contact), H
        i.best = _fp(self.best_in)
        self._p_sigma, self._p_timing = _fp(self.sigma_in), _fp(self.timing)
        o.best, o.sigma_new = _fp(self.best_out), _fp(self.sigma_out)
        o.elite_indices = self.elite.ctypes.data_as(C.POINTER(C.c_int32))
        self.key_view = np.ctypeslib.as_array(i.key)           # uint32[2] view of cin.key
        self.grfs_view = np.ctypeslib.as_array(o.grfs)
        self.pred_view = np.ctypeslib.as_array(o.predicted_state)
        self.keep = []

    def fill(self, state, reference, contact_sequence, best, key, sigma, timing, nominal, optimize_swing, noise,
             step_freqs, H):
        cs = np.asarray(contact_sequence)
        if cs.ndim != 2 or cs.shape[0] != 4 or cs.shape[1] < H:
            raise _lib.QpmpcError(_lib.ERR_INVALID, f'contact_sequence must be (4, >= {H}), got {cs.shape}')
        self.contact[...] = cs[:, :H]                       # only [:, :H] is read (FULL_STANCE tables are (4, 2H))
        try:
            self.state[...] = state
            self.reference[...] = reference
        except ValueError:                                  # e.g. (24, 1) columns
            self.state[...] = np.asarray(state).reshape(24)
            self.reference[...] = np.asarray(reference).reshape(24)
        self.best_in[...] = best
        i = self.cin
        if sigma is not None:
            self.sigma_in[...] = sigma
            i.sigma = self._p_sigma
        else:
            i.sigma = None
        if timing is not None:
            self.timing[...] = timing
            i.phase_signal = self._p_timing
        else:
            i.phase_signal = None
        i.nominal_step_freq = float(nominal) if nominal is not None else 0.0
        i.optimize_swing = int(optimize_swing) if optimize_swing is not None else 0
        if isinstance(key, np.ndarray) and key.dtype == np.uint32 and key.shape == (2,):
            self.key_view[...] = key                            # the class's own master_key: one small copy
        else:
            k0, k1 = key.tolist() if isinstance(key, np.ndarray) else key
            i.key[0], i.key[1] = int(k0) & 0xFFFFFFFF, int(k1) & 0xFFFFFFFF
        self.keep = [noise, step_freqs]                     # parity hooks only; already float32 C-contiguous
        i.noise, i.step_freqs = _fp(noise), _fp(step_freqs)


class Sampling_MPC:
    """This is a small class that implements a sampling based control law (on a B200)."""

    def __init__(self, cfg: Optional[SamplerConfig] = None, *, shard_rank: int = 0, shard_world: int = 1):
        if cfg is None:
            import quadruped_pympc.config as config      # the reference's own config module (S:15,27-33)
            cfg = SamplerConfig.from_reference(config)
        self.cfg = cfg
        self.num_parallel_computations = cfg.num_parallel_computations
        self.sampling_method = cfg.sampling_method
        self.control_parametrization = cfg.control_parametrization
        self.num_sampling_iterations = cfg.num_sampling_iterations
        self.dt = cfg.dt
        self.horizon = cfg.horizon
        self.state_dim = 24
        self.control_dim = 24
        self.reference_dim = self.state_dim
        self.max_sampling_forces_x = 10
        self.max_sampling_forces_y = 10
        self.max_sampling_forces_z = 30
        self.num_spline = cfg.num_splines
        self.num_control_parameters_single_leg = cfg.params_per_leg
        self.num_control_parameters = cfg.num_control_parameters
        if cfg.sampling_method not in _lib.METHOD:
            print("Error: sampling method not recognized")
            sys.exit(1)
        if cfg.sampling_method == 'random_sampling':
            self.sigma_random_sampling = cfg.sigma_random_sampling
        elif cfg.sampling_method == 'mppi':
            self.sigma_mppi = cfg.sigma_mppi
        self.sigma_cem_mppi = np.ones(self.num_control_parameters, np.float32) * np.float32(cfg.sigma_cem_mppi)
        self.mu = cfg.mu
        self.f_z_max = cfg.grf_max
        self.f_z_min = cfg.grf_min
        self.best_control_parameters = np.zeros((self.num_control_parameters,), dtype=np.float32)
        self.master_key = np.array([int(cfg.seed) & 0xFFFFFFFF, 0], dtype=np.uint32)
        self.adaptive = bool(cfg.optimize_step_freq)
        self.elite_k = (cfg.elite_k or (200 if self.adaptive else 10)) if cfg.sampling_method == 'cem_mppi' else 0
        self._shard = (int(shard_rank), int(shard_world))
        self._stateful = False
        self._handle = None
        self._pid = None
        self._buf = None
        self._tick_serial = 0
        self._n_local = None
        self._cem = cfg.sampling_method == 'cem_mppi'
        self.jitted_compute_control = self.compute_control

    # ------------------------------------------------------------------ handle management
    def _ensure(self):
        """Create the CUDA handle lazily in the process that ticks (fork safety, ros2/run_controller.py:225-243)."""
        if self._handle is not None and self._pid == os.getpid():
            return self._handle
        lib = _lib.load()
        c = make_c_config(self.cfg, 1, *self._shard)
        h = C.c_void_p()
        _lib.check(lib.qpmpc_create(C.byref(c), C.byref(h)))
        self._handle, self._pid = h, os.getpid()
        self._hptr = int(h.value)
        self._fast = _lib.load_fast() if self.cfg.fast_binding else None   # CPython binding of the tick call (csrc/qpfast.c)
        self._stateful = False                                  # a new handle starts without a device record
        self._shift_set = 0.0
        self._n_local = lib.qpmpc_num_local_samples(h)
        self._buf = _TickBuffers(self.num_control_parameters, self.elite_k, self.horizon)
        return h

    @property
    def kernel_name(self):
        """tick kernel variant this controller launches (tick_kernel | tick_kernel_h | tick_kernel_quad)"""
        return _lib.load().qpmpc_kernel_name(self._ensure()).decode()

    def close(self):
        if self._handle is not None and self._pid == os.getpid():
            _lib.load().qpmpc_destroy(self._handle)
        self._handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _fetch_costs(self, robot=0):
        out = np.empty(self._n_local, np.float32)
        _lib.check(_lib.load().qpmpc_get_costs(self._ensure(), robot, _fp(out)))
        return out

    # ------------------------------------------------------------------ sample-sharded controller
    @staticmethod
    def nccl_unique_id() -> bytes:
        """128-byte ncclUniqueId made on one rank; broadcast it (e.g. `distributed.broadcast_bytes`) to the others."""
        buf = (C.c_ubyte * 128)()
        _lib.check(_lib.load().qpmpc_nccl_unique_id(_lib.find_libnccl().encode(), buf))
        return bytes(buf)

    def init_nccl(self, unique_id: bytes):
        """Join the NCCL communicator of a sample-sharded controller (one rank per GPU, `shard_world` ranks)."""
        if len(unique_id) != 128:
            raise _lib.QpmpcError(_lib.ERR_INVALID, 'ncclUniqueId must be 128 bytes')
        buf = (C.c_ubyte * 128).from_buffer_copy(unique_id)
        _lib.check(_lib.load().qpmpc_shard_init_nccl(self._ensure(), _lib.find_libnccl().encode(), buf))
        return self

    def peer_export(self) -> bytes:
        """128-byte description of this shard's mailbox (CUDA IPC handle); all-gather the blobs of every rank (e.g.
        `distributed.all_gather_bytes`) and hand them to `peer_attach` in rank order."""
        buf = (C.c_ubyte * _lib.PEER_BLOB_BYTES)()
        _lib.check(_lib.load().qpmpc_shard_peer_export(self._ensure(), buf))
        return bytes(buf)

    def peer_attach(self, blobs):
        """Open every rank's mailbox: from now on a sharded tick is ONE launch per rank with the record exchange fused
        into the kernel (stores over NVLink peer memory) instead of partial + ncclAllGather + finish."""
        blobs = list(blobs)
        if len(blobs) != self._shard[1] or any(len(b) != _lib.PEER_BLOB_BYTES for b in blobs):
            raise _lib.QpmpcError(_lib.ERR_INVALID, f'need {self._shard[1]} blobs of {_lib.PEER_BLOB_BYTES} bytes')
        buf = (C.c_ubyte * (len(blobs) * _lib.PEER_BLOB_BYTES)).from_buffer_copy(b''.join(blobs))
        _lib.check(_lib.load().qpmpc_shard_peer_attach(self._ensure(), buf))
        return self

    def peer_loopback(self):
        """Test hook: run the fused exchange on one GPU; absent peers are a sink, their records come from `peer_inject`."""
        _lib.check(_lib.load().qpmpc_shard_peer_loopback(self._ensure()))
        return self

    def peer_inject(self, rank: int, record):
        """Test hook: deliver peer `rank`'s shard record (and the stamp of the upcoming tick) before ticking."""
        _lib.check(_lib.load().qpmpc_shard_peer_inject(self._ensure(), int(rank), _fp(_f32(record))))
        return self

    def peer_timeout(self, seconds: float):
        """How long the in-kernel exchange waits for the peers before a sharded tick fails with ERR_NCCL (default 2 s)."""
        _lib.check(_lib.load().qpmpc_shard_peer_timeout(self._ensure(), float(seconds)))
        return self

    def debug_peer_wait(self, enable=True):
        """(start, end) of the last peer exchange on this GPU's %globaltimer (ns); `enable` arms the recording for the
        following sharded ticks (costs one extra synchronous copy per tick -- debugging only)."""
        out = (C.c_uint64 * 2)()
        _lib.check(_lib.load().qpmpc_debug_peer_wait(self._ensure(), int(bool(enable)), out))
        return int(out[0]), int(out[1])

    def shard_partial(self, state, reference, contact_sequence, best, key, sigma=None, timing=None, nominal=None,
                      optimize_swing=None, noise=None, step_freqs=None) -> np.ndarray:
        """Rollouts + local reduction of this shard's slice; returns the shard record (floats) for a caller-side gather
        (tests, or transports other than NCCL)."""
        h, lib, b = self._ensure(), _lib.load(), self._buf
        b.fill(state, reference, contact_sequence, best, key, sigma, timing, nominal, optimize_swing,
               None if noise is None else _f32(noise), None if step_freqs is None else _f32(step_freqs), self.horizon)
        self._tick_serial += 1
        _lib.check(lib.qpmpc_shard_partial(h, C.byref(b.cin)))
        rec = np.empty(lib.qpmpc_shard_record_floats(h), np.float32)
        _lib.check(lib.qpmpc_shard_get_record(h, _fp(rec)))
        return rec

    def shard_finish(self, records):
        """Merge the `shard_world` records (in rank order) and finish the tick; returns the usual tuple."""
        h, lib, b = self._ensure(), _lib.load(), self._buf
        if len(records) != self._shard[1]:
            raise _lib.QpmpcError(_lib.ERR_INVALID, f'need {self._shard[1]} shard records, got {len(records)}')
        for slot, rec in enumerate(records):
            _lib.check(lib.qpmpc_shard_copy_record(h, slot, _fp(_f32(rec))))
        _lib.check(lib.qpmpc_shard_finish(h, C.byref(b.cout)))
        return self._unpack(b)

    # ------------------------------------------------------------------ reference surface (host side)
    def with_newkey(self):
        """S:498-501 (jax.random.split) -> advance the Philox counter."""
        self.master_key = np.array([self.master_key[0], (int(self.master_key[1]) + 1) & 0xFFFFFFFF], dtype=np.uint32)
        return self

    def get_key(self):
        return self.master_key

    def with_newsigma(self, sigma):
        """S:506-508"""
        self.sigma_cem_mppi = sigma
        return self

    def get_sigma(self):
        return self.sigma_cem_mppi

    def reset(self):
        """S:1096-1097"""
        print("Resetting the controller")

    def _spline_host(self, leg_params, step, horizon_leg):
        """Host evaluation of the three spline functions (S:181-268) for `shift_solution`; float32 like jnp on f32."""
        t = np.float32
        p = np.asarray(leg_params, dtype=t)
        H, ns = self.horizon, self.num_spline
        if self.control_parametrization not in ('linear_spline', 'cubic_spline'):
            i = int(np.int16(step))
            return p[i], p[i + H], p[i + 2 * H]
        bounds = np.linspace(0, H, ns + 1).astype(t)
        index = int(np.max(np.where(t(step) >= bounds, np.arange(ns + 1), 0)))
        q = t(step) / t(horizon_leg / ns) - t(index)
        if self.control_parametrization == 'linear_spline':
            sh = ns + 1
            return tuple((t(1) - q) * p[index + sh * a] + q * p[index + sh * a + 1] for a in range(3))
        s = 10 * index
        a = t(2) * q * q * q - t(3) * q * q + t(1)
        b = q * q * q - t(2) * q * q + q
        c = t(-2) * q * q * q + t(3) * q * q
        d = q * q * q - q * q
        out = []
        for o in (0, 4, 8):
            p0, p1, p2, p3 = p[s + o], p[s + o + 1], p[s + o + 2], p[s + o + 3]
            phi = t(0.5) * ((p2 - p1) + (p1 - p0))
            phin = t(0.5) * ((p3 - p2) + (p2 - p1))
            out.append(a * p1 + b * phi + c * p2 + d * phin)
        return tuple(out)

    def shift_solution(self, best_control_parameters, step):
        """S:513-561: writes the spline value at `step` into knots [0], [2], [4] of every leg (sic)."""
        best = np.array(best_control_parameters, dtype=np.float32)
        L = self.num_control_parameters_single_leg
        for leg in range(4):
            ctrl = best[leg * L:(leg + 1) * L].copy()
            fx, fy, fz = self._spline_host(ctrl.copy(), step, self.horizon)
            ctrl[0], ctrl[2], ctrl[4] = fx, fy, fz
            best[leg * L:(leg + 1) * L] = ctrl
        return best

    def prepare_state_and_reference(self, state_current, reference_state, current_contact, previous_contact,
                                    mpc_frequency=100):
        """S:563-627: dicts -> 24-vectors; swing feet replaced by the reference footholds; knots of a leg that just
        lifted off are zeroed; optional shift of the previous solution."""
        if self.cfg.shift_solution:
            index_shift = 1.0 / mpc_frequency
            self.best_control_parameters = self.shift_solution(self.best_control_parameters, index_shift)
        state_current_jax = np.concatenate((
            state_current["position"], state_current["linear_velocity"], state_current["orientation"],
            state_current["angular_velocity"], state_current["foot_FL"], state_current["foot_FR"],
            state_current["foot_RL"], state_current["foot_RR"])).reshape((24,))
        names = ("ref_foot_FL", "ref_foot_FR", "ref_foot_RL", "ref_foot_RR")
        for leg in range(4):
            if current_contact[leg] == 0.0:
                state_current_jax[12 + 3 * leg:15 + 3 * leg] = np.asarray(reference_state[names[leg]]).reshape((3,))
        reference_state_jax = np.concatenate((
            reference_state["ref_position"], reference_state["ref_linear_velocity"],
            reference_state["ref_orientation"], reference_state["ref_angular_velocity"],
            *[np.asarray(reference_state[n]).reshape((3,)) for n in names])).reshape((24,))
        self.best_control_parameters = np.array(self.best_control_parameters)
        L = self.num_control_parameters_single_leg
        for leg in range(4):
            if previous_contact[leg] == 1 and current_contact[leg] == 0:
                self.best_control_parameters[leg * L:(leg + 1) * L] = 0.0
        return state_current_jax, reference_state_jax

    # ------------------------------------------------------------------ the tick
    def compute_control(self, state, reference, contact_sequence, best_control_parameters, key, *args,
                        noise=None, step_freqs=None, iterations=1, **kwargs):
        """One MPC tick = `jitted_compute_control` of the reference.

        random_sampling / mppi (S:629-932, A:629-962):
            (state, reference, contact_sequence, best, key, timing, nominal_step_frequency, optimize_swing)
            -> (nmpc_GRFs, nmpc_footholds, nmpc_predicted_state, best, best_cost, best_freq, costs)
        cem_mppi (S:934-1094; A:964-1131 additionally takes timing, nominal_step_frequency, optimize_swing):
            (state, reference, contact_sequence, best, key, sigma, [timing, nominal_step_frequency, [optimize_swing]])
            -> (nmpc_GRFs, nmpc_footholds, nmpc_predicted_state, best, best_cost, best_freq, costs, new_sigma)
        """
        cem = self._cem
        if not kwargs and len(args) == (4 if cem else 3):        # the call the interface makes (I:131-168): fast path
            if cem:
                sigma, timing, nominal, optimize_swing = args
            else:
                sigma = None
                timing, nominal, optimize_swing = args
        else:
            names = ['sigma', 'timing', 'nominal_step_frequency', 'optimize_swing'] if cem else \
                    ['timing', 'nominal_step_frequency', 'optimize_swing']
            vals = dict(zip(names, args))
            if len(args) > len(names):
                raise TypeError(f'compute_control() takes at most {5 + len(names)} positional arguments')
            for k, v in kwargs.items():
                if k not in names:
                    raise TypeError(f"compute_control() got an unexpected keyword argument '{k}'")
                vals[k] = v
            sigma = vals.get('sigma')
            timing, nominal, optimize_swing = vals.get('timing'), vals.get('nominal_step_frequency'), vals.get('optimize_swing')
        if cem and sigma is None:
            raise TypeError("compute_control_cem_mppi() missing required argument 'sigma'")
        if self.adaptive and (timing is None or nominal is None or (optimize_swing is None and step_freqs is None)):
            raise TypeError('the gait-adaptive controller needs timing, nominal_step_frequency and optimize_swing '
                            '(A:326,637-638,972-974)')
        out = self._tick(state, reference, contact_sequence, best_control_parameters, key, sigma, timing, nominal,
                         optimize_swing, noise, step_freqs, iterations)
        return out

    def compute_control_iterations(self, state, reference, contact_sequence, best_control_parameters, key, *args,
                                   iterations=None, **kwargs):
        """The sampling-iteration loop of one MPC tick (I:125-168) in ONE call, state carried on the device:

            for _ in range(num_sampling_iterations):
                controller = controller.with_newkey()
                ... = controller.jitted_compute_control(state, ref, contact_sequence, controller.best_control_parameters,
                                                        controller.master_key, [controller.sigma_cem_mppi,] ...)
                controller.best_control_parameters = best ; controller = controller.with_newsigma(new_sigma)

        `key` is the key of the FIRST iteration (call `with_newkey()` once before, as the loop does); iteration i draws
        with `key[1] + i`.  Returns the last iteration's tuple, bit-identical to the loop; `master_key` is advanced by
        `iterations - 1` so that the next `with_newkey()` continues the sequence."""
        n = self.num_sampling_iterations if iterations is None else int(iterations)
        out = self.compute_control(state, reference, contact_sequence, best_control_parameters, key, *args,
                                   iterations=n, **kwargs)
        k = np.asarray(key).astype(np.uint32).ravel()
        self.master_key = np.array([k[0], (int(k[1]) + n - 1) & 0xFFFFFFFF], dtype=np.uint32)
        return out

    # ------------------------------------------------------------------ stateful controller
    _STATE_KEYS = ("position", "linear_velocity", "orientation", "angular_velocity", "foot_FL", "foot_FR", "foot_RL", "foot_RR")
    _REF_KEYS = ("ref_position", "ref_linear_velocity", "ref_orientation", "ref_angular_velocity",
                 "ref_foot_FL", "ref_foot_FR", "ref_foot_RL", "ref_foot_RR")

    def step(self, state_current, ref_state, contact_sequence, pgg_phase_signal=None, pgg_step_freq=None,
             optimize_swing=0, iterations=None, mpc_frequency=100):
        """The sampling block of SRBDControllerInterface.compute_control (I:118-168) in ONE call:

            prepare_state_and_reference(...); previous_contact_mpc = current_contact
            for _ in range(num_sampling_iterations): with_newkey(); [with_newsigma(cfg sigma) at 0]; jitted_compute_control(...)

        with best_control_parameters, sigma, the key counter and previous_contact_mpc kept in a device record between
        calls (shift_solution, the foothold substitution and the lift-off knot zeroing happen in the kernel's staging).  `state_current`
        / `ref_state` are the interface's dicts (or raw 24-vectors).  Returns the last iteration's tuple; the host
        mirrors (`best_control_parameters`, `master_key`, `sigma_cem_mppi`) are kept in step."""
        h, lib, b = self._ensure(), _lib.load(), self._buf
        shift = 1.0 / mpc_frequency if self.cfg.shift_solution else 0.0       # S:571-573, done in the kernel's staging
        if shift != getattr(self, '_shift_set', 0.0):
            _lib.check(lib.qpmpc_step_set_shift(h, shift))
            self._shift_set = shift
        n = self.num_sampling_iterations if iterations is None else int(iterations)
        if isinstance(state_current, dict):
            st = np.concatenate([np.asarray(state_current[k], np.float64).reshape(-1) for k in self._STATE_KEYS])
        else:
            st = np.asarray(state_current).reshape(24)
        if isinstance(ref_state, dict):
            rf = np.concatenate([np.asarray(ref_state[k], np.float64).reshape(-1)[:3] for k in self._REF_KEYS])
        else:
            rf = np.asarray(ref_state).reshape(24)
        if not self._stateful:
            key = (C.c_uint32 * 2)(int(self.master_key[0]), (int(self.master_key[1]) + 1) & 0xFFFFFFFF)
            best = _f32(self.best_control_parameters, (self.num_control_parameters,))
            _lib.check(lib.qpmpc_step_reset(h, key, _fp(best), None))
            self._stateful = True
        self._tick_serial += 1
        if self._fast is not None and contact_sequence is not None:
            out = self._fast_result(self._fast.tick(self._hptr, 3, st, rf, contact_sequence, None, 0, 0, None, pgg_phase_signal,
                                                    pgg_step_freq, optimize_swing, n, float(self.cfg.sigma_cem_mppi),
                                                    self.num_control_parameters, self.horizon, self.elite_k))
        else:
            b.fill(st, rf, contact_sequence, 0.0, (0, 0), None, pgg_phase_signal, pgg_step_freq, optimize_swing, None, None,
                   self.horizon)
            _lib.check(lib.qpmpc_step(h, C.byref(b.cin), n, float(self.cfg.sigma_cem_mppi), C.byref(b.cout)))
            out = self._unpack(b)
        self.best_control_parameters = out[3]
        self.master_key = np.array([self.master_key[0], (int(self.master_key[1]) + n) & 0xFFFFFFFF], dtype=np.uint32)
        if self._cem:
            self.sigma_cem_mppi = out[7]
        return out

    def step_reset(self, previous_contact=(1, 1, 1, 1)):
        """Restart the stateful controller from the host mirrors (`best_control_parameters`, `master_key`)."""
        h = self._ensure()
        key = (C.c_uint32 * 2)(int(self.master_key[0]), (int(self.master_key[1]) + 1) & 0xFFFFFFFF)
        pc = (C.c_int32 * 4)(*[int(x) for x in previous_contact])
        best = _f32(self.best_control_parameters, (self.num_control_parameters,))
        _lib.check(_lib.load().qpmpc_step_reset(h, key, _fp(best), pc))
        self._stateful = True
        return self

    def step_state(self):
        """(best, sigma, previous_contact, key of the next iteration) as stored on the device."""
        h = self._ensure()
        P = self.num_control_parameters
        best, sigma = np.zeros(P, np.float32), np.zeros(P, np.float32)
        pc, key = (C.c_int32 * 4)(), (C.c_uint32 * 2)()
        _lib.check(_lib.load().qpmpc_step_state(h, 0, _fp(best), _fp(sigma), pc, key))
        return best, sigma, np.array(pc[:]), np.array(key[:], np.uint32)

    def stance_torques(self, feet_jac, current_contact):
        """Epilogue on the device-resident GRFs of the last tick: nmpc_GRFs * current_contact
        (I:225-230) and tau = -J^T nmpc_GRFs per leg (wb_interface.py:372-375).  feet_jac (4,3,3) =
        feet_jac.leg[:, legs_qvel_idx.leg]; current_contact (4,) -> (grfs (12,), tau (12,))."""
        jac, con = _f32(feet_jac, (4, 3, 3)), _f32(current_contact, (4,))
        grf, tau = np.empty(12, np.float32), np.empty(12, np.float32)
        _lib.check(_lib.load().qpmpc_stance_torques(self._ensure(), _fp(jac), _fp(con), _fp(grf), _fp(tau)))
        return grf, tau

    def _tick(self, state, reference, contact_sequence, best, key, sigma, timing, nominal, optimize_swing, noise,
              step_freqs, iterations=1):
        h = self._ensure()
        if noise is None and step_freqs is None and self._fast is not None:
            # the real-time path: csrc/qpfast.c reads the caller's arrays in place and calls the same C-ABI entry point
            k0, k1 = key.tolist() if isinstance(key, np.ndarray) else key
            mode = 2 if iterations != 1 else (0 if self._shard[1] == 1 else 1)
            res = self._fast.tick(self._hptr, mode, state, reference, contact_sequence, best, k0, k1, sigma, timing, nominal,
                                  optimize_swing, iterations, 0.0, self.num_control_parameters, self.horizon, self.elite_k)
            self._tick_serial += 1
            return self._fast_result(res)
        lib = _lib.load()
        b = self._buf
        if noise is not None:
            noise = _f32(noise)
            if noise.shape != (self.num_parallel_computations, self.num_control_parameters):
                raise _lib.QpmpcError(_lib.ERR_INVALID, f'noise must be (N, P), got {noise.shape}')
        if step_freqs is not None:
            step_freqs = _f32(step_freqs)
            if step_freqs.shape != (self.num_parallel_computations,):
                raise _lib.QpmpcError(_lib.ERR_INVALID, f'step_freqs must be (N,), got {step_freqs.shape}')
        b.fill(state, reference, contact_sequence, best, key, sigma, timing, nominal, optimize_swing, noise, step_freqs,
               self.horizon)
        self._tick_serial += 1
        if iterations != 1:
            _lib.check(lib.qpmpc_tick_iterations(h, C.byref(b.cin), int(iterations), C.byref(b.cout)))
        elif self._shard[1] == 1:
            _lib.check(lib.qpmpc_tick(h, C.byref(b.cin), C.byref(b.cout)))
        else:
            _lib.check(lib.qpmpc_tick_sharded(h, C.byref(b.cin), C.byref(b.cout)))
        return self._unpack(b)

    def _fast_result(self, res):
        if type(res) is int:
            _lib.check(res)
        grfs, pred, best, best_cost, best_freq, self.last_best_index, sigma_new, elite = res
        out = (grfs, _ZERO_FOOTHOLDS, pred, best, np.float32(best_cost), best_freq, LazyCosts(self))
        if self._cem:
            self.last_elite_indices = elite.astype(np.int64)
            out = out + (sigma_new,)
        return out

    def _unpack(self, b):
        o = b.cout
        grfs = b.grfs_view.copy()
        pred = b.pred_view.copy()
        footholds = _ZERO_FOOTHOLDS                                 # S:748 (read-only zeros; the interface ignores them, I:170-175)
        best = b.best_out.copy()
        self.last_best_index = o.best_index
        # best_freq: a c_float field reads back as the exact float32 value already
        res = (grfs, footholds, pred, best, np.float32(o.best_cost), o.best_freq, LazyCosts(self))
        if self.sampling_method == 'cem_mppi':
            self.last_elite_indices = b.elite[:o.num_elite].astype(np.int64)
            res = res + (b.sigma_out.copy(),)
        return res

    # ------------------------------------------------------------------ the RNG-free seam (S:176-177 / A:186-187)
    def jit_vectorized_rollout(self, *args):
        """fixed gait: (state, reference, control_parameters_vec (N,P), contact_sequence) -> costs (N,)
        adaptive:   (state, reference, timing, control_parameters_vec (N,P), step_frequencies_vec (N,)) -> costs"""
        h = self._ensure()
        lib = _lib.load()
        N, P, H = self.num_parallel_computations, self.num_control_parameters, self.horizon
        if self.adaptive:
            state, reference, timing, params, freqs = args
            cs = np.ones((4, H), np.float32)        # unused by the adaptive rollout (contacts come from the phase clock)
            tm, fq = _f32(timing, (4,)), _f32(freqs, (N,))
        else:
            state, reference, params, contact_sequence = args
            cs, tm, fq = _f32(contact_sequence), None, None
        params = _f32(params, (N, P))
        st, rf = _f32(state, (24,)), _f32(reference, (24,))
        costs = np.empty(N, np.float32)
        _lib.check(lib.qpmpc_rollout(h, _fp(st), _fp(rf), _fp(cs), cs.shape[1], _fp(tm), _fp(params), _fp(fq), _fp(costs)))
        return costs

    vectorized_rollout = jit_vectorized_rollout

    # ------------------------------------------------------------------ RNG export (parity / debugging)
    def generate_noise(self, key=None, sigma=None) -> np.ndarray:
        """additional_random_parameters (N,P) the tick with `key` draws (S:647-677 | 806-812 | 951-958)."""
        h = self._ensure()
        key = np.asarray(self.master_key if key is None else key).astype(np.uint32)
        out = np.empty((self.num_parallel_computations, self.num_control_parameters), np.float32)
        sg = None
        if self.sampling_method == 'cem_mppi':
            sg = np.broadcast_to(np.asarray(self.sigma_cem_mppi if sigma is None else sigma, np.float32),
                                 (self.num_control_parameters,)).copy()
        _lib.check(_lib.load().qpmpc_generate_noise(h, 0, key.ctypes.data_as(C.POINTER(C.c_uint32)), _fp(sg), _fp(out)))
        return out

    def generate_step_freqs(self, key, nominal_step_frequency, optimize_swing) -> np.ndarray:
        h = self._ensure()
        key = np.asarray(key).astype(np.uint32)
        out = np.empty(self.num_parallel_computations, np.float32)
        _lib.check(_lib.load().qpmpc_generate_step_freqs(h, 0, key.ctypes.data_as(C.POINTER(C.c_uint32)),
                                                         float(nominal_step_frequency), int(optimize_swing), _fp(out)))
        return out
