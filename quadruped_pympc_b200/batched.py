"""`BatchedSampling_MPC`: R independent sampling controllers ticked by ONE CUDA launch (grid = blocks x robots).

The reference runs one OS process per robot, each with its own JAX runtime and controller
(simulation/batched_simulations.py:22-58); there is no device-side batching.  Here the R robots of one GPU
share a handle (`num_robots = R`), and robots are partitioned over GPUs/ranks with no communication
(`robot_base` keeps every robot's Philox stream independent of the partitioning).

Per robot the arithmetic is exactly `Sampling_MPC.compute_control` (same kernel, blockIdx.y = robot).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import numpy as np

from . import _lib
from .config import SamplerConfig
from .sampling_mpc import make_c_config, _f32, _fp

_IP = C.POINTER(C.c_int32)
_UP = C.POINTER(C.c_uint32)


def partition_robots(num_robots: int, rank: int, world: int):
    """Contiguous slice [begin, end) of the robots owned by `rank` (robot r -> rank r // ceil(R / world))."""
    per = (num_robots + world - 1) // world
    begin = min(per * rank, num_robots)
    return begin, min(begin + per, num_robots)


class BatchedSampling_MPC:
    """R controllers of one robot type; arrays carry a leading robot axis."""

    def __init__(self, cfg: SamplerConfig, num_robots: int, robot_base: int = 0):
        if cfg.sampling_method not in _lib.METHOD:
            raise _lib.QpmpcError(_lib.ERR_INVALID, 'Error: sampling method not recognized')
        self.cfg = cfg
        self.num_robots = int(num_robots)
        self.robot_base = int(robot_base)
        self.sampling_method = cfg.sampling_method
        self.num_parallel_computations = cfg.num_parallel_computations
        self.horizon = cfg.horizon
        self.num_control_parameters = cfg.num_control_parameters
        self.adaptive = bool(cfg.optimize_step_freq)
        R, P = self.num_robots, self.num_control_parameters
        self.best_control_parameters = np.zeros((R, P), np.float32)
        self.sigma_cem_mppi = np.full((R, P), np.float32(cfg.sigma_cem_mppi), np.float32)
        self.master_key = np.array([int(cfg.seed) & 0xFFFFFFFF, 0], dtype=np.uint32)
        self._handle, self._pid = None, None
        # output arrays are allocated once and reused (no allocation on the tick path)
        self._grfs = np.zeros((R, 12), np.float32)
        self._pred = np.zeros((R, 24), np.float32)
        self._best = np.zeros((R, P), np.float32)
        self._cost = np.zeros(R, np.float32)
        self._freq = np.zeros(R, np.float32)
        self._idx = np.zeros(R, np.int32)
        self._sig = np.zeros((R, P), np.float32)

    def _ensure(self):
        if self._handle is not None and self._pid == os.getpid():
            return self._handle
        lib = _lib.load()
        c = make_c_config(self.cfg, self.num_robots)
        c.robot_base = self.robot_base
        h = C.c_void_p()
        _lib.check(lib.qpmpc_create(C.byref(c), C.byref(h)))
        self._handle, self._pid = h, os.getpid()
        self._stateful = False
        self._shift_set = 0.0
        return h

    def close(self):
        if self._handle is not None and self._pid == os.getpid():
            _lib.load().qpmpc_destroy(self._handle)
        self._handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def with_newkey(self):
        self.master_key = np.array([self.master_key[0], (int(self.master_key[1]) + 1) & 0xFFFFFFFF], dtype=np.uint32)
        return self

    def compute_control(self, states, references, contact_sequences, best_control_parameters=None, key=None,
                        sigma=None, timing=None, nominal_step_frequency=None, optimize_swing=None):
        """states (R,24), references (R,24), contact_sequences (R,4,>=H), best (R,P) ->
        (nmpc_GRFs (R,12), nmpc_predicted_state (R,24), best (R,P), best_cost (R,), best_freq (R,), best_index (R,)
         [, new_sigma (R,P) for cem_mppi])"""
        h = self._ensure()
        R, P, H = self.num_robots, self.num_control_parameters, self.horizon
        st, rf = _f32(states, (R, 24)), _f32(references, (R, 24))
        cs = _f32(contact_sequences)
        if cs.ndim != 3 or cs.shape[0] != R or cs.shape[1] != 4 or cs.shape[2] < H:
            raise _lib.QpmpcError(_lib.ERR_INVALID, f'contact_sequences must be ({R}, 4, >= {H}), got {cs.shape}')
        bs = _f32(self.best_control_parameters if best_control_parameters is None else best_control_parameters, (R, P))
        cem = self.sampling_method == 'cem_mppi'
        sg = None
        if cem:
            sg = np.ascontiguousarray(np.broadcast_to(
                np.asarray(self.sigma_cem_mppi if sigma is None else sigma, np.float32), (R, P)))
        tm = _f32(timing, (R, 4)) if timing is not None else None
        if self.adaptive and tm is None:
            raise TypeError('the gait-adaptive controller needs timing (R,4)')
        nf = None if nominal_step_frequency is None else \
            np.ascontiguousarray(np.broadcast_to(np.asarray(nominal_step_frequency, np.float32), (R,)))
        osw = None if optimize_swing is None else \
            np.ascontiguousarray(np.broadcast_to(np.asarray(optimize_swing, np.int32), (R,)))
        k = np.asarray(self.master_key if key is None else key).astype(np.uint32).ravel()
        _lib.check(_lib.load().qpmpc_tick_batched_flat(
            h, _fp(st), _fp(rf), _fp(cs), cs.shape[2], _fp(bs), _fp(sg), _fp(tm), _fp(nf),
            osw.ctypes.data_as(_IP) if osw is not None else None, k.ctypes.data_as(_UP),
            _fp(self._grfs), _fp(self._pred), _fp(self._best), _fp(self._cost), _fp(self._freq),
            self._idx.ctypes.data_as(_IP), _fp(self._sig) if cem else None))
        res = (self._grfs.copy(), self._pred.copy(), self._best.copy(), self._cost.copy(), self._freq.copy(),
               self._idx.copy())
        return res + (self._sig.copy(),) if cem else res

    # ------------------------------------------------------------------ stateful batch
    def step(self, states, references, contact_sequences, timing=None, nominal_step_frequency=None, optimize_swing=None,
             iterations=1, mpc_frequency=100):
        """The sampling block of SRBDControllerInterface.compute_control (I:118-168) for all R robots in one call, with
        every robot's best / sigma / key counter / previous contact kept on the device between calls: raw states in
        (shift_solution when configured, foothold substitution and lift-off zeroing happen in the kernel), `iterations`
        sampling iterations.  states (R,24) RAW, references (R,24), contact_sequences (R,4,>=H) -> like compute_control."""
        h = self._ensure()
        lib = _lib.load()
        shift = 1.0 / mpc_frequency if self.cfg.shift_solution else 0.0       # S:571-573
        if shift != self._shift_set:
            _lib.check(lib.qpmpc_step_set_shift(h, shift))
            self._shift_set = shift
        R, P, H = self.num_robots, self.num_control_parameters, self.horizon
        if not getattr(self, '_stateful', False):
            key = (C.c_uint32 * 2)(int(self.master_key[0]), (int(self.master_key[1]) + 1) & 0xFFFFFFFF)
            _lib.check(lib.qpmpc_step_reset(h, key, _fp(_f32(self.best_control_parameters, (R, P))), None))
            self._stateful = True
        st, rf = _f32(states, (R, 24)), _f32(references, (R, 24))
        cs, ld = None, H                   # None: the tables made by gait.contact_sequences_from_phase are on the device
        if contact_sequences is not None:
            cs = _f32(contact_sequences)
            if cs.ndim != 3 or cs.shape[0] != R or cs.shape[1] != 4 or cs.shape[2] < H:
                raise _lib.QpmpcError(_lib.ERR_INVALID, f'contact_sequences must be ({R}, 4, >= {H}), got {cs.shape}')
            ld = cs.shape[2]
        tm = _f32(timing, (R, 4)) if timing is not None else None
        if self.adaptive and tm is None:
            raise TypeError('the gait-adaptive controller needs timing (R,4)')
        nf = None if nominal_step_frequency is None else \
            np.ascontiguousarray(np.broadcast_to(np.asarray(nominal_step_frequency, np.float32), (R,)))
        osw = None if optimize_swing is None else \
            np.ascontiguousarray(np.broadcast_to(np.asarray(optimize_swing, np.int32), (R,)))
        cem = self.sampling_method == 'cem_mppi'
        _lib.check(lib.qpmpc_step_flat(
            h, _fp(st), _fp(rf), _fp(cs), ld, _fp(tm), _fp(nf),
            osw.ctypes.data_as(_IP) if osw is not None else None, int(iterations), float(self.cfg.sigma_cem_mppi),
            _fp(self._grfs), _fp(self._pred), _fp(self._best), _fp(self._cost), _fp(self._freq),
            self._idx.ctypes.data_as(_IP), _fp(self._sig) if cem else None))
        self.best_control_parameters = self._best.copy()
        self.master_key = np.array([self.master_key[0], (int(self.master_key[1]) + int(iterations)) & 0xFFFFFFFF], dtype=np.uint32)
        res = (self._grfs.copy(), self._pred.copy(), self._best.copy(), self._cost.copy(), self._freq.copy(),
               self._idx.copy())
        if cem:
            self.sigma_cem_mppi = self._sig.copy()
            res = res + (self._sig.copy(),)
        return res

    def stance_torques(self, feet_jac, current_contact):
        """Batched epilogue on the device-resident GRFs of the last tick:
        nmpc_GRFs * current_contact (I:225-230) and tau = -J^T nmpc_GRFs per leg (wb_interface.py:372-375).
        feet_jac (R,4,3,3) = feet_jac.leg[:, legs_qvel_idx.leg]; current_contact (R,4) -> (grfs (R,12), tau (R,12))."""
        R = self.num_robots
        jac, con = _f32(feet_jac, (R, 4, 3, 3)), _f32(current_contact, (R, 4))
        grf, tau = np.empty((R, 12), np.float32), np.empty((R, 12), np.float32)
        _lib.check(_lib.load().qpmpc_stance_torques(self._ensure(), _fp(jac), _fp(con), _fp(grf), _fp(tau)))
        return grf, tau

    def generate_noise(self, robot: int, key=None, sigma=None) -> np.ndarray:
        """additional_random_parameters (N,P) robot `robot` of this batch draws for `key` (parity / debugging)."""
        h = self._ensure()
        k = np.asarray(self.master_key if key is None else key).astype(np.uint32).ravel()
        out = np.empty((self.num_parallel_computations, self.num_control_parameters), np.float32)
        sg = None
        if self.sampling_method == 'cem_mppi':
            sg = np.ascontiguousarray(np.broadcast_to(np.asarray(
                self.sigma_cem_mppi[robot] if sigma is None else sigma, np.float32), (self.num_control_parameters,)))
        _lib.check(_lib.load().qpmpc_generate_noise(h, self.robot_base + int(robot), k.ctypes.data_as(_UP), _fp(sg), _fp(out)))
        return out

    def fetch_costs(self, robot: int) -> np.ndarray:
        out = np.empty(self.num_parallel_computations, np.float32)
        _lib.check(_lib.load().qpmpc_get_costs(self._ensure(), int(robot), _fp(out)))
        return out
