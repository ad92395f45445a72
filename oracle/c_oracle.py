"""ctypes binding of oracle/sampling_oracle.c (TEST INFRASTRUCTURE ONLY; see sampling_oracle.py header).

Used by tests/ as a second checker, and by bench.py's ``cpu_baseline`` / ``--impl reference`` legs as the
timed CPU implementation of the tick (OpenMP over samples, all host threads).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from .sampling_oracle import OracleConfig, Q_DIAG, PGG_DUTY_ADAPTIVE

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, 'liboracle_srbd.so')
ORC_MAX_H, ORC_MAX_NS = 64, 16
PARAM_ID = {'zero_order': 0, 'linear_spline': 1, 'cubic_spline': 2}
METHOD_ID = {'random_sampling': 0, 'mppi': 1, 'cem_mppi': 2}


class OrcCfg(C.Structure):
    _fields_ = [('N', C.c_int32), ('H', C.c_int32), ('ns', C.c_int32), ('param', C.c_int32),
                ('method', C.c_int32), ('adaptive', C.c_int32), ('mass', C.c_float),
                ('inertia', C.c_float * 9), ('dts', C.c_float * ORC_MAX_H),
                ('bounds', C.c_float * (ORC_MAX_NS + 1)), ('mu', C.c_float), ('fz_min', C.c_float),
                ('fz_max', C.c_float), ('mg', C.c_float), ('qdiag', C.c_float * 24), ('mpc_dt', C.c_float),
                ('duty', C.c_float), ('sigma_rs', C.c_float * 3), ('sigma_mppi', C.c_float),
                ('freq_avail', C.c_float * 3), ('inertia_inv', C.c_float * 9), ('inv_m', C.c_float),
                ('P_leg', C.c_int32), ('P', C.c_int32), ('elite_k', C.c_int32)]


class OrcOut(C.Structure):
    _fields_ = [('grfs', C.c_float * 12), ('pred', C.c_float * 24), ('best_cost', C.c_float),
                ('best_freq', C.c_float), ('best_index', C.c_int32)]


def build(force: bool = False) -> str:
    """Compile liboracle_srbd.so next to this file (gcc, OpenMP, no FMA contraction)."""
    src = os.path.join(_HERE, 'sampling_oracle.c')
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(['gcc', '-O2', '-ffp-contract=off', '-fopenmp', '-fPIC', '-std=c11', '-shared',
                               src, '-o', _LIB_PATH, '-lm'])
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        _lib = C.CDLL(_LIB_PATH)
        _lib.orc_num_threads.restype = C.c_int
    return _lib


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float)) if a is not None else None


class COracle:
    def __init__(self, cfg: OracleConfig):
        self.cfg = cfg
        c = self.c = OrcCfg()
        c.N, c.H, c.ns = cfg.num_parallel_computations, cfg.horizon, cfg.num_splines
        c.param, c.method = PARAM_ID[cfg.control_parametrization], METHOD_ID[cfg.sampling_method]
        c.adaptive = int(cfg.optimize_step_freq)
        c.mass = cfg.mass
        c.inertia[:] = [float(x) for x in np.asarray(cfg.inertia, np.float32).ravel()]
        if cfg.use_nonuniform_discretization:
            dts = [cfg.dt_fine_grained] * cfg.horizon_fine_grained + [cfg.dt] * (cfg.horizon - cfg.horizon_fine_grained)
        else:
            dts = [cfg.dt] * cfg.horizon
        for i, d in enumerate(dts):
            c.dts[i] = d
        for i, b in enumerate(np.linspace(0, cfg.horizon, cfg.num_splines + 1).astype(np.float32)):
            c.bounds[i] = float(b)
        c.mu, c.fz_min, c.fz_max = cfg.mu, cfg.grf_min, cfg.grf_max
        c.mg = float(np.float32(cfg.mass * 9.81))
        c.qdiag[:] = [float(x) for x in Q_DIAG]
        c.mpc_dt, c.duty = cfg.dt, PGG_DUTY_ADAPTIVE
        c.sigma_rs[:] = [float(x) for x in cfg.sigma_random_sampling]
        c.sigma_mppi = cfg.sigma_mppi
        c.freq_avail[:] = [float(x) for x in cfg.step_freq_available]
        lib().orc_cfg_finalize(C.byref(c))
        self.N, self.P, self.H = c.N, c.P, c.H
        self._scratch = np.empty((self.N, self.P), np.float32)

    @staticmethod
    def _f32(a, shape=None):
        a = np.ascontiguousarray(a, dtype=np.float32)
        return a if shape is None else a.reshape(shape)

    def rollout(self, state, ref, contact, params, timing=None, freqs=None):
        state, ref = self._f32(state), self._f32(ref)
        contact = self._f32(np.asarray(contact)[:, :self.H])
        params = self._f32(params, (self.N, self.P))
        timing = self._f32(timing if timing is not None else np.zeros(4))
        freqs = self._f32(freqs) if freqs is not None else None
        costs = np.empty(self.N, np.float32)
        lib().orc_rollout(C.byref(self.c), _fp(state), _fp(ref), _fp(contact), _fp(timing), _fp(params),
                          _fp(freqs), _fp(costs))
        return costs

    def compute_control(self, state, ref, contact, best, noise, timing=None, freqs=None):
        state, ref = self._f32(state), self._f32(ref)
        contact = self._f32(np.asarray(contact)[:, :self.H])
        best, noise = self._f32(best), self._f32(noise, (self.N, self.P))
        timing = self._f32(timing if timing is not None else np.zeros(4))
        freqs = self._f32(freqs) if freqs is not None else None
        costs = np.empty(self.N, np.float32)
        best_new = np.empty(self.P, np.float32)
        sigma_new = np.zeros(self.P, np.float32)
        elite = np.full(max(self.c.elite_k, 1), -1, np.int32)
        out = OrcOut()
        lib().orc_compute_control(C.byref(self.c), _fp(state), _fp(ref), _fp(contact), _fp(timing), _fp(best),
                                  _fp(noise), _fp(freqs), _fp(self._scratch), _fp(costs), _fp(best_new),
                                  _fp(sigma_new), elite.ctypes.data_as(C.POINTER(C.c_int32)), C.byref(out))
        res = dict(nmpc_GRFs=np.array(out.grfs[:], np.float32), nmpc_predicted_state=np.array(out.pred[:], np.float32),
                   best_control_parameters=best_new, best_cost=np.float32(out.best_cost),
                   best_freq=np.float32(out.best_freq), best_index=int(out.best_index), costs=costs)
        if self.cfg.sampling_method == 'cem_mppi':
            k = min(self.c.elite_k, self.N)
            res['new_sigma_cem_mppi'] = sigma_new
            res['elite_indices'] = elite[:k].astype(np.int64)
        return res

    def noise(self, seed, tick, sigma=None, robot=0, rows=None):
        r0, r1 = rows if rows is not None else (0, self.N)
        out = np.empty((r1 - r0, self.P), np.float32)
        sig = self._f32(sigma if sigma is not None else np.zeros(self.P))
        lib().orc_noise_generate(C.byref(self.c), C.c_uint32(seed), C.c_uint32(robot), C.c_uint32(tick), _fp(sig),
                                 C.c_int(r0), C.c_int(r1), _fp(out))
        return out

    def freq_choice(self, seed, tick, robot=0, rows=None):
        r0, r1 = rows if rows is not None else (0, self.N)
        out = np.empty(r1 - r0, np.int32)
        lib().orc_freq_choice(C.c_uint32(seed), C.c_uint32(robot), C.c_uint32(tick), C.c_int(r0), C.c_int(r1),
                              out.ctypes.data_as(C.POINTER(C.c_int32)))
        return out


def philox4x32_10(ctr, key):
    ctr = np.ascontiguousarray(ctr, np.uint32)
    key = np.ascontiguousarray(key, np.uint32)
    out = np.empty(4, np.uint32)
    u32p = C.POINTER(C.c_uint32)
    lib().orc_philox4x32_10(ctr.ctypes.data_as(u32p), key.ctypes.data_as(u32p), out.ctypes.data_as(u32p))
    return out


def num_threads() -> int:
    return int(lib().orc_num_threads())


def set_num_threads(n: int) -> None:
    lib().orc_set_num_threads(C.c_int(n))
