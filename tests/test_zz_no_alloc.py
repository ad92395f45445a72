"""GPU: qpmpc_stance_torques_static, the entry point that allocates nothing (include/qpmpc.h; no counterpart in the
reference).  Called through the C ABI (ctypes) and compared with the
existing qpmpc_stance_torques on the same device-resident tick results, on calls in a row: the static scratch keeps the
previous call's contents, so a later call with other inputs (and with fewer robots) must not see them.

Collected last on purpose; self-contained, so it also passes alone in a fresh process."""
import ctypes as C

import numpy as np
import pytest

import quadruped_pympc_b200 as qp
from quadruped_pympc_b200 import _lib, synthetic
from quadruped_pympc_b200.batched import BatchedSampling_MPC

pytestmark = pytest.mark.gpu

_FP = C.POINTER(C.c_float)


def _fp(a):
    return a.ctypes.data_as(_FP)


def _static(handle, jac, contact, R):
    jac = np.ascontiguousarray(jac, np.float32).reshape(R, 4, 3, 3)
    con = np.ascontiguousarray(contact, np.float32).reshape(R, 4)
    grf, tau = np.full((R, 12), np.nan, np.float32), np.full((R, 12), np.nan, np.float32)
    _lib.check(_lib.load().qpmpc_stance_torques_static(handle, _fp(jac), _fp(con), _fp(grf), _fp(tau)))
    return grf, tau


def _agree(got, want):
    # the tolerances of tests/test_gpu_batched.py::test_stance_torque_epilogue (GRFs rtol 1e-7, torques rtol 1e-5 + 1e-4)
    np.testing.assert_allclose(got[0], want[0], rtol=1e-7, atol=0)
    np.testing.assert_allclose(got[1], want[1], rtol=1e-5, atol=1e-4)
    assert np.all(np.isfinite(got[0])) and np.all(np.isfinite(got[1]))


def test_static_scratch_entry_point_equals_the_allocating_one(built):
    R, N = 6, 600
    cfg = qp.SamplerConfig.for_robot('go2', sampling_method='mppi', num_parallel_computations=N)
    b = BatchedSampling_MPC(cfg, R)
    scs = [synthetic.scenario('go2', 'trot' if r % 2 == 0 else 'pace', seed=r) for r in range(R)]
    st = np.stack([s['state'] for s in scs]).astype(np.float32)
    rf = np.stack([s['reference'] for s in scs]).astype(np.float32)
    cs = np.stack([s['contact_sequence'] for s in scs]).astype(np.float32)
    out = b.with_newkey().compute_control(st, rf, cs)
    rng = np.random.RandomState(11)
    jac1, jac2 = rng.randn(R, 4, 3, 3).astype(np.float32), rng.randn(R, 4, 3, 3).astype(np.float32)
    con1 = cs[:, :, 0]
    con2 = 1.0 - con1                                        # the other legs: every masked GRF changes
    want1, want2 = b.stance_torques(jac1, con1), b.stance_torques(jac2, con2)
    assert not np.array_equal(want1[1], want2[1])
    h = b._ensure()
    # two calls in a row, different inputs; then the first again (nothing of call 2 may be left)
    _agree(_static(h, jac1, con1, R), want1)
    _agree(_static(h, jac2, con2, R), want2)
    _agree(_static(h, jac1, con1, R), want1)
    # against the reference's numpy expressions too (I:225-230, wb_interface.py:372-375), one robot
    g, t = _static(h, jac1, con1, R)
    for leg in range(4):
        F = out[0][2, 3 * leg:3 * leg + 3].astype(np.float64) * float(con1[2, leg])
        np.testing.assert_allclose(g[2, 3 * leg:3 * leg + 3], F, rtol=1e-7, atol=0)
        np.testing.assert_allclose(t[2, 3 * leg:3 * leg + 3], -np.matmul(jac1[2, leg].astype(np.float64).T, F), rtol=1e-5, atol=1e-4)
    # a single controller right after the batch: fewer robots than the scratch last held
    c = qp.Sampling_MPC(cfg)
    sc = synthetic.scenario('go2', 'trot', seed=0)
    c.with_newkey().jitted_compute_control(sc['state'], sc['reference'], sc['contact_sequence'], np.zeros(96, np.float32),
                                           c.master_key, sc['phase_signal'], 1.4, 0)
    want = c.stance_torques(jac2[0], sc['contact_sequence'][:, 0])
    for _ in range(2):
        got = _static(c._ensure(), jac2[0], sc['contact_sequence'][:, 0], 1)
        _agree((got[0][0], got[1][0]), want)
    # outputs are optional, like the allocating entry point's
    jac, con = np.ascontiguousarray(jac1), np.ascontiguousarray(con1, np.float32)
    tau = np.empty((R, 12), np.float32)
    _lib.check(_lib.load().qpmpc_stance_torques_static(h, _fp(jac), _fp(con), None, _fp(tau)))
    np.testing.assert_allclose(tau, want1[1], rtol=1e-5, atol=1e-4)


def test_static_scratch_entry_point_rejects_more_robots_than_it_holds(built):
    lib = _lib.load()
    R = 1025                                                 # QPMPC_NO_ALLOC_MAX_ROBOTS + 1
    cfg = qp.SamplerConfig.for_robot('go2', sampling_method='mppi', num_parallel_computations=128)
    b = BatchedSampling_MPC(cfg, R)
    jac, con = np.zeros((R, 4, 3, 3), np.float32), np.zeros((R, 4), np.float32)
    grf, tau = np.empty((R, 12), np.float32), np.empty((R, 12), np.float32)
    rc = lib.qpmpc_stance_torques_static(b._ensure(), _fp(jac), _fp(con), _fp(grf), _fp(tau))
    assert rc != _lib.OK
    msg = lib.qpmpc_last_error().decode()
    assert 'QPMPC_NO_ALLOC_MAX_ROBOTS' in msg and '1025' in msg, msg
    assert lib.qpmpc_stance_torques_static(None, _fp(jac), _fp(con), _fp(grf), _fp(tau)) != _lib.OK
