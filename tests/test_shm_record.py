"""MPC -> WBC shared-memory record (layout and seqlock of ros2/run_controller.py:72-92,373-388,587-602).
CPU: the C writer against the reference's packing expressions restated here, the seqlock parity protocol, and a reader
process that never sees a torn record while the C writer hammers it.  GPU: the record written from a real tick."""
import multiprocessing as mp
import time
from multiprocessing import shared_memory

import numpy as np
import pytest

from quadruped_pympc_b200 import shm_record as sr


def _reference_payload(grfs, contact, footholds, pred, bsf, last):
    """the writer's expressions, run_controller.py:378-385 after interface :170-180,225-230 (sampling controller)"""
    arr = np.zeros(sr.N_DBL)
    g = np.asarray(grfs)
    arr[sr.IDX_GRF] = np.concatenate([g[3 * l:3 * l + 3] * contact[l] for l in range(4)])
    arr[sr.IDX_FH] = np.asarray(footholds).reshape(-1)
    arr[sr.IDX_JP] = None
    arr[sr.IDX_JV] = None
    arr[sr.IDX_JA] = None
    arr[sr.IDX_PRED] = np.asarray(pred).reshape(-1)[:12]
    arr[sr.IDX_BSF] = float(bsf)
    arr[sr.IDX_LAST] = float(last)
    return arr


def test_layout_constants():
    assert sr.N_DBL == 75 and sr.IDX_PRED == slice(60, 72) and (sr.IDX_BSF, sr.IDX_LAST, sr.IDX_STAMP) == (72, 73, 74)


def test_pack_matches_the_reference_writer(built):
    rng = np.random.RandomState(0)
    shm = shared_memory.SharedMemory(create=True, size=sr.N_DBL * 8)
    try:
        seq = mp.Value('Q', 0, lock=False)                       # run_controller.py:236
        for it in range(5):
            grfs = rng.randn(12).astype(np.float32) * 50
            contact = rng.randint(0, 2, 4).astype(np.float32)
            fh = rng.randn(12)
            pred = rng.randn(24).astype(np.float32)
            t0 = time.monotonic()
            sr.pack(shm.buf, seq, grfs, pred, float(np.float32(1.4 + it)), 0.01 * it, current_contact=contact, footholds=fh)
            t1 = time.monotonic()
            assert seq.value == 2 * (it + 1)                     # even = stable, +2 per record
            got = sr.read(shm.buf, seq)
            want = _reference_payload(grfs.astype(np.float64), contact.astype(np.float64), fh, pred, np.float32(1.4 + it), 0.01 * it)
            np.testing.assert_array_equal(got[:sr.IDX_STAMP], want[:sr.IDX_STAMP])     # NaN == NaN here
            assert t0 <= got[sr.IDX_STAMP] <= t1
        # a writer that died mid-record left the sequence odd: the next writer keeps it odd, then publishes (:374-376,388)
        seq.value = 11
        assert sr.read(shm.buf, seq) is None
        sr.pack(shm.buf, seq, grfs, pred, 2.0, 0.0)
        assert seq.value == 12
        got = sr.read(shm.buf, seq)
        np.testing.assert_array_equal(got[sr.IDX_GRF], grfs.astype(np.float64))          # contact NULL = ones
        np.testing.assert_array_equal(got[sr.IDX_FH], np.zeros(12))
    finally:
        shm.close(); shm.unlink()


def _hammer(name, seq, n, ready):
    from quadruped_pympc_b200 import shm_record as w
    shm = shared_memory.SharedMemory(name=name)
    ready.set()
    for k in range(1, n + 1):
        v = np.full(12, float(k), np.float32)
        w.pack(shm.buf, seq, v, np.full(24, float(k), np.float32), float(k), float(k), footholds=np.full(12, float(k)))
    shm.close()


def test_reader_never_sees_a_torn_record(built):
    shm = shared_memory.SharedMemory(create=True, size=sr.N_DBL * 8)
    try:
        np.ndarray((sr.N_DBL,), dtype=np.float64, buffer=shm.buf)[:] = 0.0
        seq = mp.Value('Q', 0, lock=False)
        ready = mp.Event()
        n = 200000
        proc = mp.Process(target=_hammer, args=(shm.name, seq, n, ready))
        proc.start()
        ready.wait(30)
        good = seen_last = 0
        deadline = time.time() + 60
        while time.time() < deadline:
            rec = sr.read(shm.buf, seq)
            if rec is None:
                continue
            k = rec[sr.IDX_BSF]
            body = np.concatenate([rec[sr.IDX_GRF], rec[sr.IDX_FH], rec[sr.IDX_PRED], rec[[sr.IDX_BSF, sr.IDX_LAST]]])
            assert np.all(body == k), (k, body)
            good += 1
            seen_last = k
            if k == n:
                break
        proc.join(30)
        assert proc.exitcode == 0 and seen_last == n and good > 10
        assert seq.value == 2 * n
    finally:
        shm.close(); shm.unlink()


@pytest.mark.gpu
def test_publish_from_a_tick(built):
    import quadruped_pympc_b200 as qp
    from quadruped_pympc_b200 import synthetic
    from quadruped_pympc_b200.batched import BatchedSampling_MPC
    shm = shared_memory.SharedMemory(create=True, size=sr.N_DBL * 8)
    try:
        seq = mp.Value('Q', 0, lock=False)
        cfg = qp.SamplerConfig.for_robot('go2', sampling_method='mppi', num_parallel_computations=2048)
        ctrl = qp.Sampling_MPC(cfg)
        sc = synthetic.scenario('go2', 'trot', seed=3)
        with pytest.raises(Exception):
            sr.publish(ctrl, shm.buf, seq)                       # nothing to publish before the first tick
        best = np.zeros(ctrl.num_control_parameters, np.float32)
        out = ctrl.jitted_compute_control(sc['state'], sc['reference'], sc['contact_sequence'], best, ctrl.master_key,
                                          sc['phase_signal'], sc['nominal_step_freq'], 1)
        contact = sc['contact_sequence'][:, 0]
        fh = np.asarray(sc['reference'][12:24], np.float64)
        sr.publish(ctrl, shm.buf, seq, current_contact=contact, footholds=fh, last_mpc_loop_time=0.004)
        got = sr.read(shm.buf, seq)
        want = _reference_payload(np.asarray(out[0], np.float64), np.asarray(contact, np.float64), fh, np.asarray(out[2]),
                                  out[5], 0.004)
        np.testing.assert_array_equal(got[:sr.IDX_STAMP], want[:sr.IDX_STAMP])
        assert seq.value == 2 and got[sr.IDX_STAMP] > 0
        # batched: robot r's record
        R = 5
        b = BatchedSampling_MPC(cfg, R)
        scs = [synthetic.scenario('go2', 'trot', seed=r) for r in range(R)]
        res = b.compute_control(np.stack([s['state'] for s in scs]), np.stack([s['reference'] for s in scs]),
                                np.stack([s['contact_sequence'] for s in scs]))
        for r in (0, 3, 4):
            sr.publish(b, shm.buf, seq, robot=r)
            got = sr.read(shm.buf, seq)
            np.testing.assert_array_equal(got[sr.IDX_GRF], np.asarray(res[0][r], np.float64))
            np.testing.assert_array_equal(got[sr.IDX_PRED], np.asarray(res[1][r][:12], np.float64))
        with pytest.raises(Exception):
            sr.publish(b, shm.buf, seq, robot=R)
    finally:
        shm.close(); shm.unlink()
