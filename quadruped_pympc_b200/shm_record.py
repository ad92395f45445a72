"""MPC -> WBC output record over shared memory.

The reference's ROS2 node can run the MPC in a child process that hands its result to the whole-body controller through a
75-double `multiprocessing.shared_memory` block guarded by a seqlock (`ros2/run_controller.py:72-92` layout, `:233-236`
creation, `:373-388` writer, `:587-602` reader).  `publish` is that writer, done by the C library straight from the pinned
host packet the tick's result landed in; `read` is the reference's reader protocol (used by tests and by consumers that
are not the reference's node)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib

# payload layout, ros2/run_controller.py:72-92
N_DBL = 75
IDX_GRF, IDX_FH, IDX_JP, IDX_JV, IDX_JA, IDX_PRED = (slice(0, 12), slice(12, 24), slice(24, 36), slice(36, 48),
                                                     slice(48, 60), slice(60, 72))
IDX_BSF, IDX_LAST, IDX_STAMP = 72, 73, 74

_DP = C.POINTER(C.c_double)
_U64P = C.POINTER(C.c_uint64)


def _record_ptr(buf):
    """buf: the SharedMemory.buf memoryview (or any writable buffer of >= 75 doubles)"""
    arr = np.frombuffer(buf, dtype=np.float64, count=N_DBL)
    return arr, arr.ctypes.data_as(_DP)


def _seq_ptr(seq):
    """seq: multiprocessing.Value('Q', lock=False) (a ctypes c_ulong / c_ulonglong) or a 1-element uint64 array"""
    if isinstance(seq, np.ndarray):
        return seq.ctypes.data_as(_U64P)
    return C.cast(C.addressof(seq), _U64P)


def pack(buf, seq, grfs, predicted_state, best_sample_freq, last_mpc_loop_time, current_contact=None, footholds=None):
    """pack caller-supplied values (no GPU involved)"""
    arr, rp = _record_ptr(buf)
    g = np.ascontiguousarray(grfs, np.float32).reshape(12)
    p = np.ascontiguousarray(np.asarray(predicted_state, np.float32).reshape(-1)[:24])
    if p.size < 24:
        p = np.ascontiguousarray(np.concatenate([p, np.zeros(24 - p.size, np.float32)]))
    cc = None if current_contact is None else np.ascontiguousarray(current_contact, np.float32).reshape(4)
    fh = None if footholds is None else np.ascontiguousarray(footholds, np.float64).reshape(12)
    _lib.check(_lib.load().qpmpc_shm_pack(rp, _seq_ptr(seq), g.ctypes.data_as(_lib._FP),
                                          None if cc is None else cc.ctypes.data_as(_lib._FP),
                                          None if fh is None else fh.ctypes.data_as(_DP), p.ctypes.data_as(_lib._FP),
                                          float(best_sample_freq), float(last_mpc_loop_time)))
    del arr


def publish(controller, buf, seq, current_contact=None, footholds=None, last_mpc_loop_time=0.0, robot=0):
    """write the last tick / step result of `controller` (robot `robot` of a batched one) into the record"""
    arr, rp = _record_ptr(buf)
    cc = None if current_contact is None else np.ascontiguousarray(current_contact, np.float32).reshape(4)
    fh = None if footholds is None else np.ascontiguousarray(footholds, np.float64).reshape(12)
    _lib.check(_lib.load().qpmpc_publish_shm(controller._ensure(), int(robot), rp, _seq_ptr(seq),
                                             None if cc is None else cc.ctypes.data_as(_lib._FP),
                                             None if fh is None else fh.ctypes.data_as(_DP), float(last_mpc_loop_time)))
    del arr


def read(buf, seq):
    """the reader of ros2/run_controller.py:587-602: a copy of the payload, or None while a write is in progress / torn"""
    s1 = seq.value if hasattr(seq, 'value') else int(seq[0])
    if s1 % 2:
        return None
    tmp = np.frombuffer(buf, dtype=np.float64, count=N_DBL).copy()
    s2 = seq.value if hasattr(seq, 'value') else int(seq[0])
    if s1 != s2 or s2 % 2:
        return None
    return tmp
