"""Batched contact tables from gait phase clocks, on the device (first half).

Mirrors what the callers of the sampling controller do once per MPC tick and per robot with
`PeriodicGaitGenerator.compute_contact_sequence(contact_sequence_dts, contact_sequence_lenghts)`
(helpers/periodic_gait_generator.py:93-118): R robots in one launch, float64 phase arithmetic like the numpy original.
The tables stay on the device for the next tick / step of the same controller (pass `contact_sequences=None` there)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib

# phase offsets per gait, helpers/periodic_gait_generator.py:23-39 (GaitType order of helpers/quadruped_utils.py:12-22)
PHASE_OFFSETS = {
    'trot': (0.5, 1.0, 1.0, 0.5), 'pace': (0.8, 0.3, 0.8, 0.3), 'bounding': (0.5, 0.5, 0.0, 0.0),
    'circular_crawl': (0.0, 0.25, 0.75, 0.5), 'bf_diagonal_crawl': (0.0, 0.25, 0.5, 0.75),
    'back_diagonal_crawl': (0.0, 0.5, 0.75, 0.25), 'front_diagonal_crawl': (0.5, 1.0, 0.75, 1.25),
}

_DP = C.POINTER(C.c_double)


def contact_sequences_from_phase(controller, phase_signal, step_freq, duty_factor, phase_offset, init=None,
                                 contact_sequence_dts=None, contact_sequence_lenghts=None, fetch=True):
    """phase_signal (R,4) float64 = each robot's `pgg.phase_signal`, step_freq (R,) or scalar, `init` (R,4) bool or None
    = `pgg._init`; `controller`: a `BatchedSampling_MPC` / `Sampling_MPC` (R = its robots).  Returns the (R,4,H) tables
    (or None with fetch=False) and arms them for the controller's next `compute_control` / `step` with
    `contact_sequences=None`."""
    h = controller._ensure()
    R = getattr(controller, 'num_robots', 1)
    H = controller.horizon
    ph = np.ascontiguousarray(np.asarray(phase_signal, np.float64).reshape(R, 4))
    fr = np.ascontiguousarray(np.broadcast_to(np.asarray(step_freq, np.float64), (R,)))
    off = np.ascontiguousarray(np.asarray(PHASE_OFFSETS[phase_offset] if isinstance(phase_offset, str) else phase_offset, np.float64))
    ini = None if init is None else np.ascontiguousarray(np.asarray(init).reshape(R, 4).astype(np.uint8))
    if contact_sequence_dts is None:
        contact_sequence_dts, contact_sequence_lenghts = [controller.cfg.dt], [H]
        if controller.cfg.use_nonuniform_discretization:          # the interface's lists for the fine-grained head
            contact_sequence_dts = [controller.cfg.dt_fine_grained, controller.cfg.dt]
            contact_sequence_lenghts = [controller.cfg.horizon_fine_grained, H]
    dts = np.ascontiguousarray(np.asarray(contact_sequence_dts, np.float64))
    lens = np.ascontiguousarray(np.asarray(contact_sequence_lenghts, np.int32))
    out = np.empty((R, 4, H), np.float32) if fetch else None
    _lib.check(_lib.load().qpmpc_pgg_contact_sequence(
        h, ph.ctypes.data_as(_DP), None if ini is None else ini.ctypes.data_as(C.POINTER(C.c_uint8)),
        fr.ctypes.data_as(_DP), float(duty_factor), off.ctypes.data_as(_DP), dts.ctypes.data_as(_DP),
        lens.ctypes.data_as(C.POINTER(C.c_int32)), int(dts.size), None if out is None else out.ctypes.data_as(_lib._FP)))
    return out


def footholds_reference(controller, base_position, base_ori_euler_xyz, base_xy_lin_vel, ref_base_xy_lin_vel, hips_position,
                        lift_off_z, stance_time, com_height_nominal, hip_height, hip_offset=0.1, gravity=9.81,
                        com_pos_offset_b=None, reset_history=False):
    """`FootholdReferenceGenerator.compute_footholds_reference` (helpers/foothold_reference_generator.py:53-157) for the R
    robots of `controller` in one launch, float64.  base_position / base_ori_euler_xyz (R,3), base_xy_lin_vel /
    ref_base_xy_lin_vel (R,2), hips_position (R,4,3) in FL FR RL RR order, lift_off_z (R,4) = the z of each robot's
    `frg.lift_off_positions`, stance_time (R,) or scalar.  The moving-average velocity history of each robot lives on the
    device between calls (`reset_history=True` = freshly constructed generators).  Returns ref_feet (R,4,3), world frame."""
    h = controller._ensure()
    R = getattr(controller, 'num_robots', 1)

    def arr(x, shape):
        return np.ascontiguousarray(np.broadcast_to(np.asarray(x, np.float64), shape))

    pos, eul = arr(np.reshape(base_position, (R, 3)), (R, 3)), arr(np.reshape(base_ori_euler_xyz, (R, 3)), (R, 3))
    vel, rvel = arr(np.reshape(base_xy_lin_vel, (R, 2)), (R, 2)), arr(np.reshape(ref_base_xy_lin_vel, (R, 2)), (R, 2))
    hips, loz = arr(np.reshape(hips_position, (R, 4, 3)), (R, 4, 3)), arr(np.reshape(lift_off_z, (R, 4)), (R, 4))
    st = arr(stance_time, (R,))
    off = None if com_pos_offset_b is None else arr(com_pos_offset_b, (3,))
    out = np.empty((R, 4, 3), np.float64)
    p = lambda a: a.ctypes.data_as(_DP)
    _lib.check(_lib.load().qpmpc_footholds_reference(
        h, p(pos), p(eul), p(vel), p(rvel), p(hips), p(loz), p(st), float(com_height_nominal), float(hip_height),
        float(hip_offset), float(gravity), None if off is None else p(off), int(bool(reset_history)), p(out)))
    return out
