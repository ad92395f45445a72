"""Synthetic, seeded inputs of the named shapes for tests and bench.py.

Contact tables follow the reference's host-side numpy gait generator (helpers/periodic_gait_generator.py:24-27,48-76,
93-118): per-leg phase offsets, phase += dt*f (mod 1), contact = phase < duty; column 0 is the current contact, the
following columns advance by the MPC dt.  States / references are drawn as SURVEY 8d prescribes.
"""
from __future__ import annotations

import numpy as np

GAITS = {   # offsets, duty, step_freq : helpers/periodic_gait_generator.py:24-27 + config.py:216-218
    'trot': ((0.5, 1.0, 1.0, 0.5), 0.65, 1.4),
    'pace': ((0.8, 0.3, 0.8, 0.3), 0.7, 1.4),
    'full_stance': ((0.0, 0.0, 0.0, 0.0), 2.0, 1.4),
}
GEOMETRY = {  # nominal height z0, half length L, half width W (hip_height lives in un-vendored gym-quadruped)
    'go2': (0.30, 0.20, 0.13), 'aliengo': (0.35, 0.24, 0.15), 'hyqreal1': (0.55, 0.44, 0.30),
}


def phase_signal(gait: str, sim_steps: int = 0, sim_dt: float = 0.002):
    offsets, duty, f = GAITS[gait]
    ph = np.asarray(offsets, dtype=np.float64).copy()
    for _ in range(sim_steps):
        ph = (ph + sim_dt * f) % 1.0
    return ph


def contact_sequence(gait: str, horizon: int = 12, mpc_dt: float = 0.02, sim_steps: int = 0):
    """(4, H) table of 0/1 floats + the phase signal it was generated from."""
    offsets, duty, f = GAITS[gait]
    ph0 = phase_signal(gait, sim_steps)
    ph = ph0.copy()
    seq = np.zeros((4, horizon), dtype=np.float64)
    ph = ph % 1.0                                   # run(0.0, f)
    seq[:, 0] = (ph < duty).astype(np.float64)
    for i in range(1, horizon):
        ph = (ph + mpc_dt * f) % 1.0
        seq[:, i] = (ph < duty).astype(np.float64)
    return seq, ph0


def make_state(robot: str = 'go2', seed: int = 0):
    """(state24, reference24) float64, SURVEY 8d distributions."""
    z0, L, W = GEOMETRY[robot]
    rng = np.random.RandomState(seed)
    s = np.zeros(24)
    s[2] = z0
    s[3:6] = rng.uniform(-0.3, 0.3, 3)
    s[6:8] = rng.uniform(-0.05, 0.05, 2)
    s[8] = rng.uniform(-np.pi, np.pi)
    s[9:12] = rng.uniform(-0.2, 0.2, 3)
    feet = np.array([[L, W, 0.0], [L, -W, 0.0], [-L, W, 0.0], [-L, -W, 0.0]]) + rng.uniform(-0.03, 0.03, (4, 3))
    s[12:] = feet.ravel()
    r = np.zeros(24)
    r[2] = z0
    r[3] = 0.3
    r[12:] = s[12:]
    return s, r


def scenario(robot='go2', gait='trot', horizon=12, mpc_dt=0.02, seed=0, sim_steps=None):
    """dict(state, reference, contact_sequence, phase_signal, nominal_step_freq)"""
    if sim_steps is None:
        sim_steps = 5 * (seed % 70)                 # ticks happen every 5 sim steps (100 Hz MPC in a 500 Hz sim)
    s, r = make_state(robot, seed)
    cs, ph = contact_sequence(gait, horizon, mpc_dt, sim_steps)
    return dict(state=s, reference=r, contact_sequence=cs, phase_signal=ph, nominal_step_freq=GAITS[gait][2])
