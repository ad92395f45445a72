"""Golden vectors for the foothold reference (second half), made by running the reference's own
FootholdReferenceGenerator (helpers/foothold_reference_generator.py) HERE, in this container; /root/reference does not
travel to the GPU box, the .npz does.

    python tools/make_golden_footholds.py          ->  tests/golden/gait_footholds.npz

The reference module imports mujoco, gym_quadruped (LegsAttr) and the package's simulation config at import time; none of
them is installed, and none is used by compute_footholds_reference beyond LegsAttr as a four-field container and
config.gravity_constant.  They are stubbed below (a container class with FL/FR/RL/RR, item access and `+=` broadcast as
gym_quadruped's LegsAttr has them); the arithmetic executed is the reference's own, unmodified.
"""
import importlib.util
import os
import sys
import types

import numpy as np

REF = '/root/reference/quadruped_pympc'
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class LegsAttr:
    order = ['FL', 'FR', 'RL', 'RR']

    def __init__(self, FL, FR, RL, RR, **kw):
        self.FL, self.FR, self.RL, self.RR = FL, FR, RL, RR

    def __getitem__(self, k):
        return getattr(self, k if isinstance(k, str) else self.order[k])

    def __setitem__(self, k, v):
        setattr(self, k if isinstance(k, str) else self.order[k], v)

    def __iter__(self):
        return iter([self.FL, self.FR, self.RL, self.RR])

    def __add__(self, other):
        return LegsAttr(self.FL + other, self.FR + other, self.RL + other, self.RR + other)

    __iadd__ = __add__


def load_reference():
    for name in ('mujoco', 'gym_quadruped', 'gym_quadruped.utils'):
        sys.modules.setdefault(name, types.ModuleType(name))
    qu = types.ModuleType('gym_quadruped.utils.quadruped_utils')
    qu.LegsAttr = LegsAttr
    sys.modules['gym_quadruped.utils.quadruped_utils'] = qu
    pkg = types.ModuleType('quadruped_pympc'); pkg.__path__ = [REF]
    cfg = types.ModuleType('quadruped_pympc.config'); cfg.gravity_constant = 9.81      # config.py: gravity_constant = 9.81
    pkg.config = cfg
    helpers = types.ModuleType('quadruped_pympc.helpers'); helpers.__path__ = [REF + '/helpers']
    hq = types.ModuleType('quadruped_pympc.helpers.quadruped_utils')
    hq.GaitType = type('GaitType', (), {})
    sys.modules.update({'quadruped_pympc': pkg, 'quadruped_pympc.config': cfg, 'quadruped_pympc.helpers': helpers,
                        'quadruped_pympc.helpers.quadruped_utils': hq})
    spec = importlib.util.spec_from_file_location('quadruped_pympc.helpers.foothold_reference_generator',
                                                  REF + '/helpers/foothold_reference_generator.py')
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def main():
    with open(REF + '/config.py') as f:
        assert 'gravity_constant = 9.81' in f.read()
    mod = load_reference()
    rng = np.random.default_rng(20250)
    robots, ticks = 24, 50
    hip_heights = {'go2': 0.28, 'aliengo': 0.35, 'hyqreal1': 0.5, 'mini_cheetah': 0.225}      # hip heights spanning the supported robots' range (config.py:16)
    names = list(hip_heights)
    rec = {k: [] for k in ('base_position', 'euler', 'vel', 'ref_vel', 'hips', 'ref_feet')}
    meta = {k: [] for k in ('stance_time', 'hip_height', 'com_height', 'lift_off_z', 'com_off_b')}
    for r in range(robots):
        hh = hip_heights[names[r % 4]]
        ch = hh * rng.uniform(0.85, 1.05)
        st = rng.uniform(0.15, 0.6)
        loz = rng.normal(0.0, 0.02, 4)
        lo = LegsAttr(*[np.array([0.0, 0.0, z]) for z in loz])
        gen = mod.FootholdReferenceGenerator(stance_time=st, lift_off_positions=lo, hip_height=hh)
        off = np.zeros(3) if r % 3 else rng.normal(0, 0.02, 3)
        gen.com_pos_offset_b = off.copy()
        pos = rng.normal(0, 1.0, 3); pos[2] = ch
        eul = np.array([rng.normal(0, 0.1), rng.normal(0, 0.1), rng.uniform(-np.pi, np.pi)])
        big = r % 5 == 0                                  # drive the clips: large commanded / actual velocities
        for k in ('stance_time', 'hip_height', 'com_height', 'lift_off_z', 'com_off_b'):
            meta[k].append({'stance_time': st, 'hip_height': hh, 'com_height': ch, 'lift_off_z': loz, 'com_off_b': off}[k])
        seq = {k: [] for k in rec}
        for t in range(ticks):
            vel = rng.normal(0, 2.0 if big else 0.4, 2)
            rvel = rng.normal(0, 3.0 if big else 0.4, 2)
            hips = pos[None, :] + np.array([[0.2, 0.13, 0], [0.2, -0.13, 0], [-0.2, 0.13, 0], [-0.2, -0.13, 0]]) \
                + rng.normal(0, 0.01, (4, 3))
            out = gen.compute_footholds_reference(base_position=pos, base_ori_euler_xyz=eul, base_xy_lin_vel=vel,
                                                  ref_base_xy_lin_vel=rvel, hips_position=LegsAttr(*hips),
                                                  com_height_nominal=ch)
            for k, v in (('base_position', pos), ('euler', eul), ('vel', vel), ('ref_vel', rvel), ('hips', hips),
                         ('ref_feet', np.stack([out.FL, out.FR, out.RL, out.RR]))):
                seq[k].append(np.array(v, np.float64))
            pos = pos + np.append(vel, 0.0) * 0.002
            eul = eul + rng.normal(0, 0.01, 3)
        for k in rec:
            rec[k].append(np.stack(seq[k]))
    out = {k: np.stack(v) for k, v in rec.items()}          # (robots, ticks, ...)
    out.update({k: np.asarray(v, np.float64) for k, v in meta.items()})
    out['gravity'] = np.float64(9.81)
    out['hip_offset'] = np.float64(0.1)
    path = os.path.join(ROOT, 'tests', 'golden', 'gait_footholds.npz')
    np.savez_compressed(path, **out)
    print(path, {k: v.shape for k, v in out.items()}, os.path.getsize(path))


if __name__ == '__main__':
    main()
