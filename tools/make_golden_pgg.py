"""tests/golden/gait_pgg.npz: contact tables produced by the REFERENCE'S OWN PeriodicGaitGenerator
(helpers/periodic_gait_generator.py, executed in place from /root/reference; gym_quadruped / mujoco are stubbed, they are
only imported for type names).  Run in the build container:   python tools/make_golden_pgg.py"""
import os
import sys
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REFERENCE = os.environ.get('QPMPC_REFERENCE', '/root/reference')


def main():
    sys.path.insert(0, REFERENCE)
    mj, mjv = types.ModuleType('mujoco'), types.ModuleType('mujoco.viewer')
    mjv.Handle = type('Handle', (), {})
    mj.viewer = mjv
    sys.modules['mujoco'], sys.modules['mujoco.viewer'] = mj, mjv
    gq, gu, gqu = types.ModuleType('gym_quadruped'), types.ModuleType('gym_quadruped.utils'), types.ModuleType('gym_quadruped.utils.quadruped_utils')
    gqu.LegsAttr = type('LegsAttr', (), {})
    gq.utils, gu.quadruped_utils = gu, gqu
    sys.modules.update({'gym_quadruped': gq, 'gym_quadruped.utils': gu, 'gym_quadruped.utils.quadruped_utils': gqu})
    from quadruped_pympc.helpers.periodic_gait_generator import PeriodicGaitGenerator
    from quadruped_pympc.helpers.quadruped_utils import GaitType

    rng = np.random.RandomState(7)
    H = 12
    cases = []
    gaits = [GaitType.TROT, GaitType.PACE, GaitType.BOUNDING, GaitType.CIRCULARCRAWL, GaitType.BFDIAGONALCRAWL,
             GaitType.BACKDIAGONALCRAWL, GaitType.FRONTDIAGONALCRAWL]
    for k in range(160):
        g = gaits[k % len(gaits)]
        duty = float(rng.choice([0.5, 0.6, 0.65, 0.7, 0.8, 0.9]))
        freq = float(rng.choice([0.8, 1.3, 1.4, 2.0, 2.4]))
        pgg = PeriodicGaitGenerator(duty_factor=duty, step_freq=freq, gait_type=g.value, horizon=H)
        for _ in range(int(rng.randint(0, 400))):              # simulation steps at 500 Hz, as the simulation loop does
            pgg.run(0.002, freq)
        init = None
        if k % 5 == 4:                                          # start-and-stop style restart: legs still in their init delay
            init = [bool(b) for b in rng.randint(0, 2, 4)]
            pgg.set_phase_signal(np.array(pgg.phase_signal), init)
        if k % 3 == 2:
            dts, lens = [0.01, 0.02], [int(rng.randint(1, 6)), H]
        else:
            dts, lens = [0.02], [H]
        phase0, init0 = np.array(pgg.phase_signal), list(pgg._init)
        table = pgg.compute_contact_sequence(contact_sequence_dts=dts, contact_sequence_lenghts=lens)
        assert np.array_equal(phase0, pgg.phase_signal)        # the look-ahead restores the generator (:117)
        cases.append(dict(phase=phase0, init=np.array(init0, bool), freq=freq, duty=duty, offset=np.array(pgg.phase_offset, float),
                          dts=np.array(dts + [0.0] * (2 - len(dts))), lens=np.array(lens + [0] * (2 - len(lens))), n_dts=len(dts),
                          table=table))
    out = {k: np.stack([c[k] for c in cases]) for k in cases[0]}
    path = os.path.join(ROOT, 'tests', 'golden', 'gait_pgg.npz')
    np.savez_compressed(path, **out)
    print('wrote', path, {k: v.shape for k, v in out.items()}, 'stance fraction', float(out['table'].mean()))


if __name__ == '__main__':
    main()
