"""Executable proof of the INTEGRATION.md patch (shared by tests/test_reference_interface.py and
tools/make_interface_trace.py).

What runs here is the REFERENCE'S OWN `SRBDControllerInterface` (quadruped_pympc/interfaces/srbd_controller_interface.py,
unmodified source read from $QPMPC_REFERENCE or /root/reference) with the four-line patch of INTEGRATION.md section 1
applied to its source text in memory, `mpc_params['sampling_backend'] = 'b200'`, and `quadruped_pympc_b200.Sampling_MPC`
behind it.  The C library is the real libqpmpc_b200.so for everything that needs no device (config defaults) and a
recording stand-in for `qpmpc_create` / `qpmpc_tick` / `qpmpc_get_costs` / `qpmpc_destroy`, so the test sees exactly what
crosses the C ABI for every tick the interface drives.  Modules the reference imports but does not vendor
(`gym_quadruped`, `mujoco.viewer`) are stubbed with the few names it touches on this path.
"""
import ctypes as C
import importlib
import os
import re
import sys
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REFERENCE = os.environ.get('QPMPC_REFERENCE', '/root/reference')
INTERFACE_PY = os.path.join(REFERENCE, 'quadruped_pympc', 'interfaces', 'srbd_controller_interface.py')


def reference_available():
    return os.path.exists(INTERFACE_PY)


class LegsAttr:
    """Stand-in for gym_quadruped.utils.quadruped_utils.LegsAttr (un-vendored dependency): four named leg slots."""

    def __init__(self, FL, FR, RL, RR):
        self.FL, self.FR, self.RL, self.RR = FL, FR, RL, RR

    def to_list(self):
        return [self.FL, self.FR, self.RL, self.RR]


def install_stubs():
    if ROOT not in sys.path:
        sys.path.insert(0, ROOT)
    if REFERENCE not in sys.path:
        sys.path.insert(0, REFERENCE)
    mj, mjv = types.ModuleType('mujoco'), types.ModuleType('mujoco.viewer')
    mjv.Handle = type('Handle', (), {})
    mj.viewer = mjv
    sys.modules.setdefault('mujoco', mj)
    sys.modules.setdefault('mujoco.viewer', mjv)
    gq, rc = types.ModuleType('gym_quadruped'), types.ModuleType('gym_quadruped.robot_cfgs')
    gu, gqu = types.ModuleType('gym_quadruped.utils'), types.ModuleType('gym_quadruped.utils.quadruped_utils')

    class RobotConfig:                                    # only attribute access happens in config.py:11-16
        leg_joints = None
        feet_geom_names = None
        qpos0_js = None
        hip_height = 0.3

    rc.RobotConfig, rc.get_robot_config = RobotConfig, (lambda robot_name: RobotConfig())
    gqu.LegsAttr = LegsAttr
    gq.robot_cfgs, gq.utils, gu.quadruped_utils = rc, gu, gqu
    for name, mod in (('gym_quadruped', gq), ('gym_quadruped.robot_cfgs', rc), ('gym_quadruped.utils', gu),
                      ('gym_quadruped.utils.quadruped_utils', gqu)):
        sys.modules.setdefault(name, mod)
    import quadruped_pympc.config as config
    return config


def integration_patch():
    """(config line, patched interface block) = the two python code blocks of INTEGRATION.md section 1."""
    text = open(os.path.join(ROOT, 'INTEGRATION.md')).read()
    blocks = re.findall(r'```python\n(.*?)```', text, flags=re.S)
    return blocks[0], blocks[1]


def patched_interface_module():
    """The reference's interface module with the INTEGRATION.md block substituted for srbd_controller_interface.py:77-83."""
    src = open(INTERFACE_PY).read()
    _, new_block = integration_patch()
    m = re.search(r'        elif self\.type == "sampling":\n.*?            self\.controller = Sampling_MPC\(\)\n', src, flags=re.S)
    assert m, 'the block INTEGRATION.md patches (srbd_controller_interface.py:77-83) was not found in the reference'
    assert new_block.startswith('        elif self.type == "sampling":') and new_block.rstrip().endswith('self.controller = Sampling_MPC()')
    patched = src[:m.start()] + new_block + src[m.end():]
    mod = types.ModuleType('srbd_controller_interface_patched')
    mod.__file__ = INTERFACE_PY
    exec(compile(patched, INTERFACE_PY + ' (+ INTEGRATION.md patch)', 'exec'), mod.__dict__)
    return mod


class RecordingLib:
    """libqpmpc_b200 with qpmpc_create / qpmpc_tick / qpmpc_get_costs / qpmpc_destroy replaced by recorders.  A tick's
    result is a deterministic function of its inputs (so that what the interface carries from one tick to the next --
    best_control_parameters, sigma -- is visible in the next call's arguments)."""

    def __init__(self, real):
        self._real = real
        self.creates, self.ticks = [], []

    def __getattr__(self, name):
        return getattr(self._real, name)

    def qpmpc_create(self, cfg_ref, handle_ref):
        c = cfg_ref._obj
        self.creates.append({f: (list(getattr(c, f)) if hasattr(getattr(c, f), '__len__') else getattr(c, f))
                             for f, _ in c._fields_})
        self._P = {0: 3 * c.horizon, 1: 3 * (c.num_splines + 1), 2: 12 * c.num_splines}[c.parametrization] * 4
        self._N, self._H, self._cem = c.num_samples, c.horizon, c.method == 2
        handle_ref._obj.value = 0xB200
        return 0

    def qpmpc_num_local_samples(self, h):
        return self._N

    def qpmpc_destroy(self, h):
        return None

    def qpmpc_get_costs(self, h, robot, dst):
        np.ctypeslib.as_array(dst, shape=(self._N,))[...] = 1.0
        return 0

    def qpmpc_tick(self, h, in_ref, out_ref):
        i, o = in_ref._obj, out_ref._obj
        arr = lambda p, shape: np.ctypeslib.as_array(p, shape=shape).copy() if p else None
        rec = dict(state=arr(i.state, (24,)), reference=arr(i.reference, (24,)),
                   contact_sequence=arr(i.contact_sequence, (4, i.contact_ld))[:, :self._H], best=arr(i.best, (self._P,)),
                   sigma=arr(i.sigma, (self._P,)), phase_signal=arr(i.phase_signal, (4,)),
                   nominal_step_freq=float(i.nominal_step_freq), optimize_swing=int(i.optimize_swing),
                   key=np.array([i.key[0], i.key[1]], np.uint32), noise_injected=bool(i.noise), freqs_injected=bool(i.step_freqs))
        self.ticks.append(rec)
        n = len(self.ticks)
        o.grfs[:] = [float(10 * n + k) for k in range(12)]
        o.predicted_state[:] = [float(100 * n + k) for k in range(24)]
        o.best_cost, o.best_freq, o.best_index, o.num_elite = 7.0 * n, 1.4 + 0.1 * n, n, 0
        np.ctypeslib.as_array(o.best, shape=(self._P,))[...] = rec['best'] + np.float32(n)
        if self._cem:
            np.ctypeslib.as_array(o.sigma_new, shape=(self._P,))[...] = np.float32(0.5 + 0.25 * n)
        return 0


def make_inputs(robot, gait, tick, horizon=12):
    """The dicts the WB interface hands to SRBDControllerInterface.compute_control (wb_interface.py:262-330 shapes):
    reference footholds are (1, 3) rows, the contact table is (4, H)."""
    from quadruped_pympc_b200 import synthetic
    sc = synthetic.scenario(robot, gait, horizon=horizon, seed=3, sim_steps=25 * tick)   # the gait advances between ticks
    s, r = sc['state'].copy(), sc['reference'].copy()
    s[0:2] += 0.01 * tick
    r[12:24] += 0.02                                            # reference footholds differ from the current feet
    state_current = dict(position=s[0:3], linear_velocity=s[3:6], orientation=s[6:9], angular_velocity=s[9:12],
                         foot_FL=s[12:15], foot_FR=s[15:18], foot_RL=s[18:21], foot_RR=s[21:24])
    ref_state = dict(ref_position=r[0:3], ref_linear_velocity=r[3:6], ref_orientation=r[6:9], ref_angular_velocity=r[9:12],
                     ref_foot_FL=r[12:15].reshape(1, 3), ref_foot_FR=r[15:18].reshape(1, 3),
                     ref_foot_RL=r[18:21].reshape(1, 3), ref_foot_RR=r[21:24].reshape(1, 3))
    return state_current, ref_state, sc['contact_sequence'], sc['phase_signal'], sc['nominal_step_freq']


CASES = [   # name, robot, gait, mpc_params overrides
    ('mppi', 'go2', 'trot', dict(sampling_method='mppi', control_parametrization='cubic_spline', optimize_step_freq=False,
                                 num_sampling_iterations=1)),
    ('random_sampling_adaptive', 'aliengo', 'trot', dict(sampling_method='random_sampling', control_parametrization='linear_spline',
                                                         optimize_step_freq=True, num_sampling_iterations=1)),
    ('cem_mppi', 'go2', 'pace', dict(sampling_method='cem_mppi', control_parametrization='cubic_spline', optimize_step_freq=False,
                                     num_sampling_iterations=2)),
]
N_TRACE, TICKS = 512, 4


def run_case(name, robot, gait, overrides):
    """Drive the patched reference interface for TICKS ticks; returns (recording lib, per-tick interface outputs, inputs)."""
    import quadruped_pympc_b200 as qp
    from quadruped_pympc_b200 import _lib
    from quadruped_pympc_b200.config import ROBOT_INERTIALS
    config = install_stubs()
    mass, inertia = ROBOT_INERTIALS[robot]
    config.robot, config.mass, config.inertia = robot, mass, np.array(inertia)
    mp = config.mpc_params
    mp.update(type='sampling', num_parallel_computations=N_TRACE, grf_max=mass * 9.81, grf_min=0, num_splines=2, horizon=12,
              dt=0.02, use_nonuniform_discretization=False, shift_solution=False, sigma_cem_mppi=3, device='gpu')
    mp.update(overrides)
    cfg_line, _ = integration_patch()
    key, val = re.match(r"\s*'(\w+)':\s*'(\w+)'", cfg_line).groups()     # 'sampling_backend': 'b200'
    mp[key] = val
    real = _lib.load()
    rec = RecordingLib(real)
    _lib._lib = rec                                   # what _lib.load() hands out from now on
    fast_saved, _lib._fast = _lib._fast, False        # the ctypes binding is the one a stand-in library can observe (the
    try:                                              # CPython binding of the same call is compared with it on the GPU)
        mod = patched_interface_module()
        iface = mod.SRBDControllerInterface()
        assert type(iface.controller) is qp.Sampling_MPC, 'the patched interface did not construct the B200 controller'
        outs, inputs = [], []
        for tick in range(TICKS):
            sc, rs, cs, ph, f = make_inputs(robot, gait, tick)
            inputs.append((sc, rs, cs, ph, f))
            outs.append(iface.compute_control(sc, rs, cs, np.eye(3), ph, f, 1))
        iface.controller.close()
    finally:
        _lib._lib = real
        _lib._fast = fast_saved
        mp.pop(key, None)
    return rec, outs, inputs
