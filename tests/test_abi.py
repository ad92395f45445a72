"""CPU: the C-ABI library builds for sm_100a, loads, exports every symbol include/qpmpc.h declares, validates
configurations, and fails loudly (no CPU fallback) when there is no CUDA device.  No compute calls here."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import quadruped_pympc_b200 as qp
from quadruped_pympc_b200 import _lib
from quadruped_pympc_b200.sampling_mpc import make_c_config

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_functions():
    src = open(os.path.join(ROOT, 'include', 'qpmpc.h')).read()
    src = re.sub(r'/\*.*?\*/', '', src, flags=re.S)
    return sorted(set(re.findall(r'\b(qpmpc_[a-z0-9_]+)\s*\(', src)))


def test_header_and_binding_list_agree():
    assert _header_functions() == sorted(_lib.EXPORTS)


def test_library_exports_every_declared_symbol(built):
    lib = _lib.load()
    for name in _header_functions():
        assert hasattr(lib, name), f'{name} declared in include/qpmpc.h but not exported'
    assert lib.qpmpc_abi_version() == _lib.QPMPC_ABI_VERSION


def test_library_is_built_for_sm_100a(built):
    import subprocess
    out = subprocess.run(['cuobjdump', '--list-elf', _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert 'sm_100a' in out, out


def test_config_struct_layout_matches_defaults(built):
    """qpmpc_config_defaults writes through the C struct; reading it back through ctypes checks the field layout."""
    lib = _lib.load()
    c = _lib.QpmpcConfig()
    lib.qpmpc_config_defaults(C.byref(c))
    assert (c.abi_version, c.num_samples, c.horizon, c.num_splines) == (2, 10000, 12, 2)
    assert c.parametrization == _lib.PARAMETRIZATION['cubic_spline'] and c.method == _lib.METHOD['random_sampling']
    assert c.seed == 42 and abs(c.mass - 15.019) < 1e-6 and abs(c.pgg_duty_factor - 0.65) < 1e-7
    assert list(c.q_diag) == [0, 0, 1500, 200, 200, 200, 500, 500, 0, 20, 20, 50]
    assert abs(c.dts[11] - 0.02) < 1e-9 and c.robot_base == 0
    np.testing.assert_allclose(list(c.step_freq_available), [1.4, 2.0, 2.4], rtol=1e-6)


def test_no_cpu_fallback(built):
    """Without a CUDA device every compute entry point must fail (QPMPC_ERR_CUDA), never run somewhere else."""
    import torch
    if torch.cuda.is_available():
        pytest.skip('a GPU is present')
    assert _lib.load().qpmpc_device_count(C.byref(C.c_int(0))) == _lib.ERR_CUDA
    ctrl = qp.Sampling_MPC(qp.SamplerConfig.for_robot('go2', num_parallel_computations=64))
    with pytest.raises(qp.QpmpcError) as e:
        ctrl.compute_control(np.zeros(24), np.zeros(24), np.ones((4, 12)), np.zeros(96), ctrl.master_key,
                             np.zeros(4), 1.4, 0)
    assert e.value.code == _lib.ERR_CUDA
    tf = C.c_float(0)
    assert _lib.load().qpmpc_measure_fp32_peak(0, C.byref(tf)) == _lib.ERR_CUDA


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, 'quadruped_pympc_b200')
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(('.py', '.cu', '.cuh', '.h')):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r'^\s*(from|import)\s+oracle\b', text, flags=re.M), f'{f} imports oracle/'
                assert 'liboracle' not in text, f'{f} links the oracle'


def test_unknown_sampling_method_exits_like_the_reference():
    """S:107-110: print + sys.exit(1)."""
    with pytest.raises(SystemExit):
        qp.Sampling_MPC(qp.SamplerConfig.for_robot('go2', sampling_method='nope'))


@pytest.mark.parametrize('param,P', [('zero_order', 144), ('linear_spline', 36), ('cubic_spline', 96)])
def test_parameter_vector_sizes(param, P):
    """S:52-93"""
    cfg = qp.SamplerConfig.for_robot('go2', control_parametrization=param)
    assert cfg.num_control_parameters == P
    assert qp.Sampling_MPC(cfg).num_control_parameters == P


def test_make_c_config_carries_the_reference_keys(built):
    cfg = qp.SamplerConfig.for_robot('aliengo', sampling_method='cem_mppi', optimize_step_freq=True,
                                     use_nonuniform_discretization=True, num_parallel_computations=50000)
    c = make_c_config(cfg, num_robots=3, shard_rank=1, shard_world=4)
    assert (c.num_samples, c.method, c.adaptive, c.num_robots, c.shard_rank, c.shard_world) == (50000, 2, 1, 3, 1, 4)
    assert abs(c.dts[0] - 0.01) < 1e-9 and abs(c.dts[1] - 0.01) < 1e-9 and abs(c.dts[2] - 0.02) < 1e-9
    assert abs(c.grf_max - 24.637 * 9.81) < 1e-3


def test_from_reference_config_module():
    """SamplerConfig.from_reference reads the same keys Sampling_MPC.__init__ reads from quadruped_pympc.config."""
    import types
    mod = types.SimpleNamespace(mass=24.637, inertia=np.eye(3), mpc_params={
        'horizon': 12, 'dt': 0.02, 'grf_max': 24.637 * 9.81, 'grf_min': 0, 'mu': 0.5,
        'use_nonuniform_discretization': False, 'horizon_fine_grained': 2, 'dt_fine_grained': 0.01,
        'optimize_step_freq': False, 'step_freq_available': [1.4, 2.0, 2.4], 'sampling_method': 'mppi',
        'control_parametrization': 'linear_spline', 'num_splines': 2, 'num_parallel_computations': 10000,
        'num_sampling_iterations': 1, 'sigma_cem_mppi': 3, 'sigma_mppi': 3, 'sigma_random_sampling': [0.2, 3, 10],
        'shift_solution': False, 'device': 'gpu'})
    cfg = qp.SamplerConfig.from_reference(mod)
    assert cfg.sampling_method == 'mppi' and cfg.num_control_parameters == 36 and cfg.mass == 24.637


def test_robot_ids_must_fit_the_philox_counter_field(built):
    """The global robot id is bits 9..31 of Philox counter word 0: robot_base + num_robots > 2^23 would alias another
    robot's noise stream, so qpmpc_create rejects it (before it needs a device)."""
    lib = _lib.load()
    c = make_c_config(qp.SamplerConfig.for_robot('go2'), num_robots=4)
    c.robot_base = (1 << 23) - 3
    h = C.c_void_p()
    assert lib.qpmpc_create(C.byref(c), C.byref(h)) == _lib.ERR_INVALID
    assert b'2^23' in lib.qpmpc_last_error()


def test_cpython_binding_builds_and_refuses_to_tick_without_a_device(built):
    """csrc/qpfast.c (the CPython binding of the tick call) is built next to the library; without a handle it raises
    instead of computing anything."""
    fast = _lib.load_fast()
    assert fast is not None and hasattr(fast, 'tick')
    with pytest.raises(ValueError):
        fast.tick(0, 0, np.zeros(24), np.zeros(24), np.ones((4, 12)), np.zeros(96, np.float32), 42, 1, None, None, 1.4, 0, 1, 0.0, 96, 12, 0)
