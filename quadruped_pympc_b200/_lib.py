"""ctypes binding of libqpmpc_b200.so (the C ABI in include/qpmpc.h) + the nvcc build recipe.

Product code.  There is deliberately no fallback: if the shared library is missing or no CUDA device is
present, every compute call raises (QpmpcError) instead of silently running somewhere else.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(_HERE)
CSRC = os.path.join(_HERE, 'csrc')
LIB_DIR = os.path.join(_HERE, 'lib')
LIB_PATH = os.path.join(LIB_DIR, 'libqpmpc_b200.so')
HEADER = os.path.join(ROOT, 'include', 'qpmpc.h')

PEER_BLOB_BYTES = 128
QPMPC_ABI_VERSION = 2
MAX_HORIZON, MAX_SPLINES, MAX_ELITE = 64, 16, 256
PARAMETRIZATION = {'zero_order': 0, 'linear_spline': 1, 'cubic_spline': 2}
METHOD = {'random_sampling': 0, 'mppi': 1, 'cem_mppi': 2}
IO_MODE = {'memcpy': 0, 'mapped': 1, 'direct': 2}
OK, ERR_INVALID, ERR_CUDA, ERR_NCCL, ERR_STATE = 0, -1, -2, -3, -4


class QpmpcError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f'qpmpc error {code}: {msg}')
        self.code = code


class QpmpcConfig(C.Structure):
    _fields_ = [
        ('abi_version', C.c_int32), ('num_samples', C.c_int32), ('horizon', C.c_int32), ('num_splines', C.c_int32),
        ('parametrization', C.c_int32), ('method', C.c_int32), ('adaptive', C.c_int32), ('num_robots', C.c_int32),
        ('device', C.c_int32), ('io_mode', C.c_int32), ('use_graph', C.c_int32), ('shard_rank', C.c_int32),
        ('shard_world', C.c_int32), ('elite_k', C.c_int32), ('seed', C.c_uint32), ('mass', C.c_float),
        ('inertia', C.c_float * 9), ('dts', C.c_float * MAX_HORIZON), ('mu', C.c_float), ('grf_min', C.c_float),
        ('grf_max', C.c_float), ('q_diag', C.c_float * 12), ('sigma_random_sampling', C.c_float * 3),
        ('sigma_mppi', C.c_float), ('step_freq_available', C.c_float * 3), ('pgg_duty_factor', C.c_float),
        ('robot_base', C.c_int32), ('lanes_per_sample', C.c_int32),
    ]


_FP = C.POINTER(C.c_float)


class QpmpcTickIn(C.Structure):
    _fields_ = [
        ('state', _FP), ('reference', _FP), ('contact_sequence', _FP), ('contact_ld', C.c_int32), ('best', _FP),
        ('sigma', _FP), ('phase_signal', _FP), ('nominal_step_freq', C.c_float), ('optimize_swing', C.c_int32),
        ('key', C.c_uint32 * 2), ('noise', _FP), ('step_freqs', _FP),
    ]


class QpmpcTickOut(C.Structure):
    _fields_ = [
        ('grfs', C.c_float * 12), ('predicted_state', C.c_float * 24), ('best_cost', C.c_float),
        ('best_freq', C.c_float), ('best_index', C.c_int32), ('num_elite', C.c_int32), ('best', _FP),
        ('sigma_new', _FP), ('elite_indices', C.POINTER(C.c_int32)),
    ]


# every symbol include/qpmpc.h declares (tests check the header and this list against the built library)
EXPORTS = [
    'qpmpc_abi_version', 'qpmpc_last_error', 'qpmpc_device_count', 'qpmpc_config_defaults', 'qpmpc_create',
    'qpmpc_destroy', 'qpmpc_num_parameters', 'qpmpc_num_local_samples', 'qpmpc_tick', 'qpmpc_tick_iterations', 'qpmpc_step', 'qpmpc_step_flat', 'qpmpc_step_reset',
    'qpmpc_step_state', 'qpmpc_tick_batched',
    'qpmpc_tick_batched_flat',
    'qpmpc_pgg_contact_sequence', 'qpmpc_footholds_reference', 'qpmpc_step_set_shift', 'qpmpc_kernel_name', 'qpmpc_debug_host_times', 'qpmpc_shm_pack', 'qpmpc_publish_shm', 'qpmpc_stance_torques', 'qpmpc_stance_torques_static', 'qpmpc_get_costs', 'qpmpc_rollout', 'qpmpc_generate_noise', 'qpmpc_generate_step_freqs',
    'qpmpc_shard_record_floats', 'qpmpc_shard_partial', 'qpmpc_shard_record_device_ptr', 'qpmpc_shard_finish',
    'qpmpc_shard_copy_record', 'qpmpc_shard_get_record', 'qpmpc_nccl_unique_id', 'qpmpc_shard_init_nccl',
    'qpmpc_shard_peer_export', 'qpmpc_shard_peer_attach', 'qpmpc_shard_peer_loopback', 'qpmpc_shard_peer_inject',
    'qpmpc_tick_sharded', 'qpmpc_debug_peer_wait', 'qpmpc_shard_peer_timeout', 'qpmpc_upload_resident', 'qpmpc_bench_resident', 'qpmpc_bench_resident_wall', 'qpmpc_fetch_resident_out',
    'qpmpc_kernel_launch_count', 'qpmpc_measure_fp32_peak', 'qpmpc_debug_phase_times',
]

NVCC_COMPILE_FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo', '-O3', '-std=c++17',
                      '-Xcompiler', '-fPIC,-ffp-contract=off', '-diag-suppress', '177,550']


def build(force: bool = False, verbose: bool = False) -> str:
    """nvcc-compile the CUDA library in-tree for sm_100a (cross-compiles without a GPU).  The translation units
    (host API + six kernel-instantiation units) are compiled in parallel, then linked into one shared library."""
    files = sorted(os.listdir(CSRC))
    srcs = [os.path.join(CSRC, f) for f in files]
    newest = max(os.path.getmtime(p) for p in srcs + [HEADER] if not p.endswith('qpfast.c'))
    if not force and os.path.exists(LIB_PATH) and os.path.getmtime(LIB_PATH) >= newest:
        if not os.path.exists(fast_path()) or os.path.getmtime(fast_path()) < os.path.getmtime(os.path.join(CSRC, 'qpfast.c')):
            build_fast()
        return LIB_PATH
    os.makedirs(LIB_DIR, exist_ok=True)
    obj_dir = os.path.join(LIB_DIR, 'obj')
    os.makedirs(obj_dir, exist_ok=True)
    units = [f for f in files if f.endswith('.cu')]
    procs = []
    for u in units:
        obj = os.path.join(obj_dir, u[:-3] + '.o')
        cmd = ['nvcc'] + NVCC_COMPILE_FLAGS + os.environ.get('QPMPC_NVCC_EXTRA', '').split() + \
              (['-Xptxas', '-v'] if verbose else []) + \
              ['-c', os.path.join(CSRC, u), '-o', obj]
        procs.append((u, obj, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    objs = []
    for u, obj, p in procs:
        out, _ = p.communicate()
        if verbose or p.returncode != 0:
            print(out)
        if p.returncode != 0:
            raise subprocess.CalledProcessError(p.returncode, f'nvcc -c {u}')
        objs.append(obj)
    subprocess.check_call(['nvcc', '-gencode', 'arch=compute_100a,code=sm_100a', '-shared', '-o', LIB_PATH] + objs + ['-ldl'])
    build_fast()
    return LIB_PATH


def fast_path():
    import sysconfig
    return os.path.join(LIB_DIR, '_qpfast' + (sysconfig.get_config_var('EXT_SUFFIX') or '.so'))


def build_fast() -> str:
    """gcc-compile the CPython binding of the tick call (csrc/qpfast.c: marshalling only, links libqpmpc_b200.so)."""
    import sysconfig
    import numpy
    out = fast_path()
    subprocess.check_call(['gcc', '-O2', '-fPIC', '-shared', '-I' + sysconfig.get_paths()['include'], '-I' + numpy.get_include(),
                           os.path.join(CSRC, 'qpfast.c'), '-o', out, '-L' + LIB_DIR, '-lqpmpc_b200', '-Wl,-rpath,$ORIGIN'])
    return out


_fast = None


def load_fast():
    """The CPython binding of qpmpc_tick / qpmpc_step (lib/_qpfast*.so), or None when it has not been built for this
    interpreter (the ctypes path is then used: same C ABI, ~8 us more Python per tick)."""
    global _fast
    if _fast is None:
        _fast = False
        p = fast_path()
        if os.path.exists(p):
            load()                                          # libqpmpc_b200.so first (the module links against it)
            import importlib.util
            spec = importlib.util.spec_from_file_location('_qpfast', p)
            try:
                mod = importlib.util.module_from_spec(spec)
                spec.loader.exec_module(mod)
                _fast = mod
            except ImportError:
                _fast = False
    return _fast or None


_lib = None


def load():
    """dlopen libqpmpc_b200.so and declare the prototypes.  Raises if the library has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise QpmpcError(ERR_CUDA, f'{LIB_PATH} is missing: run `python -c "import __graft_entry__ as g; g.build()"` '
                                   f'(there is no CPU fallback for the sampling path)')
    lib = C.CDLL(LIB_PATH)
    H = C.c_void_p
    lib.qpmpc_abi_version.restype = C.c_int
    lib.qpmpc_last_error.restype = C.c_char_p
    lib.qpmpc_device_count.argtypes = [C.POINTER(C.c_int)]
    lib.qpmpc_config_defaults.argtypes = [C.POINTER(QpmpcConfig)]
    lib.qpmpc_config_defaults.restype = None
    lib.qpmpc_create.argtypes = [C.POINTER(QpmpcConfig), C.POINTER(H)]
    lib.qpmpc_destroy.argtypes = [H]
    lib.qpmpc_destroy.restype = None
    lib.qpmpc_num_parameters.argtypes = [H]
    lib.qpmpc_num_local_samples.argtypes = [H]
    lib.qpmpc_tick.argtypes = [H, C.POINTER(QpmpcTickIn), C.POINTER(QpmpcTickOut)]
    lib.qpmpc_tick_iterations.argtypes = [H, C.POINTER(QpmpcTickIn), C.c_int32, C.POINTER(QpmpcTickOut)]
    lib.qpmpc_step.argtypes = [H, C.POINTER(QpmpcTickIn), C.c_int32, C.c_float, C.POINTER(QpmpcTickOut)]
    _IP, _UP = C.POINTER(C.c_int32), C.POINTER(C.c_uint32)
    lib.qpmpc_step_flat.argtypes = [H, _FP, _FP, _FP, C.c_int32, _FP, _FP, _IP, C.c_int32, C.c_float,
                                    _FP, _FP, _FP, _FP, _FP, _IP, _FP]
    lib.qpmpc_step_reset.argtypes = [H, C.POINTER(C.c_uint32), _FP, C.POINTER(C.c_int32)]
    lib.qpmpc_step_state.argtypes = [H, C.c_int, _FP, _FP, C.POINTER(C.c_int32), C.POINTER(C.c_uint32)]
    lib.qpmpc_tick_batched.argtypes = [H, C.POINTER(QpmpcTickIn), C.POINTER(QpmpcTickOut)]
    lib.qpmpc_tick_batched_flat.argtypes = [H, _FP, _FP, _FP, C.c_int32, _FP, _FP, _FP, _FP, _IP, _UP,
                                            _FP, _FP, _FP, _FP, _FP, _IP, _FP]
    _DP = C.POINTER(C.c_double)
    lib.qpmpc_step_set_shift.argtypes = [H, C.c_float]
    lib.qpmpc_debug_host_times.argtypes = [H, C.c_int32, C.POINTER(C.c_uint64)]
    lib.qpmpc_kernel_name.argtypes = [H]
    lib.qpmpc_kernel_name.restype = C.c_char_p
    lib.qpmpc_shm_pack.argtypes = [_DP, C.POINTER(C.c_uint64), _FP, _FP, _DP, _FP, C.c_double, C.c_double]
    lib.qpmpc_publish_shm.argtypes = [H, C.c_int, _DP, C.POINTER(C.c_uint64), _FP, _DP, C.c_double]
    lib.qpmpc_footholds_reference.argtypes = [H] + [_DP] * 7 + [C.c_double] * 4 + [_DP, C.c_int32, _DP]
    lib.qpmpc_pgg_contact_sequence.argtypes = [H, _DP, C.POINTER(C.c_uint8), _DP, C.c_double, _DP, _DP, _IP, C.c_int32, _FP]
    lib.qpmpc_stance_torques.argtypes = [H, _FP, _FP, _FP, _FP]
    lib.qpmpc_stance_torques_static.argtypes = [H, _FP, _FP, _FP, _FP]
    lib.qpmpc_get_costs.argtypes = [H, C.c_int, _FP]
    lib.qpmpc_rollout.argtypes = [H, _FP, _FP, _FP, C.c_int32, _FP, _FP, _FP, _FP]
    lib.qpmpc_generate_noise.argtypes = [H, C.c_int, C.POINTER(C.c_uint32), _FP, _FP]
    lib.qpmpc_generate_step_freqs.argtypes = [H, C.c_int, C.POINTER(C.c_uint32), C.c_float, C.c_int32, _FP]
    lib.qpmpc_shard_record_floats.argtypes = [H]
    lib.qpmpc_shard_partial.argtypes = [H, C.POINTER(QpmpcTickIn)]
    lib.qpmpc_shard_record_device_ptr.argtypes = [H, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]
    lib.qpmpc_shard_finish.argtypes = [H, C.POINTER(QpmpcTickOut)]
    lib.qpmpc_shard_copy_record.argtypes = [H, C.c_int, _FP]
    lib.qpmpc_shard_get_record.argtypes = [H, _FP]
    lib.qpmpc_nccl_unique_id.argtypes = [C.c_char_p, C.POINTER(C.c_ubyte)]
    lib.qpmpc_shard_init_nccl.argtypes = [H, C.c_char_p, C.POINTER(C.c_ubyte)]
    lib.qpmpc_shard_peer_export.argtypes = [H, C.POINTER(C.c_ubyte)]
    lib.qpmpc_shard_peer_attach.argtypes = [H, C.POINTER(C.c_ubyte)]
    lib.qpmpc_shard_peer_loopback.argtypes = [H]
    lib.qpmpc_shard_peer_inject.argtypes = [H, C.c_int, _FP]
    lib.qpmpc_tick_sharded.argtypes = [H, C.POINTER(QpmpcTickIn), C.POINTER(QpmpcTickOut)]
    lib.qpmpc_debug_peer_wait.argtypes = [H, C.c_int32, C.POINTER(C.c_uint64)]
    lib.qpmpc_shard_peer_timeout.argtypes = [H, C.c_double]
    lib.qpmpc_upload_resident.argtypes = [H, C.POINTER(QpmpcTickIn)]
    lib.qpmpc_bench_resident.argtypes = [H, C.c_int, C.c_int, _FP]
    lib.qpmpc_bench_resident_wall.argtypes = [H, C.c_int, C.c_int, _FP, _FP]
    lib.qpmpc_fetch_resident_out.argtypes = [H, C.POINTER(QpmpcTickOut)]
    lib.qpmpc_kernel_launch_count.argtypes = [H, C.POINTER(C.c_uint64)]
    lib.qpmpc_measure_fp32_peak.argtypes = [C.c_int, _FP]
    lib.qpmpc_debug_phase_times.argtypes = [H, C.POINTER(C.c_uint64)]
    if lib.qpmpc_abi_version() != QPMPC_ABI_VERSION:
        raise QpmpcError(ERR_INVALID, 'libqpmpc_b200.so ABI version mismatch; rebuild')
    _lib = lib
    return lib


def check(rc: int) -> None:
    if rc != OK:
        raise QpmpcError(rc, load().qpmpc_last_error().decode())


def device_count() -> int:
    n = C.c_int(0)
    check(load().qpmpc_device_count(C.byref(n)))
    return n.value


def find_libnccl() -> str:
    """Path of the torch-bundled libnccl.so.2 (nvidia-nccl-cu12 wheel); '' lets dlopen search the default paths."""
    try:
        import nvidia.nccl  # type: ignore
        base = os.path.dirname(getattr(nvidia.nccl, '__file__', None) or list(nvidia.nccl.__path__)[0])
        cand = os.path.join(base, 'lib', 'libnccl.so.2')
        if os.path.exists(cand):
            return cand
    except Exception:
        pass
    return ''
