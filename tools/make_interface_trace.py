"""Record what crosses the C ABI when the REFERENCE'S SRBDControllerInterface (with the INTEGRATION.md patch applied)
drives quadruped_pympc_b200.Sampling_MPC, and freeze it as tests/golden/interface_trace.npz.  Needs /root/reference (or
$QPMPC_REFERENCE); the fixture is replayed on the GPU against the oracle by tests/test_gpu_interface_loop.py.
    python tools/make_interface_trace.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))
import interface_harness as ih


def trace_arrays():
    out = {}
    for case in ih.CASES:
        name = case[0]
        rec, outs, inputs = ih.run_case(*case)
        c = rec.creates[0]
        out[f'{name}/config'] = np.array([c['num_samples'], c['horizon'], c['num_splines'], c['parametrization'], c['method'],
                                          c['adaptive'], c['seed']], np.int64)
        out[f'{name}/mass_mu_grf'] = np.array([c['mass'], c['mu'], c['grf_min'], c['grf_max']], np.float64)
        for k in ('state', 'reference', 'contact_sequence', 'best', 'key'):
            out[f'{name}/{k}'] = np.stack([t[k] for t in rec.ticks])
        out[f'{name}/phase_signal'] = np.stack([t['phase_signal'] if t['phase_signal'] is not None else np.zeros(4, np.float32)
                                                for t in rec.ticks])
        out[f'{name}/sigma'] = np.stack([t['sigma'] if t['sigma'] is not None else np.zeros_like(t['best']) for t in rec.ticks])
        out[f'{name}/nominal_step_freq'] = np.array([t['nominal_step_freq'] for t in rec.ticks], np.float32)
        out[f'{name}/optimize_swing'] = np.array([t['optimize_swing'] for t in rec.ticks], np.int32)
    return out


if __name__ == '__main__':
    if not ih.reference_available():
        sys.exit(f'reference tree not found at {ih.REFERENCE}')
    path = os.path.join(ROOT, 'tests', 'golden', 'interface_trace.npz')
    np.savez_compressed(path, **trace_arrays())
    print('wrote', path, os.path.getsize(path), 'bytes')
