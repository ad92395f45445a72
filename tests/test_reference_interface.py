"""CPU: the four-line patch of INTEGRATION.md, executed.  The reference's own SRBDControllerInterface (source read from
$QPMPC_REFERENCE or /root/reference, the INTEGRATION.md block substituted for srbd_controller_interface.py:77-83 in
memory) drives quadruped_pympc_b200.Sampling_MPC; a recording stand-in for the C library shows what crosses the ABI.
Skipped (not failed) when the reference tree is absent -- e.g. on the GPU box; the recorded trace is committed as
tests/golden/interface_trace.npz and replayed there against the oracle (tests/test_gpu_interface_loop.py)."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import interface_harness as ih

pytestmark = pytest.mark.skipif(not ih.reference_available(), reason='reference tree not present')

FOOT = ('FL', 'FR', 'RL', 'RR')


@pytest.mark.parametrize('case', ih.CASES, ids=[c[0] for c in ih.CASES])
def test_patched_reference_interface_drives_the_b200_controller(built, case):
    name, robot, gait, overrides = case
    rec, outs, inputs = ih.run_case(*case)
    iters = overrides['num_sampling_iterations']
    cem = overrides['sampling_method'] == 'cem_mppi'
    # ---- one handle, configured from the reference's config module -------------------------------------------------
    assert len(rec.creates) == 1
    c = rec.creates[0]
    assert (c['num_samples'], c['horizon'], c['method'], c['adaptive']) == \
           (ih.N_TRACE, 12, {'random_sampling': 0, 'mppi': 1, 'cem_mppi': 2}[overrides['sampling_method']],
            int(overrides['optimize_step_freq']))
    assert c['parametrization'] == {'linear_spline': 1, 'cubic_spline': 2}[overrides['control_parametrization']]
    assert len(rec.ticks) == ih.TICKS * iters
    prev_contact = np.array([1, 1, 1, 1])
    carried_best, n = np.zeros_like(rec.ticks[0]['best']), 0
    for tick, (sc, rs, cs, ph, f) in enumerate(inputs):
        current = cs[:, 0]
        # prepare_state_and_reference (I:120-122 -> S:563-627): swing feet <- reference footholds
        want_state = np.concatenate([sc['position'], sc['linear_velocity'], sc['orientation'], sc['angular_velocity']] +
                                    [(rs['ref_foot_' + l].reshape(3) if current[k] == 0 else sc['foot_' + l]) for k, l in enumerate(FOOT)])
        want_ref = np.concatenate([rs['ref_position'], rs['ref_linear_velocity'], rs['ref_orientation'], rs['ref_angular_velocity']] +
                                  [rs['ref_foot_' + l].reshape(3) for l in FOOT])
        # knots of a leg that just lifted off are zeroed (S:612-625)
        L = carried_best.size // 4
        for leg in range(4):
            if prev_contact[leg] == 1 and current[leg] == 0:
                carried_best[leg * L:(leg + 1) * L] = 0.0
        prev_contact = current
        for it in range(iters):
            t = rec.ticks[n]
            n += 1
            np.testing.assert_array_equal(t['state'], want_state.astype(np.float32))
            np.testing.assert_array_equal(t['reference'], want_ref.astype(np.float32))
            np.testing.assert_array_equal(t['contact_sequence'], cs[:, :12].astype(np.float32))
            np.testing.assert_array_equal(t['key'], np.array([42, n], np.uint32))        # with_newkey() per iteration (I:126)
            np.testing.assert_array_equal(t['best'], carried_best)                        # I:135/155: best carried
            assert not t['noise_injected'] and not t['freqs_injected']
            if cem:                                                                       # I:128-129, 148
                want_sigma = np.float32(3.0) if it == 0 else np.float32(0.5 + 0.25 * (n - 1))
                np.testing.assert_array_equal(t['sigma'], np.full_like(t['best'], want_sigma))
            else:                                                                         # I:159-168
                assert t['sigma'] is None
                np.testing.assert_allclose(t['phase_signal'], ph.astype(np.float32))
                assert t['nominal_step_freq'] == np.float32(f) and t['optimize_swing'] == 1
            carried_best = t['best'] + np.float32(n)                                      # what the stand-in returned
        # ---- what the interface returns (I:170-180, 225-240) built from the LAST iteration's outputs ---------------
        grfs, footholds, jp, jv, ja, best_freq, pred = outs[tick]
        fake = np.array([10.0 * n + k for k in range(12)], np.float32)
        for k, l in enumerate(FOOT):
            np.testing.assert_array_equal(getattr(grfs, l), fake[3 * k:3 * k + 3] * current[k])
            np.testing.assert_array_equal(getattr(footholds, l), rs['ref_foot_' + l][0])
        assert jp is None and jv is None and ja is None
        assert best_freq == np.float32(1.4 + 0.1 * n)
        np.testing.assert_array_equal(pred, np.array([100.0 * n + k for k in range(24)], np.float32))


def test_committed_trace_is_what_the_reference_interface_produces(built):
    """tests/golden/interface_trace.npz (replayed on the GPU) == a fresh recording."""
    sys.path.insert(0, os.path.join(ih.ROOT, 'tools'))
    import make_interface_trace as mt
    fresh = mt.trace_arrays()
    z = np.load(os.path.join(ih.ROOT, 'tests', 'golden', 'interface_trace.npz'))
    assert sorted(z.files) == sorted(fresh)
    for k in z.files:
        np.testing.assert_array_equal(z[k], fresh[k], err_msg=k)
