import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box with -m gpu)')


@pytest.fixture(scope='session')
def built():
    """Build (or reuse) the native artefacts once per session."""
    import __graft_entry__ as g
    g.build()
    return True


def pytest_sessionfinish(session, exitstatus):
    """Where the knot / GRF checks sat relative to the plain north_star tolerance (helpers._assert_knots): summary on the
    terminal and, when gpurun_out/ exists or QPMPC_PARITY_LOG is set, as a file."""
    try:
        import helpers
    except Exception:
        return
    log = helpers.PARITY_LOG
    if not log:
        return
    import numpy as np
    lines = []
    for name in ('knots', 'GRFs'):
        rows = [r for r in log if r[0] == name]
        if not rows:
            continue
        ratios = np.array([r[1] for r in rows])
        fb = [r for r in rows if r[2]]
        lines.append(f'{name}: {len(rows)} vectors checked; max|err| / (1e-4|x| + 1 ulp): median {np.median(ratios):.3f}, '
                     f'p99 {np.percentile(ratios, 99):.3f}, max {ratios.max():.3f}; {len(fb)} needed the softmax-amplification '
                     f'fallback' + (f' (worst excess {max(r[3] for r in fb):.3e})' if fb else ''))
    for r in log:
        if r[2]:
            lines.append(f'  fallback: {r[0]} x{r[1]:.1f} (excess {r[3]:.2e}, held to atol {r[5]:.2e}) in {r[4]}')
    text = '\n'.join(lines)
    print('\n[parity margin] ' + text.replace('\n', '\n[parity margin] '))
    path = os.environ.get('QPMPC_PARITY_LOG')
    if path is None and os.path.isdir(os.path.join(ROOT, 'gpurun_out')):
        path = os.path.join(ROOT, 'gpurun_out', 'parity_tolerance_summary.log')
    if path:
        with open(path, 'w') as f:
            f.write(text + '\n')
