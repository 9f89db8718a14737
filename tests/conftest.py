import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, 'tests', 'golden')


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box with -m gpu)')


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason='no CUDA device')
    for it in items:
        if 'gpu' in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope='session', autouse=True)
def _report_boundary_flips():
    """Writes every fp32 iteration-count flip seen by helpers.compare_plist (and how many comparisons could have
    flipped) to profiles/r02_boundary_flips_{gpu,cpu}.json — north_star: boundary flips are REPORTED.  On the GPU box
    the same file also lands in gpurun_out/ so that it travels back."""
    yield
    import json
    import torch
    import helpers
    if not helpers.COMPARED:
        return
    kind = 'gpu' if torch.cuda.is_available() else 'cpu'
    rec = {'what': 'fp32 iteration-count flips against the reference golden trajectories (tests/helpers.py::compare_plist, '
                   'bound: 1 iteration or the measured spread of the reference algorithm\'s own fp32 count, tests/golden/fp32_count_spread.json); ours = ' + ('CUDA path' if kind == 'gpu' else 'CPU oracle'),
           'comparisons': len(helpers.COMPARED), 'flips': helpers.FLIPS, 'compared': helpers.COMPARED}
    for d in ('profiles', 'gpurun_out'):
        path = os.path.join(ROOT, d)
        if d == 'gpurun_out' and kind == 'cpu':
            continue
        try:
            os.makedirs(path, exist_ok=True)
            with open(os.path.join(path, f'r02_boundary_flips_{kind}.json'), 'w') as f:
                json.dump(rec, f, indent=1)
        except OSError:
            pass
