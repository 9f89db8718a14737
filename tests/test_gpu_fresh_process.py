"""smoke() in a process of its own.

Device-side state that depends on how the driver lays a module out at load time (the address of a `__device__` table, of a
static scratch array) differs from process to process: round 2 had a table lookup that assumed a 512-byte aligned module global,
held in every pytest process and failed in smoke()'s.  This test gives the suite a second, differently laid out process."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
def test_smoke_passes_in_a_fresh_process():
    env = dict(os.environ)
    env.pop('NLLS_FUSED_VARIANT', None)
    r = subprocess.run([sys.executable, '-c', 'import __graft_entry__ as g; g.smoke()'], cwd=ROOT, env=env, capture_output=True,
                       text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert 'smoke OK' in r.stdout
