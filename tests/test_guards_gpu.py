"""Out-of-bounds writes of the kernels, checked without compute-sanitizer (closed on this GPU pool): with LSC_GUARD=1
every device buffer of a context lies between two 64 KiB guard zones (csrc/dev_guard.cu); after the parity subset of
scripts/sanitize_cases.py -- ragged grids, every boundary type, the IEEE fix kernels, the split y sweep, slab teams,
the table EoS -- not one guard byte may have changed.  Runs in a child process (the variable is read at create time)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CHILD = r"""
import os, sys
sys.path.insert(0, %r); sys.path.insert(0, os.path.join(%r, "scripts"))
import sanitize_cases as sc
from lsc_internship_b200._capi import guard_check
import lsc_internship_b200 as lsc
for name, fn in sc.CASES.items():
    fn()
# contexts that stay alive: their buffers are still watched when the guards are read
m = lsc.MusclNew(130, 70, 0, 1, 0, 1, bc_x="periodic", bc_y="periodic", problem="ot"); m.init_problem(); m.step(12)
k = lsc.MusclNew(36, 125, 0, 1, -1, 1, bc_x="transmissive", bc_y="reflective", problem="kh"); k.init_problem(); k.step(12)
e = lsc.SlicEuler2d(131, 67, bc="transmissive", problem="quadrants"); e.init_problem(); e.step(12)
bad, watched = guard_check()
assert watched >= 15 and bad == 0, (bad, watched)
print("GUARDS_OK", watched)
""" % (ROOT, ROOT)


def test_no_kernel_writes_outside_its_buffers():
    env = dict(os.environ, LSC_GUARD="1")
    r = subprocess.run([sys.executable, "-c", CHILD], env=env, capture_output=True, text=True, timeout=600)
    assert "GUARDS_OK" in r.stdout, r.stdout[-1500:] + r.stderr[-1500:]
