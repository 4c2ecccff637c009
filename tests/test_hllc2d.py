"""First-order sibling scheme (MHD_codes/HLLC_MHD_2D.cpp = H): the CUDA path through the C ABI
(lsc_hllc2d_*) against the REAL program -- golden fixtures made by tests/golden/make_golden_hllc2d.py from
oracle/_ref/hllc2d_ref, and the program run live where oracle/_ref travelled to the box.  Bit equality throughout."""
import glob
import hashlib
import json
import os
import subprocess

import numpy as np
import pytest

from oracle_util import have_ref, run_hllc2d_ref

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
GOLDEN = sorted(p for p in glob.glob(os.path.join(HERE, "golden", "hllc2d_*.npz")) if "program" not in p)
DOM = (0.0, 1.0, 0.0, 1.0)
gpu = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a).view(np.uint64)


def test_golden_present():
    assert len(GOLDEN) >= 5 and os.path.exists(os.path.join(HERE, "golden", "hllc2d_program_64.npz"))


def test_initial_condition_is_periodic_orszag_tang():
    """lsc_hllc2d_initial_condition is host code (no GPU): H's cell centres x0 + (i - 0.5) dx with ONE ghost layer, so
    on the unit square the ghost layer holds the periodic images up to the rounding of sin()."""
    import ctypes
    import lsc_internship_b200 as lsc
    from lsc_internship_b200._capi import Hllc2dConfig
    cfg = Hllc2dConfig(40, 24, 0.0, 1.0, 0.0, 1.0, 5.0 / 3.0, 0.75, 0)
    u = np.empty((42, 26, 9))
    assert lsc.lib().lsc_hllc2d_initial_condition(ctypes.byref(cfg), u.ctypes.data_as(ctypes.POINTER(ctypes.c_double))) == 0
    g = 5.0 / 3.0
    assert np.all(u[..., 0] == g * g) and np.all(u[..., 3] == 0) and np.all(u[..., 8] == 0)
    assert np.allclose(u[0], u[40], atol=1e-12) and np.allclose(u[:, 25], u[:, 1], atol=1e-12)
    y = 0.0 + (np.arange(26) - 0.5) * ((1.0 - 0.0) / 24)
    assert np.array_equal(u[7, :, 5], -np.sin(2.0 * 3.14159265358979323846 * y))


@gpu
@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[7:-4] for p in GOLDEN])
def test_cuda_matches_reference_golden(path):
    import lsc_internship_b200 as lsc
    z = np.load(path)
    m = json.loads(str(z["meta"]))
    s = lsc.HllcMhd2d(m["nx"], m["ny"], *m["dom"], C=m["cfl"])
    s.init_problem()
    log = []
    for _ in range(m["nsteps"]):
        s.step(1)
        log.append(s.scalars()[:3])
    u = s.download()
    s.close()
    assert np.array_equal(bits(np.array(log)), bits(z["log"])), "per-step (t, dt, ch) differs from the reference"
    assert np.array_equal(bits(u[1:-1, 1:-1]), bits(z["u"][1:-1, 1:-1])), "interior differs"
    assert np.array_equal(bits(u), bits(z["u"])), "ghost layer differs"


@gpu
@pytest.mark.skipif(not have_ref("hllc2d_ref"), reason="oracle/_ref/hllc2d_ref not built")
@pytest.mark.parametrize("nx,ny,n", [(256, 200, 15), (800, 800, 3)])   # 800 x 800: the committed size of H
def test_cuda_matches_live_reference(nx, ny, n):
    import lsc_internship_b200 as lsc
    ur, logr, _, _ = run_hllc2d_ref(nx, ny, DOM, 1e9, 0.75, n)
    s = lsc.HllcMhd2d(nx, ny)
    s.upload(s.initial_condition())
    s.step(n)
    u = s.download()
    assert s.scalars()[:3] == tuple(logr[-1])
    s.close()
    for k in range(9):  # BASELINE tolerance, then every bit
        den = np.max(np.abs(ur[1:-1, 1:-1, k]))
        assert np.max(np.abs(u[1:-1, 1:-1, k] - ur[1:-1, 1:-1, k])) <= 1e-12 * max(den, 1e-300)
    assert np.array_equal(bits(u), bits(ur))


@gpu
def test_program_and_cpp_driver_are_drop_ins(tmp_path):
    """HllcMhd2d.run() and drivers/hllc_mhd_2d run like the reference program (64 x 64 to t = 0.1): the same log
    lines, a byte-identical MHD.dat."""
    import lsc_internship_b200 as lsc
    z = np.load(os.path.join(HERE, "golden", "hllc2d_program_64.npz"))
    s = lsc.HllcMhd2d(64, 64, tStop=0.1)
    s.init_problem()
    lines = []
    n = s.run(outdir=str(tmp_path), log=lambda k, t, dt, ch: lines.append((t, dt, ch)))
    s.close()
    assert n == len(z["log"]) and np.array_equal(bits(np.array(lines)), bits(z["log"]))
    assert hashlib.sha256(open(tmp_path / "MHD.dat", "rb").read()).hexdigest() == str(z["dat_sha256"])
    exe = os.path.join(ROOT, "drivers", "_bin", "hllc_mhd_2d")
    d = tmp_path / "drv"
    d.mkdir()
    r = subprocess.run([exe, "--nx", "64", "--ny", "64", "--tstop", "0.1"], cwd=d, check=True, capture_output=True, text=True)
    assert r.stdout == str(z["stdout"])
    assert hashlib.sha256(open(d / "MHD.dat", "rb").read()).hexdigest() == str(z["dat_sha256"])
