#!/usr/bin/env python3
"""Small parity subset for compute-sanitizer (scripts/gpu_sanitize.sh): every kernel of the product path on shapes
that exercise its hand-rolled intra-block protocols (shared-memory tile reuse after one barrier, ghost face states
moved between threads, fix lists appended with atomics, split y sweep with the ghost fill on a second stream).
Each case is still checked bit for bit against the CPU oracle, so a sanitizer run is also a parity run."""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import lsc_internship_b200 as lsc  # noqa: E402
from oracle_util import (OracleEuler, OracleMhd, euler_cfg, euler_ic, mhd_cfg, mhd_ic, tabeos_ic,  # noqa: E402
                         write_eos_table)


def bits(a):
    return np.ascontiguousarray(a).view(np.uint64)


def mhd_case(prob, nx, ny, dom, bcx, bcy, n, poke=False, overlap=None, slabs=0):
    cfg = mhd_cfg(nx, ny, dom, bcx=bcx, bcy=bcy)
    u0 = mhd_ic(cfg, prob)
    if poke:  # operands outside the fast division's window: both fix kernels run
        u0[2:4, 7:30, 3] = 1e-200
        u0[nx // 2, 5:9, 7] = -1e-180
        u0[nx + 1, 12:20, 3] = 1e-200
    o = OracleMhd(cfg, u0)
    o.step(n)
    if slabs:
        s = lsc.MusclNewGroup(slabs, nxCells=nx, nyCells=ny, x0=dom[0], x1=dom[1], y0=dom[2], y1=dom[3], bc_x=bcx,
                              bc_y=bcy, problem=prob)
    else:
        s = lsc.MusclNew(nx, ny, dom[0], dom[1], dom[2], dom[3], bc_x=bcx, bc_y=bcy, problem=prob)
        if overlap is not None:
            s.set_overlap(overlap)
    s.upload(u0)
    s.step(n)
    u = s.download()
    if poke and not slabs:
        assert s.ieee_fallback_count() >= 2
    s.close()
    assert np.array_equal(bits(u), bits(o.state()[0])), "state differs from the oracle"


def euler_case(bc, ic, nx, ny, n, table=None):
    cfg = euler_cfg(nx, ny, (0.0, 1.0, 0.0, 1.0), bc=bc)
    s = lsc.SlicEuler2d(nx, ny, bc=bc, problem=ic)
    if table is not None:
        s.set_eos_table(*table)
        u0 = tabeos_ic(cfg, table)
    else:
        u0 = euler_ic(cfg, ic)
    o = OracleEuler(cfg, u0)
    if table is not None:
        o.set_table(*table)
    s.upload(u0)
    o.step(n)
    s.step(n)
    u = s.download()
    s.close()
    assert np.array_equal(bits(u), bits(o.state()[0])), "state differs from the oracle"


KH = (0.0, 1.0, -1.0, 1.0)
OT = (0.0, 1.0, 0.0, 1.0)
CASES = {
    "mhd_ot_130x70": lambda: mhd_case("ot", 130, 70, OT, "periodic", "periodic", 3),
    "mhd_kh_36x125_trans_refl": lambda: mhd_case("kh", 36, 125, KH, "transmissive", "reflective", 3),
    "mhd_kh_33x127_per_trans": lambda: mhd_case("kh", 33, 127, KH, "periodic", "transmissive", 3),
    "mhd_fix_kernels_overlap2": lambda: mhd_case("ot", 48, 140, OT, "periodic", "periodic", 3, poke=True, overlap=2),
    "mhd_slabs3": lambda: mhd_case("kh", 70, 50, KH, "transmissive", "transmissive", 3, slabs=3),
    "euler_cylinder_200x136": lambda: euler_case("transmissive", "cylinder", 200, 136, 3),
    "euler_ot_126": lambda: euler_case("periodic", "ot", 126, 126, 2),
}


def tabeos():
    with tempfile.TemporaryDirectory() as td:
        path = os.path.join(td, "Plasma19_EoS.txt")
        write_eos_table(path)
        table = lsc.SlicEuler2d.read_eos_table(path)
    euler_case("transmissive", "tabeos", 33, 21, 2, table=table)


CASES["tabeos_33x21"] = tabeos

if __name__ == "__main__":
    names = sys.argv[1:] or list(CASES)
    for nm in names:
        CASES[nm]()
        print("OK", nm, flush=True)
    print("SANITIZE_CASES_OK", len(names))
