"""CPU tests of the C-ABI library: it loads, exports every symbol include/lsc_b200.h declares,
and refuses to run without a GPU (no CPU fallback)."""
import ctypes
import os
import re

import pytest

import lsc_internship_b200 as lsc
from lsc_internship_b200 import _capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "lsc_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(lsc_[a-z0-9_]+)\s*\(", src)) - {"lsc_mhd_dump_cb", "lsc_mhd_log_cb"})


def test_library_exports_every_declared_symbol():
    L = lsc.lib()
    syms = header_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(L, s), f"{s} declared in include/lsc_b200.h but not exported"
    assert set(syms) == set(_capi.EXPORTS)


def test_config_struct_layout():
    assert ctypes.sizeof(_capi.MhdConfig) == 4 + 4 + 6 * 8 + 7 * 4 + 4  # trailing pad to 8


@pytest.mark.parametrize("bc", ["periodic", "transmissive"])
@pytest.mark.parametrize("problem", [0, 1, 2])
def test_euler_slab_initial_condition_is_the_global_one(bc, problem):
    """lsc_euler_initial_condition for rank r of P (host code, no GPU): every slab equals the corresponding
    columns of the whole-grid array -- interior columns and the halo columns that are a neighbour's interior --
    and at a periodic wrap / physical edge holds what the reference's boundary fill (S:596) puts there."""
    import numpy as np
    L = lsc.lib()
    nx, ny, P = 37, 21, 3
    bcv = {"periodic": _capi.BC_PERIODIC, "transmissive": _capi.BC_TRANSMISSIVE_E}[bc]
    dp = ctypes.POINTER(ctypes.c_double)
    whole_cfg = _capi.EulerConfig(nx, ny, 0.0, 1.0, 0.0, 1.0, 1.4, 0.8, 0.0, bcv, 0, 0, 1)
    whole = np.empty((nx + 4, ny + 4, 4))
    assert L.lsc_euler_initial_condition(ctypes.byref(whole_cfg), problem, whole.ctypes.data_as(dp)) == 0
    covered = 0
    for r in range(P):
        b0, b1 = nx * r // P, nx * (r + 1) // P
        cfg = _capi.EulerConfig(nx, ny, 0.0, 1.0, 0.0, 1.0, 1.4, 0.8, 0.0, bcv, 0, r, P)
        slab = np.empty((b1 - b0 + 4, ny + 4, 4))
        assert L.lsc_euler_initial_condition(ctypes.byref(cfg), problem, slab.ctypes.data_as(dp)) == 0
        assert np.array_equal(slab.view(np.uint64), whole[b0:b1 + 4].view(np.uint64)), f"rank {r}"
        covered += b1 - b0
    assert covered == nx


def test_no_cpu_fallback_or_bad_config():
    L = lsc.lib()
    if L.lsc_device_count() == 0:
        with pytest.raises(lsc.LscError) as e:
            lsc.MusclNew(nxCells=16, nyCells=16)
        assert e.value.code == -2
    else:
        with pytest.raises(lsc.LscError) as e:
            lsc.MusclNew(nxCells=2, nyCells=16)
        assert e.value.code == -1


def test_product_package_does_not_touch_oracle():
    """The product path must not import, link or execute anything under oracle/."""
    pkg = os.path.join(ROOT, "lsc_internship_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")) or f == "Makefile":
                txt = open(os.path.join(dp, f), errors="replace").read()
                assert not re.search(r"liblsc_oracle|oracle_util|oracle/|lsc_oracle_|import oracle", txt), \
                    f"{f} reaches into oracle/"
