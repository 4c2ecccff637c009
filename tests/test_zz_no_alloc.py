"""lsc_mhd_state_hash_noalloc: the state fingerprint without any device allocation on the call path (scratch in a
file-scope __device__ array of state_hash.cu).  Not part of the reference: added for callers that cannot call
cudaMalloc.  Collected last on purpose."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def both_hashes(lsc, s):
    L = lsc.lib()
    ref, new = ctypes.c_ulonglong(), ctypes.c_ulonglong(0xDEADBEEF)
    assert L.lsc_mhd_state_hash(s._h, ctypes.byref(ref)) == 0
    assert L.lsc_mhd_state_hash_noalloc(s._h, ctypes.byref(new)) == 0, lsc.lib().lsc_last_error()
    return ref.value, new.value


def test_noalloc_hash_equals_the_allocating_entry_point():
    """Integer result: equality, not a tolerance.  Two calls in a row on one state, then on changed states and on
    contexts of other shapes -- the shared scratch keeps the previous call's partial sums (2304 of them after the
    large context, 2 for the small one that follows) and none of that may leak into the next result."""
    import lsc_internship_b200 as lsc
    from lsc_internship_b200._capi import state_hash_host
    s = lsc.MusclNew(200, 136, 0, 1, 0, 1, bc_x="periodic", bc_y="periodic", problem="ot")
    s.init_problem()
    s.step(3)
    ref, new = both_hashes(lsc, s)
    assert new == ref
    ref2, new2 = both_hashes(lsc, s)                   # second call in a row, same state
    assert (ref2, new2) == (ref, ref)
    assert ref == state_hash_host(s.download()[2:-2, 2:-2])
    s.step(2)
    ref3, new3 = both_hashes(lsc, s)                   # changed state
    assert new3 == ref3 and ref3 != ref
    big = lsc.MusclNew(1500, 1100, 0, 1, 0, 1, bc_x="periodic", bc_y="periodic", problem="ot")
    big.init_problem()
    rb, nb = both_hashes(lsc, big)
    assert nb == rb
    big.close()
    ref4, new4 = both_hashes(lsc, s)                   # the small context again, after the large one
    assert (ref4, new4) == (ref3, ref3)
    assert s.state_hash_noalloc() == ref3              # the Python mirror's method
    s.close()


def test_noalloc_hash_of_a_slab_carries_the_global_column_offset():
    import lsc_internship_b200 as lsc
    nx, ny = 96, 70
    whole = lsc.MusclNew(nx, ny, 0, 1, 0, 1, bc_x="periodic", bc_y="periodic", problem="ot")
    whole.init_problem()
    u0 = whole.download()
    want = whole.state_hash()
    assert whole.state_hash_noalloc() == want
    whole.close()
    total = 0
    for b0, b1 in ((0, 40), (40, 96)):
        part = lsc.MusclNew(nx, ny, 0, 1, 0, 1, bc_x="periodic", bc_y="periodic", problem="ot",
                            rank=0 if b0 == 0 else 1, nranks=2, slab=(b0, b1 - b0))
        assert (part.i0, part.nxl) == (b0, b1 - b0)
        part.upload(np.ascontiguousarray(u0[part.i0:part.i0 + part.nxl + 4]))
        r, n = both_hashes(lsc, part)
        assert n == r
        total += n
        part.close()
    assert total & ((1 << 64) - 1) == want


def test_noalloc_hash_refuses_a_slab_taller_than_its_scratch():
    import lsc_internship_b200 as lsc
    from lsc_internship_b200._capi import LscError
    s = lsc.MusclNew(4, 1048576 + 1, 0, 1, 0, 1, bc_x="periodic", bc_y="periodic", problem="ot")
    h = ctypes.c_ulonglong()
    rc = lsc.lib().lsc_mhd_state_hash_noalloc(s._h, ctypes.byref(h))
    assert rc == -1                                    # LSC_EINVAL, before anything is launched
    msg = lsc.lib().lsc_last_error().decode()
    assert "1048577" in msg and "1048576" in msg
    with pytest.raises(LscError):
        s.state_hash_noalloc()
    s.close()
