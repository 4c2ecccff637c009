"""world_size-2/3 gloo tests (CPU) of the host-side logic of the multi-GPU path: slab partition arithmetic,
neighbour plan (ring for periodic x, open ends otherwise), the halo column pattern (two edge columns each way)
and the max-reduction that makes dt identical on every rank.  The device side of the same path is covered on
a GPU by tests/test_mhd_gpu.py::test_slab_decomposition_is_bit_invariant."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from lsc_internship_b200.mhd import neighbours, slab_range
from oracle_util import mhd_cfg, mhd_ic


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, nx, ny, bcx, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        cfg = mhd_cfg(nx, ny, bcx=bcx)
        full = mhd_ic(cfg, "ot")                     # every rank can build the analytic full grid
        i0, nxl = slab_range(nx, rank, world)
        slab = full[i0:i0 + nxl + 4].copy()           # what lsc_mhd_initial_condition gives this rank
        slab[:2] = -7.0                               # poison the halos: they must come from the neighbours
        slab[-2:] = -7.0
        left, right = neighbours(rank, world, bcx)
        # the exchange of lsc_comm.cu: send first/last two interior columns, receive into the ghost columns
        reqs = []
        lo_send = torch.from_numpy(slab[2:4].copy())
        hi_send = torch.from_numpy(slab[nxl:nxl + 2].copy())
        lo_recv = torch.empty_like(lo_send)
        hi_recv = torch.empty_like(hi_send)
        if left >= 0:
            reqs += [dist.isend(lo_send, left, tag=0), dist.irecv(lo_recv, left, tag=1)]
        if right >= 0:
            reqs += [dist.isend(hi_send, right, tag=1), dist.irecv(hi_recv, right, tag=0)]
        for r in reqs:
            r.wait()
        if left >= 0:
            slab[:2] = lo_recv.numpy()
        else:                                         # transmissive image, MUSCL_NEW.cpp:617-622
            slab[:2] = slab[2:4]
        if right >= 0:
            slab[-2:] = hi_recv.numpy()
        else:
            slab[-2:] = slab[nxl:nxl + 2]
        # expected: the full grid after the reference's x boundary fill
        exp = full.copy()
        if bcx == "periodic":
            exp[0], exp[1], exp[nx + 2], exp[nx + 3] = exp[nx], exp[nx + 1], exp[2], exp[3]
        else:
            exp[0], exp[1], exp[nx + 2], exp[nx + 3] = exp[2], exp[3], exp[nx], exp[nx + 1]
        ok_halo = np.array_equal(slab, exp[i0:i0 + nxl + 4])
        # max-allreduce: identical on all ranks and equal to the global maximum
        local = torch.tensor([float(np.abs(slab[2:-2, 2:-2, 1]).max())], dtype=torch.float64)
        dist.all_reduce(local, op=dist.ReduceOp.MAX)
        ok_max = float(local) == float(np.abs(full[2:-2, 2:-2, 1]).max())
        # slabs tile the interior exactly
        widths = [torch.zeros(2, dtype=torch.int64) for _ in range(world)]
        dist.all_gather(widths, torch.tensor([i0, nxl]))
        ok_tile = widths[0][0] == 0 and all(int(widths[k][0] + widths[k][1]) == int(widths[k + 1][0])
                                            for k in range(world - 1)) and int(widths[-1].sum()) == nx
        out[rank] = bool(ok_halo and ok_max and ok_tile)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,nx,bcx", [(2, 32, "periodic"), (2, 30, "transmissive"), (3, 35, "periodic")])
def test_slab_plan_with_gloo(world, nx, bcx):
    port = _free_port()
    with mp.Manager() as m:
        out = m.dict()
        mp.spawn(_worker, args=(world, port, nx, 16, bcx, out), nprocs=world, join=True)
        assert all(out.get(r) for r in range(world)), dict(out)


def test_slab_range_matches_library_arithmetic():
    for nx in (16, 35, 16384):
        for p in (1, 2, 3, 8):
            tot = 0
            for r in range(p):
                i0, w = slab_range(nx, r, p)
                assert i0 == tot and w >= nx // p
                tot += w
            assert tot == nx
