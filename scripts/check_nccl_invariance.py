"""Run under torchrun (one rank per GPU): the NCCL slab path must reproduce the single-GPU run bit for bit.
Each rank advances its slab of a small grid with halo exchange + max-allreduce, then advances the WHOLE grid
alone on its own GPU and compares its slab's columns.  Prints 'NCCL_INVARIANCE_OK' on rank 0 if all ranks agree."""
import ctypes
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lsc_internship_b200 as lsc  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
L = lsc.lib()
ok = True
for prob, nx, ny, dom, bcx, bcy in [("ot", 192, 128, (0, 1, 0, 1), "periodic", "periodic"),
                                     ("kh", 128, 160, (0, 1, -1, 1), "transmissive", "reflective")]:
    kw = dict(nxCells=nx, nyCells=ny, x0=dom[0], x1=dom[1], y0=dom[2], y1=dom[3], bc_x=bcx, bc_y=bcy, problem=prob)
    s = lsc.MusclNew(device=local, rank=rank, nranks=world, **kw)
    idbuf = ctypes.create_string_buffer(128)
    if rank == 0:
        lsc._capi.check(L.lsc_comm_unique_id(ctypes.cast(idbuf, ctypes.c_void_p)))
    t = torch.tensor(list(idbuf.raw), dtype=torch.uint8, device="cuda")
    dist.broadcast(t, 0)
    s.comm_init(bytes(t.cpu().tolist()))
    s.init_problem()
    s.step(9)
    mine = s.download()
    sc = s.scalars()
    one = lsc.MusclNew(device=local, **kw)
    one.init_problem()
    one.step(9)
    full = one.download()
    same = np.array_equal(mine[2:-2, 2:-2].view(np.uint64), full[s.i0 + 2:s.i0 + 2 + s.nxl, 2:-2].view(np.uint64)) \
        and sc == one.scalars()
    if not same:
        a, b = mine[2:-2, 2:-2], full[s.i0 + 2:s.i0 + 2 + s.nxl, 2:-2]
        bad = np.argwhere(a.view(np.uint64) != b.view(np.uint64))
        print(f"[rank {rank}] {prob} {bcx}/{bcy}: scalars {sc} vs {one.scalars()}; {len(bad)} differing values, "
              f"first {bad[:5].tolist()}, columns {sorted(set(bad[:, 0].tolist()))[:12]}", flush=True)
    ok = ok and same
    s.close()
    one.close()
# the Euler SLIC slabs: same protocol with 4 variables
for prob, nx, ny, bc in [("ot", 160, 96, "periodic"), ("cylinder", 144, 120, "transmissive")]:
    kw = dict(nxCells=nx, nyCells=ny, bc=bc, problem=prob)
    s = lsc.SlicEuler2d(device=local, rank=rank, nranks=world, **kw)
    idbuf = ctypes.create_string_buffer(128)
    if rank == 0:
        lsc._capi.check(L.lsc_comm_unique_id(ctypes.cast(idbuf, ctypes.c_void_p)))
    t = torch.tensor(list(idbuf.raw), dtype=torch.uint8, device="cuda")
    dist.broadcast(t, 0)
    s.comm_init(bytes(t.cpu().tolist()))
    s.init_problem()
    s.step(9)
    mine = s.download()
    sc = s.scalars()
    one = lsc.SlicEuler2d(device=local, **kw)
    one.init_problem()
    one.step(9)
    full = one.download()
    same = np.array_equal(mine[2:-2, 2:-2].view(np.uint64), full[s.i0 + 2:s.i0 + 2 + s.nxl, 2:-2].view(np.uint64)) \
        and sc == one.scalars()
    if not same:
        print(f"[rank {rank}] euler {prob} {bc}: scalars {sc} vs {one.scalars()}", flush=True)
    ok = ok and same
    s.close()
    one.close()
flag = torch.tensor([1 if ok else 0], device="cuda")
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    print("NCCL_INVARIANCE_OK" if int(flag.item()) == 1 else "NCCL_INVARIANCE_FAILED")
dist.destroy_process_group()
sys.exit(0 if int(flag.item()) == 1 else 1)
