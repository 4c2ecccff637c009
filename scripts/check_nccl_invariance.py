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
# uploaded random states whose ghost columns are NOT boundary images (every rank builds the same array): the first
# step has to refill / exchange them as the reference's applyBoundaryConditions does (M:838, S:604)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
rng = np.random.default_rng(4242)
for scheme, nx, ny, bcs in [("mhd", 96, 40, ("periodic", "periodic")), ("mhd", 72, 33, ("transmissive", "reflective")),
                            ("euler", 88, 37, ("periodic",)), ("euler", 64, 50, ("transmissive",))]:
    shape = (nx + 4, ny + 4)
    rho = rng.uniform(0.5, 2.0, shape)
    v = rng.uniform(-0.5, 0.5, (3,) + shape)
    B = rng.uniform(-0.5, 0.5, (3,) + shape)
    p = rng.uniform(0.8, 2.0, shape)
    if scheme == "mhd":
        u = np.zeros(shape + (9,))
        u[..., 0] = rho
        for k in range(3):
            u[..., 1 + k] = rho * v[k]
            u[..., 5 + k] = B[k]
        u[..., 4] = p / (5.0 / 3.0 - 1) + 0.5 * rho * (v ** 2).sum(0) + 0.5 * (B ** 2).sum(0)
        u[..., 8] = rng.uniform(-0.05, 0.05, shape)
        kw = dict(nxCells=nx, nyCells=ny, x0=0, x1=1, y0=0, y1=1, bc_x=bcs[0], bc_y=bcs[1], problem="ot")
        cls = lsc.MusclNew
    else:
        u = np.stack([rho, rho * v[0], rho * v[1], p / 0.4 + 0.5 * rho * (v[0] ** 2 + v[1] ** 2)], axis=-1)
        kw = dict(nxCells=nx, nyCells=ny, bc=bcs[0])
        cls = lsc.SlicEuler2d
    s = cls(device=local, rank=rank, nranks=world, **kw)
    idbuf = ctypes.create_string_buffer(128)
    if rank == 0:
        lsc._capi.check(L.lsc_comm_unique_id(ctypes.cast(idbuf, ctypes.c_void_p)))
    t = torch.tensor(list(idbuf.raw), dtype=torch.uint8, device="cuda")
    dist.broadcast(t, 0)
    s.comm_init(bytes(t.cpu().tolist()))
    s.upload(np.ascontiguousarray(u[s.i0:s.i0 + s.nxl + 4]))
    s.step(4)
    mine, sc = s.download(), s.scalars()
    one = cls(device=local, **kw)
    one.upload(u)
    one.step(4)
    full = one.download()
    same = np.array_equal(mine[2:-2, 2:-2].view(np.uint64), full[s.i0 + 2:s.i0 + 2 + s.nxl, 2:-2].view(np.uint64)) \
        and sc == one.scalars()
    if not same:
        print(f"[rank {rank}] uploaded random state, {scheme} {bcs}: scalars {sc} vs {one.scalars()}", flush=True)
    ok = ok and same
    s.close()
    one.close()
flag = torch.tensor([1 if ok else 0], device="cuda")
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    print("NCCL_INVARIANCE_OK" if int(flag.item()) == 1 else "NCCL_INVARIANCE_FAILED")
dist.destroy_process_group()
sys.exit(0 if int(flag.item()) == 1 else 1)
