"""Under torchrun (one rank per GPU): where a multi-GPU step spends its time -- sweeps, NCCL halo exchange,
max-allreduce, everything else (host readback + launch gaps + waiting for the slowest rank at the allreduce).
usage: torchrun ... scripts/comm_breakdown.py [N] [steps]"""
import ctypes
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lsc_internship_b200 as lsc  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
L = lsc.lib()
s = lsc.MusclNew(n, n, 0, 1, 0, 1, bc_x="periodic", bc_y="periodic", problem="ot", device=local, rank=rank, nranks=world)
idbuf = ctypes.create_string_buffer(128)
if rank == 0:
    lsc._capi.check(L.lsc_comm_unique_id(ctypes.cast(idbuf, ctypes.c_void_p)))
t = torch.tensor(list(idbuf.raw), dtype=torch.uint8, device="cuda")
dist.broadcast(t, 0)
s.comm_init(bytes(t.cpu().tolist()))
s.init_problem()
s.step(3)
s.set_profiling(True)
dist.barrier()
torch.cuda.synchronize()
ms = s.step_timed(steps)
mx, my, kx, ky = s.kernel_times()
mh, mr, kh, kr = s.comm_times()
print(f"[rank {rank}/{world}] step {ms/steps:7.3f} ms | x {mx/kx:6.3f} y {my/ky:6.3f} halo {mh/max(kh,1):6.3f} "
      f"allreduce(incl. wait for peers) {mr/max(kr,1):6.3f} other {(ms-mx-my-mh-mr)/steps:6.3f}", flush=True)
s.close()
dist.destroy_process_group()
