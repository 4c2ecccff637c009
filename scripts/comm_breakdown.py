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
# the two orders alternate (6 rounds each) so that clock / power drift hits both alike
rounds = 6
acc = {0: [0.0] * 7, 1: [0.0] * 7}
for r in range(2 * rounds):
    mode = r % 2
    s.set_overlap(mode)
    s.step(1)
    s.kernel_times()
    s.comm_times()
    dist.barrier()
    torch.cuda.synchronize()
    ms = s.step_timed(steps)
    mx, my, kx, ky = s.kernel_times()
    mh, mr, kh, kr = s.comm_times()
    tm = torch.tensor([ms], device="cuda")
    dist.all_reduce(tm, op=dist.ReduceOp.MAX)
    for q, v in enumerate((ms, tm.item(), mx, my, mh, mr, float(steps))):
        acc[mode][q] += v
for mode, name in ((0, "serial "), (1, "overlap")):
    ms, mmax, mx, my, mh, mr, n = acc[mode]
    # serial: the exchange sits on the solver's stream (its time adds to the step); overlap: it runs on the
    # second stream behind the slab-edge columns, so "other" = step - x - y - allreduce is what is still exposed
    exposed = (ms - mx - my - mr - (mh if mode == 0 else 0.0)) / n
    print(f"[rank {rank}/{world}] {name} step {ms/n:7.3f} ms (max over ranks {mmax/n:7.3f}) | x {mx/n:6.3f} "
          f"y {my/n:6.3f} halo {mh/n:6.3f} allreduce(incl. wait for peers) {mr/n:6.3f} other {exposed:6.3f}", flush=True)
s.close()
dist.destroy_process_group()
