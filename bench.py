#!/usr/bin/env python3
"""bench.py -- cell-updates/s (fp64) of the 2-D MUSCL-Hancock + HLLC + GLM MHD step on B200.

    python bench.py --gpus N --steps K --warmup W            (N>1: launched under torchrun)
    python bench.py --impl reference --gpus N --steps K --warmup W
    python bench.py --workload euler|tabeos ...              (secondary BASELINE configs 3 and 5, same contract)

Workload (BASELINE.json configs[3]): Orszag-Tang vortex, 16384 x 16384, periodic, fp64; at N>1 the
SAME grid slab-decomposed along x (strong scaling), NCCL halo exchange + one max-allreduce per step.
A "step" is one full time step (x sweep + y sweep) of the whole grid.

One JSON line on rank 0.  `value` = cell-updates/s with the state resident in HBM, timed with CUDA
events on the solver's stream, max over ranks; the time loop is device-resident, so the K timed steps
are ONE enqueued batch with one host synchronisation at its end, and no per-kernel events sit inside it.
`e2e` = the same metric through the public C ABI with HOST buffers: upload of the pinned host state + K steps
+ download, all inside the timed region.  `roofline` is for the dominant kernel = the slower of the two sweeps
(both move 216 algorithmic bytes per cell), its launch time measured live with CUDA events on the launching
stream in a SECOND pass of the same K steps right behind the timed one.  `sim.state_hash` is a
decomposition-independent 64-bit fingerprint of the final interior (lsc_mhd_state_hash, summed over ranks):
equal across the N = 1/2/4/8 lines of a scaling run exactly when every bit of the state agrees.
`cpu_baseline` (N = 1 only, as the tier rule asks) is the REAL reference (oracle/_ref/muscl_ref, compiled
from the reference sources with g++ -O3) on one host core.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "cell_updates_per_s"
UNIT = "cell-updates/s"
B_ALG = 432.0          # algorithmic bytes per cell-update
B_ALG_X = 216.0        # x sweep: read u^n 72 + read u^{n-1} 72 + write uBar 72
B_ALG_Y = 216.0        # y sweep: read uBar 72 + read u^n 72 + write u^{n+1} 72
NOMINAL_HBM_GBS = 8000.0
TRAFFIC_FILE = "r2/traffic.json"  # ncu --set full dram bytes per cell of each sweep (profiles/)


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """SM clock + throttle reasons sampled DURING the timed region: NVML polled every 5 ms from a thread
    (ctypes releases the GIL while the solver runs); nvidia-smi -lms 100 as the fallback."""

    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index):
        self.index = index
        self.rows = []          # (sm_mhz, power_w, reasons bitmask or list)
        self.max_mhz = None
        self.proc = None
        self.thread = None
        self.stop_flag = False
        self.nvml = None

    def _visible_index(self):
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            try:
                return int(vis.split(",")[self.index])
            except Exception:
                pass
        return self.index

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nvml = pynvml
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(self._visible_index())
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self.thread = threading.Thread(target=self._poll_nvml, daemon=True)
            self.thread.start()
            return
        except Exception:
            self.nvml = None
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self._visible_index()}", f"--query-gpu={q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump_smi, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _poll_nvml(self):
        n = self.nvml
        bits = [(getattr(n, "nvmlClocksEventReasonHwSlowdown", 0x8), "hw_slowdown"),
                (getattr(n, "nvmlClocksEventReasonHwThermalSlowdown", 0x40), "hw_thermal_slowdown"),
                (getattr(n, "nvmlClocksEventReasonSwThermalSlowdown", 0x20), "sw_thermal_slowdown"),
                (getattr(n, "nvmlClocksEventReasonSwPowerCap", 0x4), "sw_power_cap")]
        get_reasons = getattr(n, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
            getattr(n, "nvmlDeviceGetCurrentClocksThrottleReasons")
        while not self.stop_flag:
            try:
                mhz = float(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM))
                pw = n.nvmlDeviceGetPowerUsage(self.handle) / 1000.0
                mask = int(get_reasons(self.handle))
                self.rows.append((mhz, pw, [nm for b, nm in bits if mask & b]))
            except Exception:
                pass
            time.sleep(0.005)

    def _pump_smi(self):
        for line in self.proc.stdout:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                self.max_mhz = float(f[1])
                self.rows.append((float(f[0]), float(f[2]),
                                  [nm for nm, v in zip(self.NAMES, f[3:7]) if v.lower().startswith("active")]))
            except ValueError:
                continue

    def stop(self):
        self.stop_flag = True
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
        if self.thread:
            self.thread.join(timeout=2)
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "samples": 0,
                    "reasons": ["no samples (nvml and nvidia-smi unavailable)"]}
        sm = sorted(r[0] for r in self.rows)
        reasons = sorted({nm for r in self.rows for nm in r[2]})
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": self.max_mhz, "power_w_max": max(r[1] for r in self.rows),
                "samples": len(sm), "source": "nvml" if self.nvml else "nvidia-smi", "reasons": reasons}


# --------------------------------------------------------------------------------------
# workloads: the headline (BASELINE.json configs[3]) and the two secondary single-/multi-GPU configurations
# --------------------------------------------------------------------------------------
WORKLOADS = {
    "mhd": {
        "default_n": 16384, "b_alg": 432.0, "b_x": 216.0, "b_y": 216.0, "nvar": 9,
        "kernels": ("x_sweep_kernel", "y_sweep_kernel"),
        "name": "Orszag-Tang vortex, 2-D ideal MHD MUSCL-Hancock+HLLC+GLM, {n}x{n}, periodic, fp64, C=0.8, gamma=5/3 "
                "(BASELINE.json configs[3])",
        "ref_exe": "muscl_ref", "ref_n": 1024,
        "ref_name": "oracle/_ref/muscl_ref (MUSCL_NEW.cpp built from the reference sources, g++ -O3, serial)",
        "ref_why": "the 16384^2 grid itself needs 290 GB in the reference and cannot be allocated",
        "bytes_note": "x: read u^n 72 + u^{n-1} 72, write uBar 72; y: read uBar 72 + u^n 72, write u^{n+1} 72",
    },
    "euler": {
        "default_n": 4096, "b_alg": 160.0, "b_x": 64.0, "b_y": 96.0, "nvar": 4,
        "kernels": ("euler_x_sweep_kernel", "euler_y_sweep_kernel"),
        "name": "2-D Euler cylindrical explosion, SLIC (MUSCL-Hancock + FORCE), {n}x{n}, transmissive, fp64, C=0.8, "
                "gamma=1.4 (BASELINE.json configs[2])",
        "ref_exe": "gslic_ref", "ref_n": 1024,
        "ref_name": "oracle/_ref/gslic_ref (generaliseSLIC.cpp built from the reference sources, g++ -O3, serial)",
        "ref_why": "bounded sample of the 4096^2 workload",
        "bytes_note": "x: read u 32, write ubar 32; y: read ubar 32 + u 32, write u' 32",
    },
    "tabeos": {
        "default_n": 8192, "b_alg": 160.0, "b_x": 64.0, "b_y": 96.0, "nvar": 4,
        "kernels": ("euler_x_sweep_kernel<TableEos>", "euler_y_sweep_kernel<TableEos>"),
        "name": "tabulated-EoS Euler SLIC (TabEoS.cpp), {n}x{n}, transmissive, fp64, synthetic Plasma19-format table "
                "(BASELINE.json configs[4])",
        "ref_exe": "tabeos_ref", "ref_n": 512,
        "ref_name": "oracle/_ref/tabeos_ref (TabEoS.cpp built from the reference sources, g++ -O3, serial)",
        "ref_why": "bounded sample of the 8192^2 workload",
        "bytes_note": "as euler; the 4 MB table stays in L2",
    },
}


def make_solver(workload, n, local_rank, rank, world, arithmetic, tmpdir):
    import lsc_internship_b200 as lsc
    if workload == "mhd":
        return lsc.MusclNew(n, n, 0.0, 1.0, 0.0, 1.0, tStop=0.5, C=0.8, bc_x="periodic", bc_y="periodic", problem="ot",
                            device=local_rank, rank=rank, nranks=world, arithmetic=arithmetic)
    if workload == "euler":
        return lsc.SlicEuler2d(n, n, bc="transmissive", problem="cylinder", device=local_rank, rank=rank, nranks=world)
    from lsc_internship_b200.synthetic_eos import write_eos_table
    path = os.path.join(tmpdir, f"Plasma19_EoS_{rank}.txt")
    write_eos_table(path)
    s = lsc.SlicEuler2d(n, n, bc="transmissive", problem="tabeos", device=local_rank, rank=rank, nranks=world)
    s.set_eos_table(*lsc.SlicEuler2d.read_eos_table(path))
    return s


# --------------------------------------------------------------------------------------
# reference arm / cpu_baseline: the reference's own CPU programs (oracle/_ref/*)
# --------------------------------------------------------------------------------------
def run_reference_sample(n, steps, workload="mhd"):
    """`steps` steps of the REAL reference program on an n x n sample of the workload, 1 thread.
    Returns (cell-updates/s, loop seconds)."""
    exe = os.path.join(ROOT, "oracle", "_ref", WORKLOADS[workload]["ref_exe"])
    if not os.path.exists(exe):
        raise FileNotFoundError(exe)
    with tempfile.TemporaryDirectory() as td:
        if workload == "mhd":
            cmd = [exe, "ot", str(n), str(n), "0", "1", "0", "1", "1e9", "0.8", "periodic", "periodic", str(steps), "0"]
        elif workload == "euler":
            cmd = [exe, str(n), str(n), "0", "1", "0", "1", "1e9", "0.8", "1.4", str(steps), "0", "1"]
        else:
            sys.path.insert(0, ROOT)
            from lsc_internship_b200.synthetic_eos import write_eos_table
            write_eos_table(os.path.join(td, "Plasma19_EoS.txt"))
            cmd = [exe, str(n), str(n), "0", "1", "0", "1", "1e9", "0.8", str(steps), "0"]
        r = subprocess.run(cmd, cwd=td, capture_output=True, text=True, check=True)
    secs = float(r.stderr.split("elapsed_s=")[1].split()[0])
    return n * n * steps / secs, secs


def cpu_baseline(workload="mhd", n=None, steps=None):
    w = WORKLOADS[workload]
    n = n or w["ref_n"]
    steps = steps or (16 if workload == "mhd" else 8)
    try:
        v, secs = run_reference_sample(n, steps, workload)
        return {"value": v, "unit": UNIT, "cores": 1, "kind": "reference",
                "host_cores_available": os.cpu_count(),
                "sample": f"{w['ref_name']}: {n}x{n}, {steps} steps, {secs:.1f} s loop time; {w['ref_why']}"}
    except Exception as e:  # the oracle always exists; report loudly if it does not
        return {"value": None, "unit": UNIT, "cores": 1, "kind": "reference", "sample": f"FAILED: {e}"}


def main_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    w = WORKLOADS[args.workload]
    n = args.ref_n or w["ref_n"]
    named = args.n or w["default_n"]
    for _ in range(min(args.warmup, 1)):
        run_reference_sample(min(n, 256), 2, args.workload)
    t0 = time.time()
    v, secs = run_reference_sample(n, args.steps, args.workload)
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": secs / args.steps * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": w["name"].format(n=named) + "; reference CPU program",
                   "sample_grid": [n, n], "named_grid": [named, named]},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": 1, "kind": "reference",
                         "host_cores_available": os.cpu_count(),
                         "sample": f"each step = one time step on {n}x{n} (bounded sample of the {named}^2 workload; "
                                   f"{w['ref_why']}); {w['ref_name']}, 1 thread"},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "wall_s": time.time() - t0,
    }
    print(json.dumps(line))
    return 0


# --------------------------------------------------------------------------------------
# the CUDA arm
# --------------------------------------------------------------------------------------
def pinned_array(lib, shape):
    """numpy view of pinned host memory from lsc_host_alloc (falls back to pageable memory)."""
    import numpy as np
    nbytes = int(np.prod(shape)) * 8
    ptr = ctypes.c_void_p()
    if lib.lsc_host_alloc(ctypes.byref(ptr), nbytes) == 0 and ptr.value:
        buf = (ctypes.c_double * (nbytes // 8)).from_address(ptr.value)
        return np.frombuffer(buf, dtype=np.float64).reshape(shape), ptr, True
    return np.empty(shape, dtype=np.float64), None, False


def main_gpu(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("--gpus N>1 must be launched with torch.distributed.run (one rank per GPU)")
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    import numpy as np
    import lsc_internship_b200 as lsc
    L = lsc.lib()
    w = WORKLOADS[args.workload]
    n = args.n or w["default_n"]
    tmp = tempfile.TemporaryDirectory()
    s = make_solver(args.workload, n, local_rank, rank, world, args.arithmetic, tmp.name)
    if world > 1:
        import torch
        idbuf = ctypes.create_string_buffer(128)
        if rank == 0:
            lsc._capi.check(L.lsc_comm_unique_id(ctypes.cast(idbuf, ctypes.c_void_p)))
        t = torch.tensor(list(idbuf.raw), dtype=torch.uint8, device="cuda")
        dist.broadcast(t, 0)
        s.comm_init(bytes(t.cpu().tolist()))

    def barrier():
        if dist is not None:
            import torch
            dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(x):
        if dist is None:
            return x
        import torch
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_mod_2_64(x):
        """state-hash words of the ranks added modulo 2^64 (two 32-bit halves through an int64 all-gather)"""
        if dist is None:
            return x
        import torch
        mine = torch.tensor([x & 0xffffffff, x >> 32], dtype=torch.int64, device="cuda")
        parts = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(parts, mine)
        return sum(int(p[0].item()) + (int(p[1].item()) << 32) for p in parts) & ((1 << 64) - 1)

    s.init_problem()
    if args.warmup > 0:
        s.step(args.warmup)

    # ---- slab calibration (MHD, N > 1): every step ends in a max-allreduce, so the slowest GPU paces all ranks, and
    # under the board power cap individual GPUs differ by a few per cent.  Measure each rank's sweep time over the
    # warm-up steps and size the slabs in proportion to its speed (explicit bounds, lsc_mhd_config.slab_i0/slab_nx).
    # The decomposition does not touch the arithmetic: sim.state_hash is the same for any widths.
    slabs = None
    if world > 1 and args.workload == "mhd" and args.balance:
        import torch
        s.set_profiling(True)
        s.kernel_times()
        s.step(max(args.warmup, 3))
        mx, my, kx, ky = s.kernel_times()
        s.set_profiling(False)
        mine = torch.tensor([(mx + my) / max(kx, 1), float(s.nxl)], dtype=torch.float64, device="cuda")
        allr = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
        times = [float(a[0].item()) for a in allr]
        widths = [float(a[1].item()) for a in allr]
        speed = [wd / tm for wd, tm in zip(widths, times)]          # columns per ms of each GPU
        share = [n * sp / sum(speed) for sp in speed]
        new_w = [max(4, int(round(x))) for x in share]
        new_w[new_w.index(max(new_w))] += n - sum(new_w)             # rounding remainder to the widest slab
        slabs = {"sweep_ms_even_split": times, "widths": new_w,
                 "spread_even_split": (max(times) - min(times)) / max(times)}
        if max(abs(a - b) for a, b in zip(new_w, widths)) >= 2:
            s.close()
            i0 = sum(new_w[:rank])
            s = lsc.MusclNew(n, n, 0.0, 1.0, 0.0, 1.0, tStop=0.5, C=0.8, bc_x="periodic", bc_y="periodic",
                             problem="ot", device=local_rank, rank=rank, nranks=world, arithmetic=args.arithmetic,
                             slab=(i0, new_w[rank]))
            idbuf = ctypes.create_string_buffer(128)
            if rank == 0:
                lsc._capi.check(L.lsc_comm_unique_id(ctypes.cast(idbuf, ctypes.c_void_p)))
            t = torch.tensor(list(idbuf.raw), dtype=torch.uint8, device="cuda")
            dist.broadcast(t, 0)
            s.comm_init(bytes(t.cpu().tolist()))
        else:
            slabs["widths"] = [int(x) for x in widths]
        # start over from the initial condition: every N then hashes the state after exactly warmup + K steps
        s.init_problem()
        if args.warmup > 0:
            s.step(args.warmup)

    # ---- device-resident throughput: K steps as one enqueued batch, CUDA events on the solver's stream
    sampler = ClockSampler(local_rank)
    barrier()
    launches0 = L.lsc_kernel_launch_count()
    sampler.start()
    ms = s.step_timed(args.steps)
    barrier()
    clocks = sampler.stop()
    launches = L.lsc_kernel_launch_count() - launches0
    ms = max_over_ranks(ms)
    cells = float(n) * float(n)
    value = cells * args.steps / (ms * 1e-3)
    scal = s.scalars()
    state_hash = sum_mod_2_64(s.state_hash())
    loop = s.loop_stats() if hasattr(s, "loop_stats") else None

    # ---- second pass, untimed for `value`: the same K steps with CUDA events around every sweep launch
    s.set_profiling(True)
    s.kernel_times()
    s.step(args.steps)
    ms_x, ms_y, nxk, nyk = s.kernel_times()
    comm = s.comm_times() if (world > 1 and hasattr(s, "comm_times")) else None
    s.set_profiling(False)

    # ---- end to end through the C ABI with host buffers: upload + K steps + download
    host, hptr, pinned = pinned_array(L, s.shape)
    s.download(host)  # untimed: this is the caller's host-resident state
    state_bytes = host.nbytes
    barrier()
    t0 = time.perf_counter()
    s.upload(host)
    s.step(args.steps)
    s.download(host)
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    e2e_value = cells * args.steps / e2e_s
    finite = bool(np.isfinite(host[2:-2, 2:-2]).all())
    if hptr is not None:
        del host
        L.lsc_host_free(hptr)
    s.close()
    tmp.cleanup()

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return 0

    peak, peak_src = measured_peaks()
    cells_local = cells / world
    # dominant kernel = the sweep furthest below the roofline it is measured against (lowest achieved GB/s)
    per_x, per_y = ms_x / max(nxk, 1), ms_y / max(nyk, 1)
    gbs_x = cells_local * w["b_x"] / (per_x * 1e-3) / 1e9
    gbs_y = cells_local * w["b_y"] / (per_y * 1e-3) / 1e9
    if gbs_x <= gbs_y:
        dom_name, dom_ms, dom_b, achieved = w["kernels"][0], per_x, w["b_x"], gbs_x
    else:
        dom_name, dom_ms, dom_b, achieved = w["kernels"][1], per_y, w["b_y"], gbs_y
    traffic, traffic_src = None, None
    try:  # dram bytes per cell of that kernel from the committed ncu --set full capture, scaled to this launch
        with open(os.path.join(ROOT, "profiles", TRAFFIC_FILE)) as f:
            tj = json.load(f)
        traffic = tj[dom_name]["bytes_per_cell"] * cells_local
        traffic_src = tj["source"]
    except Exception:
        pass
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": w["name"].format(n=n),
                   "grid": [n, n], "parallelism": f"slab-x{world}", "slabs": slabs,
                   "cache": "inputs larger than L2 (3 states x %.2f GB per GPU vs 126 MB L2), no flush needed"
                            % (cells_local * 8 * w["nvar"] / 1e9),
                   "arithmetic": ("reference: fp64, no FMA contraction, IEEE division, reference operation order "
                                  "(bit-identical to the reference)") if args.arithmetic == "reference" else
                                 ("relaxed (NOT the parity path, SURVEY 8f N3): fp64 with FMA contraction and "
                                  "1-multiply quotients; deviation measured in tests/test_mhd_gpu.py")},
        "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": state_bytes / args.steps,
                "d2h_bytes_per_step": state_bytes / args.steps,
                "timed_region": f"upload(host AoS, {state_bytes / 1e9:.2f} GB/rank, "
                                f"{'pinned' if pinned else 'pageable'}) + {args.steps} steps (one enqueued batch, one "
                                f"host synchronisation) + download(host AoS); wall clock, max over ranks",
                "seconds": e2e_s, "result_finite": finite},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "kernel": dom_name, "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                     "algorithmic_bytes_per_cell": dom_b, "cells_per_launch": cells_local,
                     "launch_ms": dom_ms, "x_sweep_ms": per_x, "y_sweep_ms": per_y,
                     "x_sweep_GBs": gbs_x, "y_sweep_GBs": gbs_y, "bytes": w["bytes_note"],
                     "kernel_times_from": "second pass of the same K steps with CUDA events around each sweep launch "
                                          "(not inside the region `value` is timed over)",
                     "non_sweep_ms_per_step": ms / args.steps - per_x - per_y,
                     "whole_step_frac_of_measured": value * w["b_alg"] / world / (peak * 1e9),
                     "whole_step_frac_of_nominal_8TBs": value * w["b_alg"] / world / (NOMINAL_HBM_GBS * 1e9),
                     "note": "bound by the fp64 pipe / issue port, not HBM (no FMA contraction, IEEE div/sqrt): "
                             "ncu digests under profiles/"},
        "sim": {"t": scal[0], "dt": scal[1], "ch": scal[2] if len(scal) > 3 else None, "steps_taken": scal[-1],
                "state_hash": f"{state_hash:016x}",
                "state_hash_of": "interior of the state after warmup + K steps (lsc_*_state_hash, summed over ranks "
                                 "modulo 2^64): equal for every N exactly when all bits agree"},
    }
    if loop is not None:
        line["sim"]["host_syncs_of_the_loop"] = loop[0]
        line["sim"]["exp_table_misses"] = loop[1]
    if comm is not None:
        line["roofline"]["nccl_halo_ms_per_step"] = comm[0] / max(comm[2], 1)
        line["roofline"]["nccl_allreduce_ms_per_step"] = comm[1] / max(comm[3], 1)
    try:  # the 16384^2 run is power-capped and boxes settle at different clocks: the same metric per SM clock
        mhz = clocks.get("sm_mhz") if isinstance(clocks, dict) else None
        if mhz:
            line["roofline"]["cell_updates_per_gpu_per_sm_clock"] = value / world / (float(mhz) * 1e6)
    except Exception:
        pass
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(args.workload)
    print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="cuda", choices=["cuda", "reference"])
    ap.add_argument("--workload", default="mhd", choices=sorted(WORKLOADS),
                    help="mhd = the headline (BASELINE configs[3]); euler / tabeos = configs[2] / configs[4]")
    ap.add_argument("--grid", "--n", dest="n", type=int, default=0, help="grid edge (default: the named size)")
    ap.add_argument("--ref-n", type=int, default=0, help="grid edge of the bounded CPU reference sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--balance", action="store_true",
                    help="N > 1 (mhd): size the slabs by the measured sweep time of each GPU instead of the even split "
                         "(measured on 8 x B200: 54.8 G against 55.2 G for the even split -- the 1-2 %% spread between "
                         "GPUs is not stable enough from pass to pass to be calibrated away; off by default)")
    ap.add_argument("--arithmetic", default="reference", choices=["reference", "relaxed"],
                    help="reference = bit-identical IEEE path (default, the one that is judged); relaxed = N3 variant")
    args = ap.parse_args()
    if args.impl == "reference":
        return main_reference(args)
    return main_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
