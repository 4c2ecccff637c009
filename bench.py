#!/usr/bin/env python3
"""bench.py -- cell-updates/s (fp64) of the 2-D MUSCL-Hancock + HLLC + GLM MHD step on B200.

    python bench.py --gpus N --steps K --warmup W            (N>1: launched under torchrun)
    python bench.py --impl reference --gpus N --steps K --warmup W

Workload (BASELINE.json configs[3]): Orszag-Tang vortex, 16384 x 16384, periodic, fp64; at N>1 the
SAME grid slab-decomposed along x (strong scaling), NCCL halo exchange + one max-allreduce per step.
A "step" is one full time step (x sweep + y sweep) of the whole grid.

One JSON line on rank 0.  `value` = cell-updates/s with the state resident in HBM, timed with CUDA
events on the solver's stream, max over ranks.  `e2e` = the same metric through the public C ABI with
HOST buffers: upload of the pinned host state + K steps (each with its 8-byte scalar readback) +
download, all inside the timed region.  `roofline` is for the dominant kernel (x sweep), its launch
time measured live with CUDA events on the launching stream.  `cpu_baseline` is the REAL reference
(oracle/_ref/muscl_ref, compiled from the reference sources) on one host core.
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
# reference arm / cpu_baseline: the reference's own CPU program (oracle/_ref/muscl_ref)
# --------------------------------------------------------------------------------------
def run_reference_sample(n, steps):
    """OT n x n, periodic, `steps` steps of the REAL reference, 1 thread.  Returns (cell-updates/s, secs)."""
    exe = os.path.join(ROOT, "oracle", "_ref", "muscl_ref")
    if not os.path.exists(exe):
        raise FileNotFoundError(exe)
    with tempfile.TemporaryDirectory() as td:
        r = subprocess.run([exe, "ot", str(n), str(n), "0", "1", "0", "1", "1e9", "0.8", "periodic", "periodic",
                            str(steps), "0"], cwd=td, capture_output=True, text=True, check=True)
    secs = float(r.stderr.split("elapsed_s=")[1].split()[0])
    return n * n * steps / secs, secs


def cpu_baseline(n=1024, steps=16):
    try:
        v, secs = run_reference_sample(n, steps)
        return {"value": v, "unit": UNIT, "cores": 1, "kind": "reference",
                "host_cores_available": os.cpu_count(),
                "sample": f"oracle/_ref/muscl_ref (MUSCL_NEW.cpp built from the reference sources, g++ -O2, serial): "
                          f"Orszag-Tang {n}x{n} periodic, {steps} steps, {secs:.1f} s loop time; the 16384^2 grid "
                          f"itself needs 290 GB in the reference and cannot be allocated"}
    except Exception as e:  # the oracle always exists; report loudly if it does not
        return {"value": None, "unit": UNIT, "cores": 1, "kind": "reference", "sample": f"FAILED: {e}"}


def main_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    n = args.ref_n
    for _ in range(min(args.warmup, 1)):
        run_reference_sample(256, 2)
    t0 = time.time()
    v, secs = run_reference_sample(n, args.steps)
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": secs / args.steps * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "Orszag-Tang 2-D MHD, MUSCL-Hancock+HLLC+GLM, fp64, periodic; reference CPU program",
                   "sample_grid": [n, n], "named_grid": [16384, 16384]},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": 1, "kind": "reference",
                         "host_cores_available": os.cpu_count(),
                         "sample": f"each step = one time step of Orszag-Tang {n}x{n} (bounded sample of the "
                                   f"16384^2 workload, which needs 290 GB in the reference); serial program, 1 thread"},
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
    n = args.n
    s = lsc.MusclNew(n, n, 0.0, 1.0, 0.0, 1.0, tStop=0.5, C=0.8, bc_x="periodic", bc_y="periodic", problem="ot",
                     device=local_rank, rank=rank, nranks=world, arithmetic=args.arithmetic)
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

    s.init_problem()
    if args.warmup > 0:
        s.step(args.warmup)

    # ---- device-resident throughput: K steps, CUDA events on the solver's stream
    sampler = ClockSampler(local_rank)
    s.set_profiling(True)
    s.kernel_times()
    barrier()
    launches0 = L.lsc_kernel_launch_count()
    sampler.start()
    ms = s.step_timed(args.steps)
    barrier()
    clocks = sampler.stop()
    launches = L.lsc_kernel_launch_count() - launches0
    ms_x, ms_y, nxk, nyk = s.kernel_times()
    s.set_profiling(False)
    ms = max_over_ranks(ms)
    cells = float(n) * float(n)
    value = cells * args.steps / (ms * 1e-3)
    t_sim, dt_sim, ch_sim, nstep = s.scalars()

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

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return 0

    peak, peak_src = measured_peaks()
    cells_local = cells / world
    # dominant kernel = the slower of the two sweeps (both move 216 algorithmic bytes per cell)
    per_x, per_y = ms_x / max(nxk, 1), ms_y / max(nyk, 1)
    dom_name, dom_ms = ("x_sweep_kernel", per_x) if per_x >= per_y else ("y_sweep_kernel", per_y)
    achieved = cells_local * B_ALG_X / (dom_ms * 1e-3) / 1e9
    traffic, traffic_src = None, None
    try:  # dram bytes per cell of that kernel from the committed ncu --set full capture, scaled to this launch
        with open(os.path.join(ROOT, "profiles", "r1_v5_traffic.json")) as f:
            tj = json.load(f)
        traffic = tj[dom_name]["bytes_per_cell"] * cells_local
        traffic_src = tj["source"]
    except Exception:
        pass
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"Orszag-Tang vortex, 2-D ideal MHD MUSCL-Hancock+HLLC+GLM, {n}x{n}, periodic, "
                               f"fp64, C=0.8, gamma=5/3 (BASELINE.json configs[3])",
                   "grid": [n, n], "parallelism": f"slab-x{world}",
                   "cache": "inputs larger than L2 (3 states x %.1f GB per GPU vs 126 MB L2), no flush needed"
                            % (cells_local * 72 / 1e9),
                   "arithmetic": ("reference: fp64, no FMA contraction, IEEE division, reference operation order "
                                  "(bit-identical to the reference)") if args.arithmetic == "reference" else
                                 ("relaxed (NOT the parity path, SURVEY 8f N3): fp64 with FMA contraction and "
                                  "1-multiply quotients; deviation measured in tests/test_mhd_gpu.py")},
        "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": state_bytes / args.steps,
                "d2h_bytes_per_step": state_bytes / args.steps + 8,
                "timed_region": f"lsc_mhd_upload(host AoS, {state_bytes / 1e9:.2f} GB/rank, "
                                f"{'pinned' if pinned else 'pageable'}) + {args.steps} x lsc_mhd_step (8-byte "
                                f"scalar readback each) + lsc_mhd_download(host AoS); wall clock, max over ranks",
                "seconds": e2e_s, "result_finite": finite},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "kernel": dom_name, "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                     "algorithmic_bytes_per_cell": B_ALG_X, "cells_per_launch": cells_local,
                     "launch_ms": dom_ms, "x_sweep_ms": per_x, "y_sweep_ms": per_y,
                     "whole_step_frac_of_measured": value * B_ALG / world / (peak * 1e9),
                     "whole_step_frac_of_nominal_8TBs": value * B_ALG / world / (NOMINAL_HBM_GBS * 1e9),
                     "note": "bound by the fp64 pipe / issue port, not HBM (no FMA contraction, IEEE div/sqrt): ncu "
                             "sm__inst_executed_pipe_fp64 65 % (x) / 61 % (y) of peak, profiles/r1_v5_summary.md"},
        "sim": {"t": t_sim, "dt": dt_sim, "ch": ch_sim, "steps_taken": nstep},
    }
    try:  # the 16384^2 run is power-capped and boxes settle at different clocks: the same metric per SM clock
        mhz = clocks.get("sm_mhz") if isinstance(clocks, dict) else None
        if mhz:
            line["roofline"]["cell_updates_per_gpu_per_sm_clock"] = value / world / (float(mhz) * 1e6)
    except Exception:
        pass
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline()
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
    ap.add_argument("--grid", "--n", dest="n", type=int, default=16384, help="grid edge (default: the named 16384)")
    ap.add_argument("--ref-n", type=int, default=1024, help="grid edge of the bounded CPU reference sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--arithmetic", default="reference", choices=["reference", "relaxed"],
                    help="reference = bit-identical IEEE path (default, the one that is judged); relaxed = N3 variant")
    args = ap.parse_args()
    if args.impl == "reference":
        return main_reference(args)
    return main_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
