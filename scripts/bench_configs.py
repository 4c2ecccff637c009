#!/usr/bin/env python3
"""Secondary BASELINE.json configurations (bench.py measures the headline configs[3]):
   c1  Orszag-Tang MHD 256^2 to t = 0.5 (the reference's CPU-runnable case): wall time, steps
   c2  Kelvin-Helmholtz MHD 1024^2, periodic-x / reflective-y: cell-updates/s over 200 steps
   c3  Euler SLIC 4096^2, transmissive, cylindrical explosion: cell-updates/s over 100 steps + CPU reference
   c5  tabulated-EoS Euler SLIC (TabEoS.cpp) 8192^2 on the synthetic Plasma19-format table + CPU reference
One JSON line per config; device timing with CUDA events on the solver's stream."""
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import lsc_internship_b200 as lsc  # noqa: E402

PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(
    os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
which = sys.argv[1:] or ["c1", "c2", "c3", "c5"]

if "c1" in which:
    s = lsc.MusclNew(256, 256, 0, 1, 0, 1, tStop=0.5, bc_x="periodic", bc_y="periodic", problem="ot")
    s.init_problem()
    t0 = time.perf_counter()
    n = s.run(tStop=0.5)
    wall = time.perf_counter() - t0
    print(json.dumps({"config": "c1 OT MHD 256^2 periodic t=0.5", "steps": n, "t_final": s.scalars()[0], "wall_s": wall,
                      "cell_updates_per_s": 256 * 256 * n / wall,
                      "note": "65k cells is far below one wave of the GPU; per-step host readback dominates"}))
    s.close()

if "c2" in which:
    s = lsc.MusclNew(1024, 1024, 0, 1, -1, 1, bc_x="periodic", bc_y="reflective", problem="kh")
    s.init_problem()
    s.step(3)
    s.set_profiling(True)
    ms = s.step_timed(200)
    mx, my, kx, ky = s.kernel_times()
    v = 1024 * 1024 * 200 / (ms * 1e-3)
    print(json.dumps({"config": "c2 KH MHD 1024^2 periodic-x/reflective-y", "steps": 200, "ms_per_step": ms / 200,
                      "cell_updates_per_s": v, "hbm_frac_of_measured": v * 432 / (PEAK * 1e9),
                      "x_sweep_ms": mx / kx, "y_sweep_ms": my / ky, "ieee_fallback_steps": s.ieee_fallback_count()}))
    s.close()

if "c3" in which:
    n = 4096
    s = lsc.SlicEuler2d(n, n, bc="transmissive", problem="cylinder")
    s.init_problem()
    s.step(3)
    s.set_profiling(True)
    ms = s.step_timed(100)
    mx, my, kx, ky = s.kernel_times()
    v = n * n * 100 / (ms * 1e-3)
    line = {"config": "c3 Euler SLIC 4096^2 transmissive cylindrical explosion", "steps": 100, "ms_per_step": ms / 100,
            "cell_updates_per_s": v, "algorithmic_bytes_per_cell_update": 160,
            "hbm_frac_of_measured": v * 160 / (PEAK * 1e9), "x_sweep_ms": mx / kx, "y_sweep_ms": my / ky,
            "x_sweep_GBs": n * n * 64 / (mx / kx * 1e-3) / 1e9, "y_sweep_GBs": n * n * 96 / (my / ky * 1e-3) / 1e9,
            "ieee_fallback_steps": s.ieee_fallback_count()}
    s.close()
    exe = os.path.join(ROOT, "oracle", "_ref", "gslic_ref")
    if os.path.exists(exe):
        with tempfile.TemporaryDirectory() as td:
            r = subprocess.run([exe, "1024", "1024", "0", "1", "0", "1", "1e9", "0.8", "1.4", "8", "0", "1"], cwd=td,
                               capture_output=True, text=True)
        secs = float(r.stderr.split("elapsed_s=")[1].split()[0])
        line["cpu_reference"] = {"cell_updates_per_s": 1024 * 1024 * 8 / secs, "cores": 1,
                                 "sample": "oracle/_ref/gslic_ref (generaliseSLIC.cpp) 1024^2 cylinder, 8 steps"}
    print(json.dumps(line))


if "c5" in which:
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oracle_util import write_eos_table  # the synthetic table (the real Plasma19_EoS.txt is not in the reference)
    n = int(os.environ.get("LSC_C5_N", "8192"))
    with tempfile.TemporaryDirectory() as td:
        path = os.path.join(td, "Plasma19_EoS.txt")
        write_eos_table(path)
        table = lsc.SlicEuler2d.read_eos_table(path)
        s = lsc.SlicEuler2d(n, n, bc="transmissive", problem="tabeos")
        s.set_eos_table(*table)
        s.init_problem()
        s.step(3)
        s.set_profiling(True)
        ms = s.step_timed(20)
        mx, my, kx, ky = s.kernel_times()
        v = n * n * 20 / (ms * 1e-3)
        line = {"config": f"c5 tabulated-EoS Euler SLIC (TabEoS.cpp) {n}^2 transmissive, synthetic table", "steps": 20,
                "ms_per_step": ms / 20, "cell_updates_per_s": v, "algorithmic_bytes_per_cell_update": 160,
                "hbm_frac_of_measured": v * 160 / (PEAK * 1e9), "x_sweep_ms": mx / kx, "y_sweep_ms": my / ky,
                "ieee_fallback_blocks": s.ieee_fallback_count(),
                "eos_search": "warm-started" if s.eos_search_is_warm else "bisection",
                "note": "13 EoS lookups per cell-update (density axis search, energy-column search, reverse bilinear "
                        "interpolation): instruction bound, not HBM bound"}
        s.close()
        exe = os.path.join(ROOT, "oracle", "_ref", "tabeos_ref")
        if os.path.exists(exe):
            r = subprocess.run([exe, "512", "512", "0", "1", "0", "1", "1e9", "0.8", "4", "0"], cwd=td,
                               capture_output=True, text=True)
            secs = float(r.stderr.split("elapsed_s=")[1].split()[0])
            line["cpu_reference"] = {"cell_updates_per_s": 512 * 512 * 4 / secs, "cores": 1,
                                     "sample": "oracle/_ref/tabeos_ref (TabEoS.cpp) 512^2, 4 steps"}
    print(json.dumps(line))
