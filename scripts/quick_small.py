"""Small-grid timings of the device-resident loop (BASELINE configs 1 and 2): what a step costs beyond its sweeps."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lsc_internship_b200 as lsc

tag = "graphs off" if os.environ.get("LSC_NO_GRAPH") else "graphs on"
# C1: Orszag-Tang 256^2 to t = 0.5 (407 steps; dump pauses at every 0.1)
s = lsc.MusclNew(256, 256, 0, 1, 0, 1, tStop=0.5, bc_x="periodic", bc_y="periodic", problem="ot")
best = 1e9
for _ in range(3):
    s.init_problem()
    t0 = time.perf_counter()
    n = s.run(tStop=0.5)
    best = min(best, time.perf_counter() - t0)
syncs, misses = s.loop_stats()
s.init_problem()
ms407 = s.step_timed(407)
s.set_profiling(True); s.kernel_times(); s.step(50); mx, my, kx, ky = s.kernel_times(); s.set_profiling(False)
print(json.dumps({"config": "c1 OT 256^2 to t=0.5", "mode": tag, "steps": n, "run_wall_ms_best_of_3": best * 1e3,
                  "step_timed_407_ms": ms407, "us_per_step": ms407 / 407 * 1e3, "x_sweep_us": mx / kx * 1e3,
                  "y_sweep_us": my / ky * 1e3, "t_final": s.scalars()[0]}))
s.close()
# C2: Kelvin-Helmholtz 1024^2
s = lsc.MusclNew(1024, 1024, 0, 1, -1, 1, bc_x="periodic", bc_y="reflective", problem="kh")
s.init_problem(); s.step(5)
ms = min(s.step_timed(200) for _ in range(3))
s.set_profiling(True); s.kernel_times(); s.step(100); mx, my, kx, ky = s.kernel_times(); s.set_profiling(False)
x, y = mx / kx, my / ky
print(json.dumps({"config": "c2 KH 1024^2", "mode": tag, "ms_per_step": ms / 200, "x_sweep_ms": x, "y_sweep_ms": y,
                  "ratio_to_sweeps": ms / 200 / (x + y), "non_sweep_us": (ms / 200 - x - y) * 1e3,
                  "cell_updates_per_s": 1024 * 1024 * 200 / (ms * 1e-3)}))
s.close()
