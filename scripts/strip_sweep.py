"""x-sweep strip length at the slab shape of the 8-GPU run (2048 columns x 16384 rows on ONE GPU): time the sweeps for
LSC_XS_STRIP overrides (each in a fresh process; the override is read once)."""
import json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r'''
import os, sys, json
sys.path.insert(0, %r)
import lsc_internship_b200 as lsc
nx, ny = int(sys.argv[1]), int(sys.argv[2])
s = lsc.MusclNew(nx, ny, 0, 1, 0, 1, bc_x="periodic", bc_y="periodic", problem="ot")
s.init_problem(); s.step(3)
s.set_profiling(True); s.kernel_times(); s.step(30); mx, my, kx, ky = s.kernel_times()
print(json.dumps({"strip": os.environ.get("LSC_XS_STRIP", "auto"), "nx": nx, "ny": ny, "x_ms": mx / kx, "y_ms": my / ky, "hash": "%%016x" %% s.state_hash()}))
''' % ROOT
nx, ny = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (2048, 16384)
for strip in ["auto"] + sys.argv[3:] if len(sys.argv) > 3 else ["auto", "24", "32", "45", "52", "64", "69", "79", "90", "128"]:
    env = dict(os.environ)
    if strip != "auto":
        env["LSC_XS_STRIP"] = strip
    r = subprocess.run([sys.executable, "-c", CHILD, str(nx), str(ny)], env=env, capture_output=True, text=True)
    print(r.stdout.strip() or r.stderr[-300:], flush=True)
