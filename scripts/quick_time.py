"""Quick device-side timing of lsc_mhd_step on an Orszag-Tang grid (development aid).
usage: quick_time.py N[xM] STEPS [hash]   -- LSC_B200_LIB selects the build under test; NxM = N columns (x) by M
rows (y), e.g. 2048x16384 is the slab one of 8 GPUs holds of the 16384^2 grid."""
import hashlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lsc_internship_b200 as lsc
a = (sys.argv[1] if len(sys.argv) > 1 else "4096").split("x")
n, m = int(a[0]), int(a[-1])
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
s = lsc.MusclNew(n, m, 0, 1, 0, 1, bc_x="periodic", bc_y="periodic", problem="ot")
s.init_problem()
s.step(2)
s.set_profiling(True)
ms = s.step_timed(steps)
mx, my, kx, ky = s.kernel_times()
h = ""
if len(sys.argv) > 3:
    h = hashlib.sha256(s.download().tobytes()).hexdigest()[:16]
print(f"{os.environ.get('LSC_B200_LIB','default'):40s} n={n}x{m} {ms/steps:8.3f} ms/step x={mx/kx:7.3f} y={my/ky:7.3f} "
      f"{n*m*steps/ms/1e6:6.3f} Gcu/s  frac={n*m*steps/(ms*1e-3)*432/6544.3e9:.3f} {h}")
