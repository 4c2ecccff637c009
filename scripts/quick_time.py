"""Quick device-side timing of lsc_mhd_step on an Orszag-Tang grid (development aid).
usage: quick_time.py N STEPS [hash]   -- LSC_B200_LIB selects the build under test."""
import hashlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lsc_internship_b200 as lsc
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
s = lsc.MusclNew(n, n, 0, 1, 0, 1, bc_x="periodic", bc_y="periodic", problem="ot")
s.init_problem()
s.step(2)
s.set_profiling(True)
ms = s.step_timed(steps)
mx, my, kx, ky = s.kernel_times()
h = ""
if len(sys.argv) > 3:
    h = hashlib.sha256(s.download().tobytes()).hexdigest()[:16]
print(f"{os.environ.get('LSC_B200_LIB','default'):40s} n={n} {ms/steps:8.3f} ms/step x={mx/kx:7.3f} y={my/ky:7.3f} "
      f"{n*n*steps/ms/1e6:6.3f} Gcu/s  frac={n*n*steps/(ms*1e-3)*432/6544.3e9:.3f} {h}")
