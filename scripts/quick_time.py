"""Quick device-side timing of lsc_mhd_step on an Orszag-Tang grid (development aid)."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lsc_internship_b200 as lsc
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
s = lsc.MusclNew(n, n, 0, 1, 0, 1, bc_x="periodic", bc_y="periodic", problem="ot")
t0 = time.time(); s.init_problem(); print("init", time.time() - t0, "s")
s.step(2)
t0 = time.time(); s.step(steps); dt = time.time() - t0
print(f"n={n} steps={steps} {dt/steps*1e3:.3f} ms/step  {n*n*steps/dt/1e9:.3f} Gcell-updates/s  "
      f"roofline(432B, 6544GB/s)={n*n*steps/dt*432/6544.3e9:.3f}")
print(s.scalars())
