"""Quick device-side timing of the Euler SLIC step (cylinder explosion, transmissive)."""
import hashlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lsc_internship_b200 as lsc
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 50
s = lsc.SlicEuler2d(n, n, bc="transmissive", problem="cylinder")
s.init_problem()
s.step(3)
s.set_profiling(True)
ms = s.step_timed(steps)
mx, my, kx, ky = s.kernel_times()
h = hashlib.sha256(s.download().tobytes()).hexdigest()[:12] if len(sys.argv) > 3 else ""
print(f"{os.environ.get('LSC_B200_LIB','default')[-40:]:40s} euler n={n} {ms/steps:8.3f} ms/step x={mx/kx:7.3f} y={my/ky:7.3f} "
      f"{n*n*steps/ms/1e6:6.3f} Gcu/s frac160={n*n*steps/(ms*1e-3)*160/6544.3e9:.3f} redo={s.ieee_fallback_count()} {h}")
