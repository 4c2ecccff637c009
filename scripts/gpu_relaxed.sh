timeout 300 python -m pytest tests/test_mhd_gpu.py -m gpu -x -q -s -k "relaxed or golden or kat_hllc" 2>&1 | tail -6
cat > /tmp/qt_relaxed.py <<'PY'
import sys; sys.path.insert(0,'.')
import lsc_internship_b200 as lsc
for mode in ("reference","relaxed"):
    n=8192
    s = lsc.MusclNew(n, n, 0, 1, 0, 1, bc_x="periodic", bc_y="periodic", problem="ot", arithmetic=mode)
    s.init_problem(); s.step(2); s.set_profiling(True)
    ms = s.step_timed(5); mx,my,kx,ky = s.kernel_times()
    print(f"{mode:10s} n={n} {ms/5:8.3f} ms/step x={mx/kx:7.3f} y={my/ky:7.3f} {n*n*5/ms/1e6:6.3f} Gcu/s frac={n*n*5/(ms*1e-3)*432/6531e9:.3f}")
    s.close()
PY
timeout 200 python /tmp/qt_relaxed.py
