#!/bin/bash
# round 2, GPU pass H: first-order sibling scheme (N4) parity + a timing line
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_hllc2d.py -m gpu -q -x > gpurun_out/h_pytest.log 2>&1; echo "pytest exit $?"; tail -15 gpurun_out/h_pytest.log
python - <<'PY' 2>&1 | tee gpurun_out/h_timing.log
import time, lsc_internship_b200 as lsc
for n in (800, 4096):
    s = lsc.HllcMhd2d(n, n); s.init_problem(); s.step(3)
    t0 = time.perf_counter(); s.step(30); dt = time.perf_counter() - t0
    print(f"hllc2d OT {n}^2: {dt/30*1e3:.3f} ms/step, {n*n*30/dt/1e9:.2f} G cell-updates/s (host-synced loop, one HLLC flux per interface)")
    s.close()
PY
