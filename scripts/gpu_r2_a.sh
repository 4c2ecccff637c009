#!/bin/bash
# round 2, GPU pass A: parity at the named sizes, ncu --set full of the Euler / TabEoS sweeps, sanitizers
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv,noheader > gpurun_out/a_smi.txt
timeout 1500 python -m pytest tests/test_named_sizes_gpu.py -m gpu -q -x --durations=10 > gpurun_out/a_named.log 2>&1; echo "named-size tests exit $?"
tail -15 gpurun_out/a_named.log
bash scripts/gpu_ncu_euler.sh 2>&1 | tail -6
bash scripts/gpu_ncu_tabeos.sh 2>&1 | tail -6
for r in euler_full tabeos_full; do
  [ -f gpurun_out/$r.ncu-rep ] && python scripts/ncu_summary.py gpurun_out/$r.ncu-rep euler_x_sweep_kernel euler_y_sweep_kernel > gpurun_out/${r}_summary.txt 2>&1
  [ -f gpurun_out/$r.ncu-rep ] && [ $(stat -c %s gpurun_out/$r.ncu-rep) -gt 28000000 ] && rm gpurun_out/$r.ncu-rep
done
SAN_TIMEOUT=900 bash scripts/gpu_sanitize.sh
ls -la gpurun_out
