#!/bin/bash
# ncu --set full capture of the two Euler sweeps (cylinder 4096^2), after the same command ran clean without ncu
mkdir -p gpurun_out
timeout 200 python scripts/quick_time_euler.py > gpurun_out/qt_euler.log 2>&1 || { echo "euler timing failed"; tail -5 gpurun_out/qt_euler.log; exit 1; }
cat gpurun_out/qt_euler.log
timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:^euler_[xy]_sweep_kernel$' -s 4 -c 2 -o gpurun_out/euler_full -f \
  python scripts/quick_time_euler.py > gpurun_out/ncu_euler.log 2>&1
tail -3 gpurun_out/ncu_euler.log
ls -la gpurun_out/euler_full.ncu-rep
