#!/bin/bash
# ncu --set full capture of the tabulated-EoS Euler sweeps (c5 at 2048^2), after the same command ran clean without ncu
mkdir -p gpurun_out
LSC_C5_N=2048 timeout 300 python scripts/bench_configs.py c5 > gpurun_out/c5_2048.log 2>&1 || { echo "c5 failed"; tail -5 gpurun_out/c5_2048.log; exit 1; }
cut -c1-400 gpurun_out/c5_2048.log
LSC_C5_N=2048 timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:^euler_[xy]_sweep_kernel$' -s 6 -c 2 -o gpurun_out/tabeos_full -f \
  python scripts/bench_configs.py c5 > gpurun_out/ncu_tabeos.log 2>&1
tail -3 gpurun_out/ncu_tabeos.log
ls -la gpurun_out/tabeos_full.ncu-rep
