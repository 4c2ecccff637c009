#!/bin/bash
# ncu passes for the two production sweeps (after the same command ran clean without ncu)
N=${1:-4096}
mkdir -p gpurun_out
timeout 300 python bench.py --n $N --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/b$N.log 2>&1 || { echo "bench failed"; tail -5 gpurun_out/b$N.log; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_ot$N.csv \
  python bench.py --n $N --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:^[xy]_sweep_kernel$' -s 2 -c 2 -o gpurun_out/sweeps_full -f \
  python bench.py --n $N --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
ls -la gpurun_out/*.ncu-rep
