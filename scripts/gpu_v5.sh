#!/bin/bash
# v5 pass in one box: parity tests on the production build, A/B against the control build(s) in
# lsc_internship_b200/_lib_v*/ (hash at 1024^2, two alternating timing passes at 8192^2), headline bench +
# reference arm, then the ncu launch list and the full capture of the two sweeps.  Every command under its own timeout.
mkdir -p gpurun_out
timeout 420 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
: > gpurun_out/variants.log
for pass in 1 2; do
for d in lsc_internship_b200/_lib lsc_internship_b200/_lib_v*; do
  [ -f $d/liblsc_b200.so ] || continue
  echo "## $d $(cat $d/flags.txt 2>/dev/null)"
  [ $pass = 1 ] && LSC_B200_LIB=$PWD/$d/liblsc_b200.so timeout 60 python scripts/quick_time.py 1024 3 hash | cut -c60-
  LSC_B200_LIB=$PWD/$d/liblsc_b200.so timeout 120 python scripts/quick_time.py 8192 6 | cut -c60-
done 2>&1 | tee -a gpurun_out/variants.log
done
timeout 500 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench exit $?"
tail -1 gpurun_out/bench_n1.json | cut -c1-300
timeout 200 python bench.py --impl reference --steps 12 --warmup 1 > gpurun_out/bench_ref.json 2>&1; tail -1 gpurun_out/bench_ref.json | cut -c1-200
timeout 500 python scripts/bench_configs.py > gpurun_out/secondary_configs.jsonl 2> gpurun_out/secondary.err; cut -c1-300 gpurun_out/secondary_configs.jsonl
bash scripts/gpu_ncu.sh 4096
