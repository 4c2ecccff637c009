#!/bin/bash
# short GPU pass: parity tests, quick timings, optional microbenchmarks (every command under its own timeout)
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
for d in lsc_internship_b200/_lib lsc_internship_b200/_lib_v*; do
  [ -f $d/liblsc_b200.so ] || continue
  echo "## $d $(cat $d/flags.txt 2>/dev/null)"
  LSC_B200_LIB=$PWD/$d/liblsc_b200.so timeout 60 python scripts/quick_time.py 1024 2 hash | cut -c60-
  LSC_B200_LIB=$PWD/$d/liblsc_b200.so timeout 120 python scripts/quick_time.py ${1:-8192} ${2:-5} | cut -c60-
  LSC_B200_LIB=$PWD/$d/liblsc_b200.so timeout 120 python scripts/quick_time_euler.py 4096 50 hash | cut -c30-
done 2>&1 | tee gpurun_out/variants.log
timeout 120 python scripts/quick_time_euler.py 2>&1 | tee gpurun_out/qt_euler.log
if [ -x scripts/microbench/fp64_pipe ]; then timeout 120 scripts/microbench/fp64_pipe 2>&1 | tee gpurun_out/fp64_pipe.log; fi
