#!/bin/bash
# non-periodic y boundary path: MHD parity tests, then KH 1024^2 (config c2) and OT 8192^2 for the production build
# and every experimental build
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_mhd_gpu.py -m gpu -x -q > gpurun_out/pytest_mhd.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_mhd.log
tail -3 gpurun_out/pytest_mhd.log
for d in lsc_internship_b200/_lib lsc_internship_b200/_lib_v*; do
  [ -f $d/liblsc_b200.so ] || continue
  echo "## $d $(cat $d/flags.txt 2>/dev/null)"
  LSC_B200_LIB=$PWD/$d/liblsc_b200.so timeout 120 python scripts/bench_configs.py c2 | cut -c1-330
  LSC_B200_LIB=$PWD/$d/liblsc_b200.so timeout 120 python scripts/quick_time.py 8192 6 | cut -c60-
done 2>&1 | tee gpurun_out/kh.log
