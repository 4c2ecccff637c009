#!/bin/bash
# compare experimental builds (make OUT=../_lib_vN) with the production library: state hash at 1024^2, sweep times
# at 8192^2 (two alternating passes), and the MHD parity tests against each variant
mkdir -p gpurun_out
: > gpurun_out/variants.log
for pass in 1 2; do
for d in lsc_internship_b200/_lib lsc_internship_b200/_lib_v*; do
  [ -f $d/liblsc_b200.so ] || continue
  echo "## $d $(cat $d/flags.txt 2>/dev/null)"
  [ $pass = 1 ] && LSC_B200_LIB=$PWD/$d/liblsc_b200.so timeout 60 python scripts/quick_time.py 1024 2 hash | cut -c60-
  LSC_B200_LIB=$PWD/$d/liblsc_b200.so timeout 120 python scripts/quick_time.py ${1:-8192} ${2:-5} | cut -c60-
done 2>&1 | tee -a gpurun_out/variants.log
done
for d in lsc_internship_b200/_lib_v*; do
  [ -f $d/liblsc_b200.so ] || continue
  echo "## parity $d"; LSC_B200_LIB=$PWD/$d/liblsc_b200.so timeout 300 python -m pytest tests/test_mhd_gpu.py -m gpu -x -q -k "not nccl and not driver and not division" 2>&1 | tail -2
done 2>&1 | tee -a gpurun_out/variants.log
