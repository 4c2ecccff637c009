#!/bin/bash
# A/B of the production library against experimental builds lsc_internship_b200/_lib_v*/ : state hash at 1024^2,
# then alternating timing passes at 8192^2 (args: passes, steps)
mkdir -p gpurun_out
: > gpurun_out/variants.log
for pass in $(seq 1 ${1:-2}); do
for d in lsc_internship_b200/_lib lsc_internship_b200/_lib_v*; do
  [ -f $d/liblsc_b200.so ] || continue
  echo "## $d $(cat $d/flags.txt 2>/dev/null)"
  [ $pass = 1 ] && LSC_B200_LIB=$PWD/$d/liblsc_b200.so timeout 60 python scripts/quick_time.py 1024 3 hash | cut -c60-
  LSC_B200_LIB=$PWD/$d/liblsc_b200.so timeout 120 python scripts/quick_time.py 8192 ${2:-6} | cut -c60-
  [ -n "$3" ] && LSC_B200_LIB=$PWD/$d/liblsc_b200.so timeout 120 python scripts/quick_time_euler.py 4096 60 hash | cut -c30-
done 2>&1 | tee -a gpurun_out/variants.log
done
