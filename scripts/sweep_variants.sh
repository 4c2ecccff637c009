#!/bin/bash
# time every experimental build lsc_internship_b200/_lib_v*/ (development aid)
N=${1:-8192}; STEPS=${2:-4}
for d in lsc_internship_b200/_lib lsc_internship_b200/_lib_v*; do
  [ -f $d/liblsc_b200.so ] || continue
  echo "## $d $(cat $d/flags.txt 2>/dev/null)"
  LSC_B200_LIB=$PWD/$d/liblsc_b200.so python scripts/quick_time.py 1024 2 hash | cut -c60-
  LSC_B200_LIB=$PWD/$d/liblsc_b200.so python scripts/quick_time.py $N $STEPS | cut -c60-
done
