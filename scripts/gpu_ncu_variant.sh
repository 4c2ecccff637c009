#!/bin/bash
# ncu --set full of one kernel of an experimental build: gpu_ncu_variant.sh LIBDIR KERNEL_REGEX [N]
LIBDIR=$1; K=$2; N=${3:-4096}
mkdir -p gpurun_out
export LSC_B200_LIB=$PWD/$LIBDIR/liblsc_b200.so
timeout 120 python scripts/quick_time.py $N 3 | cut -c60- || exit 1
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:$K" -s 2 -c 1 -o gpurun_out/variant_full -f \
  python scripts/quick_time.py $N 3 > gpurun_out/ncu_variant.log 2>&1
tail -2 gpurun_out/ncu_variant.log
