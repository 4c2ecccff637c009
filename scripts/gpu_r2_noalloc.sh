#!/bin/bash
# (1) the allocation-free entry point's test alone in a fresh process, (2) the whole GPU suite (it runs last there),
# (3) the LSC_SCALED_GUARDS=1 build: subnormal KATs, whole suite, A/B timing against the default build.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/noalloc_alone.log 2>&1
echo "noalloc alone: exit $?"; tail -3 gpurun_out/noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu_noalloc.log 2>&1
echo "gpu suite: exit $?"; tail -3 gpurun_out/pytest_gpu_noalloc.log
V1=$PWD/lsc_internship_b200/_lib_v1/liblsc_b200.so
if [ -f $V1 ]; then
  LSC_B200_LIB=$V1 timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu_guards1.log 2>&1
  echo "gpu suite, LSC_SCALED_GUARDS=1: exit $?"; tail -3 gpurun_out/pytest_gpu_guards1.log
  bash scripts/gpu_ab.sh 3 6 euler > gpurun_out/ab_guards1.log 2>&1
  grep -v "^$" gpurun_out/ab_guards1.log | tail -24
fi
