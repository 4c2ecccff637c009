#!/bin/bash
# Round-2 follow-up: (1) the subnormal-scaling KATs against the previous build (expected to FAIL there: shows the
# tests see the double rounding) and against the current build; (2) A/B timing of the two builds; (3) the GPU suite.
mkdir -p gpurun_out
HEADLIB=$PWD/lsc_internship_b200/_lib_v0head/liblsc_b200.so
if [ -f $HEADLIB ]; then
  LSC_B200_LIB=$HEADLIB timeout 600 python -m pytest tests/test_mhd_gpu.py tests/test_euler.py -q -m gpu -k subnormal \
      > gpurun_out/subnormal_kat_previous_build.log 2>&1
  echo "previous build: exit $?" | tee -a gpurun_out/subnormal_kat_previous_build.log
fi
timeout 600 python -m pytest tests/test_mhd_gpu.py tests/test_euler.py -q -m gpu -k subnormal \
    > gpurun_out/subnormal_kat_current_build.log 2>&1
echo "current build: exit $?" | tee -a gpurun_out/subnormal_kat_current_build.log
bash scripts/gpu_ab.sh 3 6 euler > gpurun_out/ab_guard.log 2>&1
tail -40 gpurun_out/ab_guard.log
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu_guard.log 2>&1
echo "gpu suite: exit $?"; tail -5 gpurun_out/pytest_gpu_guard.log
