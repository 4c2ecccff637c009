#!/bin/bash
# 2+ GPU pass for the halo/compute overlap: parity (1-GPU tests incl. the split y sweep, NCCL invariance),
# then the device-side breakdown serial vs overlapped
N=${1:-2}; SIZE=${2:-16384}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_mhd_gpu.py tests/test_euler.py -m gpu -x -q -k "split_y or nccl or out_of_window or slab_decomposition or tabeos" > gpurun_out/pytest_overlap.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_overlap.log; tail -4 gpurun_out/pytest_overlap.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 \
  scripts/check_nccl_invariance.py 2>&1 | grep -v "^W\|^\*" | tail -8 | tee gpurun_out/nccl_invariance_n$N.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29543 \
  scripts/comm_breakdown.py $SIZE 10 2>&1 | grep "rank" | sort | tee gpurun_out/comm_breakdown_n$N.log
