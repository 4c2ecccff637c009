#!/bin/bash
# 2-GPU check of the v5 kernels: NCCL parity test, NCCL invariance script, headline bench at N=2
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_mhd_gpu.py tests/test_euler.py -m gpu -x -q -k "nccl or split_y" > gpurun_out/pytest_n2.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_n2.log; tail -3 gpurun_out/pytest_n2.log
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 \
  scripts/check_nccl_invariance.py 2>&1 | grep -v "^W\|^\*" | tail -4 | tee gpurun_out/nccl_invariance_n2.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29545 \
  bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench exit $?"
tail -1 gpurun_out/bench_n2.json | cut -c1-200
