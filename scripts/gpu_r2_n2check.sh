#!/bin/bash
# 2 GPUs: NCCL slab path after the Euler first-step ghost fix (library ICs and uploaded random states), the 2-GPU
# pytest subset, Euler bench hash at N=2 vs N=1
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 600 $TR --nproc-per-node 2 --master-port 29503 scripts/check_nccl_invariance.py > gpurun_out/n2_nccl_invariance.log 2>&1; echo "invariance exit $?"; tail -3 gpurun_out/n2_nccl_invariance.log
timeout 600 python -m pytest tests -m gpu -q -k "nccl" 2>&1 | tail -2
for n in 2 1; do
  if [ $n = 2 ]; then cmd="$TR --nproc-per-node 2 --master-port 29504 bench.py --workload euler --gpus 2"; else cmd="python bench.py --workload euler --no-cpu-baseline"; fi
  timeout 300 $cmd --grid 2048 --steps 10 --warmup 3 > gpurun_out/n2_euler_n$n.json 2> gpurun_out/n2_euler_n$n.err; echo "euler n=$n exit $?"
  grep -o '"value": [0-9.e+]*' gpurun_out/n2_euler_n$n.json | head -1; grep -o '"state_hash": "[0-9a-f]*"' gpurun_out/n2_euler_n$n.json
done
