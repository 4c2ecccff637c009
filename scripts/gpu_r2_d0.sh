#!/bin/bash
# 2-GPU smoke of everything pass D runs on 8 (small grids): catches scripting errors at 2x instead of 8x the price
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
run() { name=$1; shift; timeout 300 "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name exit $?"; cut -c1-260 gpurun_out/$name.json; grep -o '"state_hash": "[0-9a-f]*"' gpurun_out/$name.json; grep -o '"slabs": {[^}]*}' gpurun_out/$name.json; tail -2 gpurun_out/$name.err; }
run d0_bench_n2 $TR --nproc-per-node 2 --master-port 29501 bench.py --gpus 2 --grid 4096 --steps 6 --warmup 3
run d0_bench_n2_even $TR --nproc-per-node 2 --master-port 29502 bench.py --gpus 2 --grid 4096 --steps 6 --warmup 3 --no-balance
run d0_bench_n1 python bench.py --grid 4096 --steps 6 --warmup 3 --no-cpu-baseline
( cd /tmp && timeout 300 $GRAFT_REPO_ROOT/drivers/_bin/muscl_new --problem ot --nx 4096 --ny 4096 --y0 0 --bcy periodic --steps 9 --gpus 2 --no-dat --quiet --time ) 2>&1 | tail -2
timeout 600 $TR --nproc-per-node 2 --master-port 29503 scripts/check_nccl_invariance.py 2>&1 | tail -2
timeout 900 python -m pytest tests/test_mhd_gpu.py -m gpu -q -k "nccl or same_program or group_run or slab" 2>&1 | tail -3
run d0_euler_n2 $TR --nproc-per-node 2 --master-port 29504 bench.py --workload euler --gpus 2 --grid 2048 --steps 10 --warmup 3
run d0_tabeos_n2 $TR --nproc-per-node 2 --master-port 29505 bench.py --workload tabeos --gpus 2 --grid 1024 --steps 5 --warmup 3
run d0_euler_n1 python bench.py --workload euler --grid 2048 --steps 10 --warmup 3
run d0_tabeos_n1 python bench.py --workload tabeos --grid 1024 --steps 5 --warmup 3
( cd /tmp && timeout 300 $GRAFT_REPO_ROOT/drivers/_bin/slic_euler_2d --generalise --problem cylinder --nx 2048 --ny 2048 --tstop 0.002 --gpus 2 --no-dat --quiet --time ) 2>&1 | tail -2
timeout 300 python scripts/quick_small.py 2>&1 | cut -c1-300
