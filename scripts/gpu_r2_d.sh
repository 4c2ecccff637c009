#!/bin/bash
# round 2, GPU pass D (8 x B200, one box): NCCL bench with calibrated / even slabs, the drop-in program on 8 GPUs in
# ONE process (J5), NCCL bit-invariance, the Euler / TabEoS workloads under torchrun, state hashes for N = 1 and 8
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
nvidia-smi --query-gpu=index,name,clocks.sm,power.draw --format=csv,noheader > gpurun_out/d_smi.txt
run() { name=$1; shift; timeout 600 "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name exit $?"; cut -c1-330 gpurun_out/$name.json; grep -o '"state_hash": "[0-9a-f]*"' gpurun_out/$name.json; }
run d_bench_n8 $TR --nproc-per-node 8 --master-port 29501 bench.py --gpus 8 --steps 20 --warmup 3
run d_bench_n8_balanced $TR --nproc-per-node 8 --master-port 29502 bench.py --gpus 8 --steps 20 --warmup 3 --balance
run d_bench_n4 $TR --nproc-per-node 4 --master-port 29506 bench.py --gpus 4 --steps 20 --warmup 3
run d_bench_n2 $TR --nproc-per-node 2 --master-port 29507 bench.py --gpus 2 --steps 20 --warmup 3
( cd /tmp && timeout 600 $GRAFT_REPO_ROOT/drivers/_bin/muscl_new --problem ot --nx 16384 --ny 16384 --y0 0 --bcy periodic --steps 23 --gpus 8 --no-dat --quiet --time ) > gpurun_out/d_driver_gpus8.log 2>&1; echo "driver exit $?"; cat gpurun_out/d_driver_gpus8.log
timeout 600 $TR --nproc-per-node 8 --master-port 29503 scripts/check_nccl_invariance.py > gpurun_out/d_nccl_invariance_n8.log 2>&1; tail -2 gpurun_out/d_nccl_invariance_n8.log
timeout 900 python -m pytest tests/test_mhd_gpu.py tests/test_euler.py -m gpu -q -k "nccl or same_program or group_run or slab" > gpurun_out/d_pytest_multi.log 2>&1; tail -3 gpurun_out/d_pytest_multi.log
run d_euler_n8 $TR --nproc-per-node 8 --master-port 29504 bench.py --workload euler --gpus 8 --steps 100 --warmup 3
run d_tabeos_n8 $TR --nproc-per-node 8 --master-port 29505 bench.py --workload tabeos --gpus 8 --steps 20 --warmup 3
run d_euler_n1 python bench.py --workload euler --steps 100 --warmup 3
run d_tabeos_n1 python bench.py --workload tabeos --steps 20 --warmup 3
( cd /tmp && timeout 300 $GRAFT_REPO_ROOT/drivers/_bin/slic_euler_2d --generalise --problem cylinder --nx 4096 --ny 4096 --tstop 0.02 --gpus 8 --no-dat --quiet --time ) > gpurun_out/d_euler_driver_gpus8.log 2>&1; cat gpurun_out/d_euler_driver_gpus8.log
( cd /tmp && timeout 300 $GRAFT_REPO_ROOT/drivers/_bin/muscl_new --tstop 0.25 --gpus 8 --outdir /tmp > /tmp/drv_stdout.txt 2>/dev/null; cmp /tmp/drv_stdout.txt $GRAFT_REPO_ROOT/tests/golden/muscl_new_stdout_t025.txt && echo "muscl_new --gpus 8: stdout identical to the reference program's"; sha256sum /tmp/MHD0.dat /tmp/MHD3.dat ) > gpurun_out/d_driver_gpus8_dropin.log 2>&1; cat gpurun_out/d_driver_gpus8_dropin.log
run d_bench_n1 python bench.py --steps 20 --warmup 3 --no-cpu-baseline
ls gpurun_out | head -40
