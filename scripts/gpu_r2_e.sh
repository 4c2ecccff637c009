#!/bin/bash
# round 2, GPU pass E (2 GPUs): the Euler loop device-resident -- full GPU test tier, Euler workloads at N = 1 / 2, drivers
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 1700 python -m pytest tests -m gpu -q -x > gpurun_out/e_pytest.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/e_pytest.log
run() { name=$1; shift; timeout 300 "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name exit $?"; grep '^{' gpurun_out/$name.json | cut -c1-250; grep -o '"state_hash": "[0-9a-f]*"' gpurun_out/$name.json; grep -o '"non_sweep_ms_per_step": [-0-9.e]*' gpurun_out/$name.json; tail -2 gpurun_out/$name.err; }
run e_euler_n1 python bench.py --workload euler --steps 100 --warmup 3 --no-cpu-baseline
run e_euler_n2 $TR --nproc-per-node 2 --master-port 29504 bench.py --workload euler --gpus 2 --steps 100 --warmup 3
run e_tabeos_n2 $TR --nproc-per-node 2 --master-port 29505 bench.py --workload tabeos --gpus 2 --grid 4096 --steps 10 --warmup 3
timeout 600 $TR --nproc-per-node 2 --master-port 29503 scripts/check_nccl_invariance.py 2>&1 | tail -2
( cd /tmp && timeout 300 $GRAFT_REPO_ROOT/drivers/_bin/slic_euler_2d --generalise --problem cylinder --nx 4096 --ny 4096 --tstop 0.004 --gpus 2 --no-dat --quiet --time ) 2>&1 | tail -2
( cd /tmp && timeout 300 $GRAFT_REPO_ROOT/drivers/_bin/slic_euler_2d --generalise --problem cylinder --nx 4096 --ny 4096 --tstop 0.004 --no-dat --quiet --time ) 2>&1 | tail -2
