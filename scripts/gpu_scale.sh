#!/bin/bash
# strong-scaling run of bench.py at N ranks (launched like the driver does)
N=${1:-2}; STEPS=${2:-10}; SIZE=${3:-16384}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_n$N.txt 2>&1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
  bench.py --gpus $N --steps $STEPS --warmup 3 --grid $SIZE > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
echo "exit $?"; tail -2 gpurun_out/bench_n$N.err; cat gpurun_out/bench_n$N.json
