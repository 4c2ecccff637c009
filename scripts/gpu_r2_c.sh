#!/bin/bash
# round 2, GPU pass C: CUDA-graph replay of step batches -- tests, small-grid timings with and without graphs
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/c_pytest.log 2>&1; echo "pytest exit $?"
tail -4 gpurun_out/c_pytest.log
timeout 300 python scripts/quick_small.py > gpurun_out/c_small_graphs.jsonl 2>&1; cat gpurun_out/c_small_graphs.jsonl
LSC_NO_GRAPH=1 timeout 300 python scripts/quick_small.py > gpurun_out/c_small_nographs.jsonl 2>&1; cat gpurun_out/c_small_nographs.jsonl
cd /tmp && timeout 120 $GRAFT_REPO_ROOT/drivers/_bin/muscl_new --problem ot --nx 256 --ny 256 --y0 0 --bcy periodic --tstop 0.5 --quiet --no-dat --time 2>&1 | tail -2
