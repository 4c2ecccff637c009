#!/bin/bash
# round 2, GPU pass B: the device-resident time loop -- full GPU test tier, small-grid timings, bench at the named size
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -m gpu -q -x --durations=8 > gpurun_out/b_pytest.log 2>&1; echo "pytest exit $?"
tail -14 gpurun_out/b_pytest.log
timeout 300 python scripts/bench_configs.py c1 c2 > gpurun_out/b_c1c2.jsonl 2>&1; cat gpurun_out/b_c1c2.jsonl | cut -c1-600
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/b_bench_n1.json 2> gpurun_out/b_bench_n1.err; echo "bench exit $?"; cut -c1-1500 gpurun_out/b_bench_n1.json; tail -3 gpurun_out/b_bench_n1.err
