#!/bin/bash
# round 2, final single-GPU pass: what the driver runs at round end (GPU tests, smoke, both bench arms)
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -m gpu -q -x > gpurun_out/z_pytest.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/z_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 600 python bench.py --impl reference --steps 8 --warmup 1 > gpurun_out/z_bench_reference.json 2>&1; cut -c1-200 gpurun_out/z_bench_reference.json
timeout 900 python bench.py > gpurun_out/z_bench_n1.json 2> gpurun_out/z_bench_n1.err; echo "bench exit $?"; cut -c1-400 gpurun_out/z_bench_n1.json; tail -2 gpurun_out/z_bench_n1.err
