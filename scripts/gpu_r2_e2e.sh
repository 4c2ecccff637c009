#!/bin/bash
# after the two-half staging of lsc_mhd_upload / download: the GPU suite, then the bench line (e2e is the point)
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -m gpu -q -x > gpurun_out/e2e_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/e2e_pytest.log
timeout 900 python bench.py > gpurun_out/e2e_bench_n1.json 2> gpurun_out/e2e_bench_n1.err; echo "bench exit $?"
grep -o '"value": [0-9.e+]*' gpurun_out/e2e_bench_n1.json | head -2; grep -o '"seconds": [0-9.]*' gpurun_out/e2e_bench_n1.json; grep -o '"state_hash": "[0-9a-f]*"' gpurun_out/e2e_bench_n1.json; grep -o '"sm_mhz": [0-9.]*' gpurun_out/e2e_bench_n1.json
