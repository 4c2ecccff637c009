#!/bin/bash
# full single-GPU pass for the round's artefacts: parity tests, headline bench (+ reference arm), secondary
# configs, ncu launch list + full capture of the two sweeps.  Every command under its own timeout.
mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench exit $?"
tail -1 gpurun_out/bench_n1.json | cut -c1-400
timeout 400 python bench.py --impl reference --steps 12 --warmup 1 > gpurun_out/bench_ref.json 2>&1; tail -1 gpurun_out/bench_ref.json | cut -c1-300
timeout 600 python scripts/bench_configs.py > gpurun_out/secondary_configs.jsonl 2> gpurun_out/secondary.err; cat gpurun_out/secondary_configs.jsonl | cut -c1-420
timeout 120 python scripts/quick_time_euler.py 2>&1 | tee gpurun_out/qt_euler.log
bash scripts/gpu_ncu.sh 4096
