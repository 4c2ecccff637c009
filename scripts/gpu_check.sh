#!/bin/bash
# last check of a build: all GPU tests, smoke(), quick MHD and Euler timings
mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 120 python scripts/quick_time.py 8192 6 | cut -c60-
timeout 120 python scripts/quick_time_euler.py 4096 60 hash | cut -c30-
