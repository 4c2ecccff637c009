#!/bin/bash
# one GPU-box pass: parity tests, bench at the named size, ncu launch list + one --set full capture
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench exit $?"
cat gpurun_out/bench_n1.json
python bench.py --impl reference --steps 12 --warmup 1 > gpurun_out/bench_ref.json 2>&1
python scripts/quick_time.py 8192 5 > gpurun_out/qt_mhd.log 2>&1; cat gpurun_out/qt_mhd.log
python scripts/quick_time_euler.py > gpurun_out/qt_euler.log 2>&1; cat gpurun_out/qt_euler.log
python bench.py --n 4096 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/b4096.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_ot4096.csv \
  python bench.py --n 4096 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:sweep -s 4 -c 2 -o gpurun_out/sweeps_full -f \
  python bench.py --n 4096 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out
