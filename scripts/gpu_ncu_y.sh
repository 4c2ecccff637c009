mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k 'regex:^y_sweep_kernel$' -s 2 -c 1 -o gpurun_out/y_tma -f python bench.py --grid 4096 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_y.log 2>&1
tail -2 gpurun_out/ncu_y.log
