#!/bin/bash
# final evidence for round 2: launch list of the default bench command, ncu --set full of one sort pass
export PYTHONPATH=$PWD
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum
python -m pytest tests/test_gpu_sort_tree.py -m gpu -x -q -k "degenerate" 2>&1 | tail -6
ncu --metrics $M --clock-control none -c 600 --csv --log-file gpurun_out/r02_bench_default_launches_v9.csv python bench.py --steps 2 --warmup 3 --no-paths --no-parity --no-cpu-baseline > gpurun_out/ncu_bench9.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:rs_pass_kernel -s 4 -c 1 -o gpurun_out/r02_sort_pass_v9 -f python tools/bench_paths.py sort 16777216 30 > gpurun_out/ncu_sort9.log 2>&1
ls -la gpurun_out | tail -4
