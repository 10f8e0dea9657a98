#!/bin/bash
# ncu launch list of the 16 Mi Barnes-Hut step and one --set full capture of the pair-walk traversal (gpurun, 1 GPU)
export PYTHONPATH=$PWD
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum
ncu --metrics $M --clock-control none -c 400 --csv --log-file gpurun_out/r02_bh16_launches_v8.csv python tools/bench_paths.py bh 16777216 plummer > /dev/null 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:bh_force_pairs_kernel -s 2 -c 1 -o gpurun_out/r02_bh_force_pairs_16mi_v8 -f python tools/bench_paths.py bh 16777216 plummer > gpurun_out/ncu_bh8.log 2>&1
ls -la gpurun_out | tail -5
