#!/bin/bash
# ncu launch lists (with DRAM bytes) of the Barnes-Hut 16 Mi and spatial-hash 10 M steps and one full capture of the
# Barnes-Hut traversal kernel at 1 Mi; run under gpurun on one GPU, outputs in gpurun_out/.
export PYTHONPATH=$PWD
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum
python tools/bench_paths.py bh 16777216 plummer | tail -1 > gpurun_out/paths_v3.jsonl || exit 1
python tools/bench_paths.py sh 10000000 | tail -1 >> gpurun_out/paths_v3.jsonl || exit 1
python tools/bench_paths.py sort 16777216 30 | tail -1 >> gpurun_out/paths_v3.jsonl
ncu --metrics $M --clock-control none -c 400 --csv --log-file gpurun_out/bh16_launches.csv python tools/bench_paths.py bh 16777216 plummer > /dev/null 2>&1
ncu --metrics $M --clock-control none -c 400 --csv --log-file gpurun_out/sh10_launches.csv python tools/bench_paths.py sh 10000000 > /dev/null 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:bh_force_kernel -s 2 -c 1 -o gpurun_out/bh_force_1mi_v3 -f python tools/bench_paths.py bh 1048576 plummer > gpurun_out/ncu_bh.log 2>&1
ls -la gpurun_out | tail -8
