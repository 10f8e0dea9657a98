#!/bin/bash
# The three captures profiles/ holds for the headline path (run under gpurun, one GPU): the default bench command as is,
# its ncu launch list with DRAM bytes, and one ncu --set full capture of the dominant kernel at the full 1 Mi size.
set -x
mkdir -p gpurun_out
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err || exit 1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv \
  --log-file gpurun_out/direct1mi_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-reference-gpu \
  > gpurun_out/ncu_ll.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:direct_force_kernel -s 3 -c 1 \
  -o gpurun_out/r01_direct_1mi -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-reference-gpu \
  > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out
