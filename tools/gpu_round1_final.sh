#!/bin/bash
# One box: parity of the direct path, the default bench, its ncu launch list and one full capture of the force kernel,
# then the rest of the GPU suite that runs the direct path.
mkdir -p gpurun_out
exec 2>&1
set -x
timeout 300 python -m pytest tests/test_gpu_direct.py -m gpu -x -q --durations=5 > gpurun_out/t_direct.log 2>&1
rc=$?; tail -5 gpurun_out/t_direct.log
[ $rc -ne 0 ] && exit 1
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err || { tail -20 gpurun_out/bench_default.err; exit 1; }
cat gpurun_out/bench_default.json
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference_arm.json 2>> gpurun_out/bench_default.err
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv \
  --log-file gpurun_out/direct1mi_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-reference-gpu \
  > gpurun_out/ncu_ll.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:direct_force_kernel -s 3 -c 1 \
  -o gpurun_out/r01_direct_1mi_onemass -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-reference-gpu \
  > gpurun_out/ncu_full.log 2>&1
timeout 420 python -m pytest tests/test_gpu_full_size.py tests/test_gpu_multi.py tests/test_gpu_dropin.py tests/test_gpu_graph.py \
  tests/test_gpu_cpp_host.py tests/test_gpu_reference_suite.py tests/test_gpu_vs_reference.py -m gpu -x -q --durations=12 \
  > gpurun_out/t_rest.log 2>&1
tail -25 gpurun_out/t_rest.log
python __graft_entry__.py smoke 2>&1 | tail -2
ls -la gpurun_out
