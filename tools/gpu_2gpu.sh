#!/bin/bash
mkdir -p gpurun_out; exec 2>&1; set -x
timeout 200 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/t_multi2.log 2>&1; tail -3 gpurun_out/t_multi2.log
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 2 --warmup 3 > gpurun_out/bench_2gpu.json 2> gpurun_out/bench_2gpu.err
cat gpurun_out/bench_2gpu.json; tail -3 gpurun_out/bench_2gpu.err
