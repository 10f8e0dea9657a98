#!/bin/bash
# tools/mgpu_bh.sh [notest] : on a multi-GPU box, the multi-GPU tests and `bench.py --path bh` on all GPUs and on one
N=$(python -c "import torch; print(torch.cuda.device_count())")
echo "gpus: $N"
[ "$1" = notest ] || python -m pytest tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -3
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29701 bench.py --gpus $N --path bh --steps 10 --warmup 3 > gpurun_out/r02_bench_bh_16mi_${N}gpu_v2.json 2> gpurun_out/bh_${N}gpu.err; tail -c 700 gpurun_out/r02_bench_bh_16mi_${N}gpu_v2.json
[ "$1" = notest ] || { python bench.py --path bh --steps 10 --warmup 3 > gpurun_out/r02_bench_bh_16mi_1gpu_v3.json 2>/dev/null; tail -c 800 gpurun_out/r02_bench_bh_16mi_1gpu_v3.json; }
