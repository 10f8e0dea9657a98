mkdir -p gpurun_out
N=${NGPU:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
(timeout 600 python -m pytest tests/test_gpu_vs_reference.py tests/test_gpu_multi.py tests/test_gpu_direct.py -m gpu -x -q > gpurun_out/pytest14.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest14.log); tail -4 gpurun_out/pytest14.log
for ex in peer nccl; do
(NB_MGPU_EXCHANGE=$ex timeout 400 $TR --master-port 29624 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_${N}gpu_4mi_${ex}_v3.json 2> gpurun_out/bench_${N}gpu_4mi_${ex}_v3.err; echo "exit $?"); tail -1 gpurun_out/bench_${N}gpu_4mi_${ex}_v3.json | cut -c1-260
(NB_MGPU_EXCHANGE=$ex timeout 400 $TR --master-port 29625 bench.py --gpus $N --steps 20 --warmup 5 --bodies 131072 > gpurun_out/bench_${N}gpu_128k_${ex}_v3.json 2> gpurun_out/bench_${N}gpu_128k_${ex}_v3.err; echo "exit $?"); tail -1 gpurun_out/bench_${N}gpu_128k_${ex}_v3.json | cut -c1-260
done
