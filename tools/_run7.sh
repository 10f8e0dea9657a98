mkdir -p gpurun_out
N=${NGPU:-8}
(timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 tests/mgpu/peer_check.py 100003 > gpurun_out/peer_check_$N.log 2>&1; echo "exit $?" >> gpurun_out/peer_check_$N.log); tail -3 gpurun_out/peer_check_$N.log
for ex in peer nccl; do
(NB_MGPU_EXCHANGE=$ex timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus $N --steps 10 --warmup 3 --bodies 262144 > gpurun_out/bench_${N}gpu_256k_$ex.json 2> gpurun_out/bench_${N}gpu_256k_$ex.err; echo "exit $?"); tail -1 gpurun_out/bench_${N}gpu_256k_$ex.json | cut -c1-300
(NB_MGPU_EXCHANGE=$ex timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29613 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_${N}gpu_4mi_$ex.json 2> gpurun_out/bench_${N}gpu_4mi_$ex.err; echo "exit $?"); tail -1 gpurun_out/bench_${N}gpu_4mi_$ex.json | cut -c1-300
done
tail -3 gpurun_out/bench_${N}gpu_4mi_peer.err
