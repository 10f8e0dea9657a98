mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
(timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 tests/mgpu/peer_check.py 100003 > gpurun_out/peer_check.log 2>&1; echo "exit $?" >> gpurun_out/peer_check.log); tail -5 gpurun_out/peer_check.log
for ex in peer nccl; do
(NB_MGPU_EXCHANGE=$ex timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus 2 --steps 5 --warmup 3 --bodies 262144 > gpurun_out/bench_2gpu_256k_$ex.json 2> gpurun_out/bench_2gpu_256k_$ex.err; echo "exit $?"); tail -1 gpurun_out/bench_2gpu_256k_$ex.json | cut -c1-400
done
(timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29613 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench_2gpu_4mi_peer.json 2> gpurun_out/bench_2gpu_4mi_peer.err; echo "exit $?"); tail -1 gpurun_out/bench_2gpu_4mi_peer.json | cut -c1-600
tail -5 gpurun_out/bench_2gpu_4mi_peer.err
