mkdir -p gpurun_out
N=${NGPU:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
(timeout 300 $TR --master-port 29621 tests/mgpu/bh_check.py 200003 3 > gpurun_out/bh_check_${N}_small.log 2>&1; echo "exit $?" >> gpurun_out/bh_check_${N}_small.log); grep -E "bh_check|exit|Error|error" gpurun_out/bh_check_${N}_small.log | tail -4
(timeout 400 $TR --master-port 29622 tests/mgpu/bh_check.py 16777216 4 > gpurun_out/bh_check_${N}_16mi.log 2>&1; echo "exit $?" >> gpurun_out/bh_check_${N}_16mi.log); grep -E "bh_check|exit|Error|error" gpurun_out/bh_check_${N}_16mi.log | tail -4
(timeout 300 $TR --master-port 29623 tests/mgpu/peer_check.py 100003 > gpurun_out/peer_check_${N}b.log 2>&1; echo "exit $?" >> gpurun_out/peer_check_${N}b.log); grep -E "peer_check|exit" gpurun_out/peer_check_${N}b.log | tail -2
for ex in peer nccl; do
(NB_MGPU_EXCHANGE=$ex timeout 400 $TR --master-port 29624 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_${N}gpu_4mi_${ex}_v2.json 2> gpurun_out/bench_${N}gpu_4mi_${ex}_v2.err; echo "exit $?"); tail -1 gpurun_out/bench_${N}gpu_4mi_${ex}_v2.json | cut -c1-260
done
