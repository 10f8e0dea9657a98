mkdir -p gpurun_out
N=${NGPU:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
(timeout 400 $TR --master-port 29622 tests/mgpu/bh_check.py 16777216 4 > gpurun_out/bh_check_${N}_16mi.log 2>&1; echo "exit $?" >> gpurun_out/bh_check_${N}_16mi.log); grep -E "bh_check|exit|Error|error" gpurun_out/bh_check_${N}_16mi.log | tail -4
for ex in peer nccl peer; do
(NB_MGPU_EXCHANGE=$ex timeout 400 $TR --master-port 29624 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_${N}gpu_4mi_${ex}_v3.json 2> gpurun_out/bench_${N}gpu_4mi_${ex}_v3.err; echo "exit $?"); python - gpurun_out/bench_${N}gpu_4mi_${ex}_v3.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1].split('/')[-1], round(d['ms_per_step'],1), round(d['value']), 'e2e', round(d['e2e']['value']), 'kernel_ms(rank0)', round(d['roofline']['kernel_ms'],1), d['clocks'])
PY
done
