mkdir -p gpurun_out; rm -f gpurun_out/paths5.jsonl
(timeout 600 python -m pytest tests/test_gpu_vs_reference.py -m gpu -x -q -k "peer or sharded" > gpurun_out/pytest5.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest5.log); tail -4 gpurun_out/pytest5.log
(NB_SORT_MATCH=1 timeout 600 python -m pytest tests/test_gpu_sort_tree.py -m gpu -x -q -k "radix" > gpurun_out/pytest5b.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest5b.log); tail -3 gpurun_out/pytest5b.log
for m in 0 1; do for w in 16 8; do NB_SORT_MATCH=$m NB_SORT_WARPS=$w python tools/bench_paths.py sort 16777216 30 >> gpurun_out/paths5.jsonl 2>>gpurun_out/paths5.err; done; done
NB_SORT_MATCH=1 python tools/bench_paths.py sort 10000000 21 >> gpurun_out/paths5.jsonl 2>>gpurun_out/paths5.err
python - <<'PY'
import json
for l in open('gpurun_out/paths5.jsonl'):
    d=json.loads(l); d.pop('ms_all',None); print(d)
PY
tail -3 gpurun_out/paths5.err
