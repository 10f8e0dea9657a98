#!/bin/bash
# same-box A/B of radix-sort builds: tools/ab_sort.sh out.jsonl tag ...   (tag = default or a tools/build_variant.sh tag)
out=$1; shift; : > $out
for tag in "$@"; do
  lib=$PWD/nbody_b200/libnbody_b200_$tag.so; [ "$tag" = default ] && lib=$PWD/nbody_b200/libnbody_b200.so
  export NB_LIB_PATH=$lib
  if ! python -m pytest tests/test_gpu_sort_tree.py -m gpu -x -q -k "sort" > gpurun_out/sorttest_$tag.log 2>&1; then echo "{\"variant\": \"$tag\", \"tests\": \"FAILED\"}" >> $out; tail -5 gpurun_out/sorttest_$tag.log; continue; fi
  for args in "sort 16777216 30" "sort 10000000 21"; do
    echo -n "{\"variant\": \"$tag\", \"run\": \"$args\", \"result\": " >> $out
    python tools/bench_paths.py $args 2>/dev/null | tail -1 >> $out
    sed -i '$ s/$/}/' $out
  done
  echo -n "{\"variant\": \"$tag\", \"run\": \"bh 16777216 stages\", \"result\": " >> $out
  python tools/bench_paths.py bh 16777216 2>/dev/null | tail -1 | python -c "import sys,json; r=json.loads(sys.stdin.read()); print(json.dumps({k:round(v['ms'],4) for k,v in r['stages'].items()}|{'build_ms':r['build_ms'],'step_ms':r['step_ms']}))" >> $out
  sed -i '$ s/$/}/' $out
done
python - "$out" <<'PY'
import json,sys
for ln in open(sys.argv[1]):
    r=json.loads(ln)
    if 'tests' in r: print(r); continue
    q=r['result']
    print(r['variant'], r['run'], {k:(round(v,4) if isinstance(v,float) else v) for k,v in q.items() if k in ('ms','frac','bh.sort','bh.tree','build_ms','step_ms')})
PY
