#!/bin/bash
# same-box A/B of the radix-sort pass variants
out=gpurun_out/r02_sort_register_diet_ab.jsonl; : > $out
for tag in "" _rsold _rs3 _rs3b4 _rs2b16 "" _rsold _rs3; do
  export NB_LIB_PATH=$PWD/nbody_b200/libnbody_b200$tag.so
  if ! python -m pytest tests/test_gpu_sort_tree.py -m gpu -x -q -k "sort" > gpurun_out/sorttest$tag.log 2>&1; then echo "{\"variant\": \"$tag\", \"tests\": \"FAILED\"}" >> $out; tail -5 gpurun_out/sorttest$tag.log; continue; fi
  for args in "sort 16777216 30" "sort 10000000 21"; do
    echo -n "{\"variant\": \"${tag:-default}\", \"run\": \"$args\", \"result\": " >> $out
    python tools/bench_paths.py $args 2>/dev/null | tail -1 >> $out
    sed -i '$ s/$/}/' $out
  done
  echo -n "{\"variant\": \"${tag:-default}\", \"run\": \"bh 16777216 stages\", \"result\": " >> $out
  python tools/bench_paths.py bh 16777216 2>/dev/null | tail -1 | python -c "import sys,json; r=json.loads(sys.stdin.read()); print(json.dumps({k:round(v['ms'],4) for k,v in r['stages'].items()}|{'build_ms':r['build_ms'],'step_ms':r['step_ms']}))" >> $out
  sed -i '$ s/$/}/' $out
done
cat $out | cut -c1-400
