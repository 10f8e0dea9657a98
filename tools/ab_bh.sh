#!/bin/bash
# same-box A/B of Barnes-Hut traversal variants: tools/ab_bh.sh out.jsonl "tag[:WALK[:STACK]]" ...
# tag = default or the <tag> of a tools/build_variant.sh library; WALK = pairs | nodes (NB_BH_WALK); STACK = NB_BH_STACK
out=$1; shift; : > $out
for spec in "$@"; do
  IFS=: read -r tag walk stack <<< "$spec"
  lib=$PWD/nbody_b200/libnbody_b200_$tag.so; [ "$tag" = default ] && lib=$PWD/nbody_b200/libnbody_b200.so
  export NB_LIB_PATH=$lib
  if [ -n "$walk" ]; then export NB_BH_WALK=$walk; else unset NB_BH_WALK; fi
  if [ -n "$stack" ]; then export NB_BH_STACK=$stack; else unset NB_BH_STACK; fi
  for n in 16777216 1048576; do
    echo -n "{\"variant\": \"$spec\", \"n\": $n, \"result\": " >> $out
    python tools/bench_paths.py bh $n 2>/dev/null | tail -1 | python -c "import sys,json; r=json.loads(sys.stdin.read()); print(json.dumps({k:round(v['ms'],4) for k,v in r['stages'].items()}|{k:r[k] for k in ('build_ms','force_ms','step_ms','interactions_per_body','stack_deepest','stack_cap')}))" >> $out
    sed -i '$ s/$/}/' $out
  done
done
cut -c1-360 $out
