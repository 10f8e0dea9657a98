#!/bin/bash
# same-box A/B of spatial-hash builds at 10 M bodies, after the grid tests on the default build:
#   TAGS="default <tag> ..." OUT=name.jsonl tools/ab_sh.sh      (<tag> = a tools/build_variant.sh library)
out=gpurun_out/${OUT:-r02_sh_force_ab.jsonl}; : > $out
python -m pytest tests/test_gpu_grid.py tests/test_gpu_full_size_parity.py tests/test_gpu_vs_reference.py tests/test_gpu_reference_suite.py -m gpu -x -q -k "sh or grid or spatial or hash" 2>&1 | tail -4
for tag in ${TAGS:-default shold default shold}; do
  lib=$PWD/nbody_b200/libnbody_b200_$tag.so; [ "$tag" = default ] && lib=$PWD/nbody_b200/libnbody_b200.so
  echo -n "{\"variant\": \"$tag\", \"result\": " >> $out
  NB_LIB_PATH=$lib python tools/bench_paths.py sh 10000000 2>/dev/null | tail -1 | python -c "import sys,json; r=json.loads(sys.stdin.read()); print(json.dumps({k:round(v['ms'],4) for k,v in r['stages'].items()}|{k:r[k] for k in ('build_ms','force_ms','step_ms','pairs_in_cutoff')}))" >> $out
  sed -i '$ s/$/}/' $out
done
cat $out
