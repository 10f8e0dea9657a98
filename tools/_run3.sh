mkdir -p gpurun_out; rm -f gpurun_out/paths3.jsonl
(timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest3.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest3.log); tail -15 gpurun_out/pytest3.log
for w in 16 8; do NB_SORT_WARPS=$w python tools/bench_paths.py sort 16777216 30 >> gpurun_out/paths3.jsonl 2>>gpurun_out/paths3.err; done
NB_SORT_WARPS=16 python tools/bench_paths.py sort 10000000 21 >> gpurun_out/paths3.jsonl 2>>gpurun_out/paths3.err
for sc in 2048 1024 768; do NB_BH_STACK=$sc python tools/bench_paths.py bh 16777216 plummer >> gpurun_out/paths3.jsonl 2>>gpurun_out/paths3.err; done
NB_BH_STACK=1024 NB_BH_SUPER=256 python tools/bench_paths.py bh 16777216 plummer >> gpurun_out/paths3.jsonl 2>>gpurun_out/paths3.err
python tools/bench_paths.py bh 1048576 plummer >> gpurun_out/paths3.jsonl 2>>gpurun_out/paths3.err
python tools/bench_paths.py bh 16777216 sphere >> gpurun_out/paths3.jsonl 2>>gpurun_out/paths3.err
cat gpurun_out/paths3.jsonl; tail -5 gpurun_out/paths3.err
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-reference-gpu > gpurun_out/bench3.json 2> gpurun_out/bench3.err; cat gpurun_out/bench3.json; tail -3 gpurun_out/bench3.err
ncu --set full --clock-control none --import-source on -k regex:"rs_onesweep|rs_hist_all" --launch-skip 10 -c 5 -f -o gpurun_out/sort_16mi python tools/bench_paths.py sort 16777216 30 > gpurun_out/ncu_sort.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:bh_force --launch-skip 3 -c 1 -f -o gpurun_out/bh_force_1mi_v2 python tools/bench_paths.py bh 1048576 plummer > gpurun_out/ncu_bh5.log 2>&1
