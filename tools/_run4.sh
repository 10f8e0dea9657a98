mkdir -p gpurun_out; rm -f gpurun_out/paths4.jsonl
(timeout 600 python -m pytest tests/test_gpu_sort_tree.py tests/test_gpu_grid.py tests/test_gpu_vs_reference.py tests/test_gpu_full_size.py -m gpu -x -q > gpurun_out/pytest4.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest4.log); tail -5 gpurun_out/pytest4.log
for w in 16 8; do NB_SORT_WARPS=$w python tools/bench_paths.py sort 16777216 30 >> gpurun_out/paths4.jsonl 2>>gpurun_out/paths4.err; done
NB_SORT_WARPS=16 python tools/bench_paths.py sort 10000000 21 >> gpurun_out/paths4.jsonl 2>>gpurun_out/paths4.err
python tools/bench_paths.py bh 16777216 plummer >> gpurun_out/paths4.jsonl 2>>gpurun_out/paths4.err
python tools/bench_paths.py sh 10000000 >> gpurun_out/paths4.jsonl 2>>gpurun_out/paths4.err
cat gpurun_out/paths4.jsonl; tail -5 gpurun_out/paths4.err
ncu --set full --clock-control none --import-source on -k regex:"rs_onesweep" --launch-skip 8 -c 2 -f -o gpurun_out/sort_16mi_v2 python tools/bench_paths.py sort 16777216 30 > gpurun_out/ncu_sort2.log 2>&1
