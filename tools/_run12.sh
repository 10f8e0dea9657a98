mkdir -p gpurun_out; rm -f gpurun_out/paths12.jsonl
(timeout 900 python -m pytest tests/test_gpu_vs_reference.py tests/test_gpu_sort_tree.py -m gpu -x -q -k "sharded or peer or bh" > gpurun_out/pytest12.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest12.log); tail -12 gpurun_out/pytest12.log
python tools/bench_paths.py bh 16777216 plummer >> gpurun_out/paths12.jsonl 2>>gpurun_out/paths12.err
python tools/bench_paths.py bh 1048576 plummer >> gpurun_out/paths12.jsonl 2>>gpurun_out/paths12.err
cat gpurun_out/paths12.jsonl; tail -5 gpurun_out/paths12.err
