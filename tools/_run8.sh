mkdir -p gpurun_out; rm -f gpurun_out/paths8.jsonl
(timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest8.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest8.log); tail -6 gpurun_out/pytest8.log
python tools/bench_paths.py sh 10000000 >> gpurun_out/paths8.jsonl 2>>gpurun_out/paths8.err
python tools/bench_paths.py sort 16777216 30 >> gpurun_out/paths8.jsonl 2>>gpurun_out/paths8.err
python tools/bench_paths.py sort 10000000 21 >> gpurun_out/paths8.jsonl 2>>gpurun_out/paths8.err
cat gpurun_out/paths8.jsonl; tail -3 gpurun_out/paths8.err
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-reference-gpu > gpurun_out/bench8.json 2> gpurun_out/bench8.err; cut -c1-330 gpurun_out/bench8.json; tail -3 gpurun_out/bench8.err
ncu --set full --clock-control none --import-source on -k regex:"sh_force" --launch-skip 6 -c 1 -f -o gpurun_out/sh_force_v2 python tools/bench_paths.py sh 10000000 > gpurun_out/ncu_sh8.log 2>&1
