mkdir -p gpurun_out; rm -f gpurun_out/paths11.jsonl
(timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest11.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest11.log); tail -12 gpurun_out/pytest11.log
for a in "graph 4096" "graph 16384" "graph 65536" "graph 4096 bh" "graph 65536 bh" "graph 1048576 bh"; do python tools/bench_paths.py $a >> gpurun_out/paths11.jsonl 2>>gpurun_out/paths11.err; done
cat gpurun_out/paths11.jsonl; tail -5 gpurun_out/paths11.err
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-reference-gpu > gpurun_out/bench11.json 2> gpurun_out/bench11.err; cut -c1-330 gpurun_out/bench11.json; tail -3 gpurun_out/bench11.err
