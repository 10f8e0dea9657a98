mkdir -p gpurun_out; rm -f gpurun_out/paths2.jsonl
(timeout 900 python -m pytest tests/test_gpu_sort_tree.py tests/test_gpu_grid.py tests/test_gpu_full_size.py tests/test_gpu_vs_reference.py -m gpu -x -q > gpurun_out/pytest2.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest2.log); tail -8 gpurun_out/pytest2.log
for w in 16 8; do NB_SORT_WARPS=$w python tools/bench_paths.py sort 16777216 30 >> gpurun_out/paths2.jsonl 2>>gpurun_out/paths2.err; done
NB_SORT_WARPS=16 python tools/bench_paths.py sort 10000000 21 >> gpurun_out/paths2.jsonl 2>>gpurun_out/paths2.err
for sg in 64 128 256 512 1024; do NB_BH_SUPER=$sg python tools/bench_paths.py bh 16777216 plummer >> gpurun_out/paths2.jsonl 2>>gpurun_out/paths2.err; done
NB_BH_GROUP=32 NB_BH_SUPER=256 python tools/bench_paths.py bh 16777216 plummer >> gpurun_out/paths2.jsonl 2>>gpurun_out/paths2.err
python tools/bench_paths.py bh 1048576 plummer >> gpurun_out/paths2.jsonl 2>>gpurun_out/paths2.err
python tools/bench_paths.py bh 16777216 sphere >> gpurun_out/paths2.jsonl 2>>gpurun_out/paths2.err
python tools/bench_paths.py sh 10000000 >> gpurun_out/paths2.jsonl 2>>gpurun_out/paths2.err
cat gpurun_out/paths2.jsonl; tail -5 gpurun_out/paths2.err
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv --log-file gpurun_out/bh16_launches2.csv python tools/bench_paths.py bh 16777216 plummer > gpurun_out/ncu_bh2.log 2>&1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv --log-file gpurun_out/sh10_launches2.csv python tools/bench_paths.py sh 10000000 > gpurun_out/ncu_sh2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:bh_force --launch-skip 3 -c 1 -f -o gpurun_out/bh_force_1mi python tools/bench_paths.py bh 1048576 plummer > gpurun_out/ncu_bh3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"rs_onesweep|rs_hist_all|bh_com|bh_karras|bh_groups|bh_gather" --launch-skip 27 -c 9 -f -o gpurun_out/bh_build_16mi python tools/bench_paths.py bh 16777216 plummer > gpurun_out/ncu_bh4.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"sh_force|sh_ranges" --launch-skip 6 -c 2 -f -o gpurun_out/sh_10m python tools/bench_paths.py sh 10000000 > gpurun_out/ncu_sh3.log 2>&1
ls -la gpurun_out
