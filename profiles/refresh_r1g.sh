python -m pytest tests/test_gpu_stats.py -x -q 2>&1 | tail -3 > gpurun_out/stats_test.log
python bench.py > gpurun_out/bench_r1g.json 2> gpurun_out/bench_r1g.err; echo bench rc=$?
python bench_kernels.py --out gpurun_out/kernels_r1h.json > gpurun_out/kernels_r1h.log 2>&1; echo kernels rc=$?
python bench.py --steps 2 --warmup 3 --skip-e2e --skip-cpu > /dev/null 2>&1 && ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_r1g.csv python bench.py --steps 2 --warmup 3 --skip-e2e --skip-cpu > gpurun_out/ncu_list.log 2>&1; echo list rc=$?
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:fused_rowres -c 1 -o gpurun_out/rowres_r1g python bench.py --steps 2 --warmup 3 --skip-e2e --skip-cpu > gpurun_out/ncu_full.log 2>&1; echo full rc=$?
cat gpurun_out/stats_test.log
