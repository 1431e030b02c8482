python -m pytest tests/test_gpu_multi.py -x -q 2>&1 | tail -4 > gpurun_out/multi_n8.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29621 bench.py --gpus 8 --steps 30 --warmup 5 --skip-cpu > gpurun_out/bench_n8_r1g.json 2> gpurun_out/bench_n8_r1g.err; echo n8 rc=$?
python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29622 bench.py --gpus 4 --steps 30 --warmup 5 --skip-cpu > gpurun_out/bench_n4_r1g.json 2> gpurun_out/bench_n4_r1g.err; echo n4 rc=$?
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29623 bench_vocab_parallel.py > gpurun_out/vocab_parallel_n8_r1g.json 2> gpurun_out/vocab_parallel_n8_r1g.err; echo vp rc=$?
cat gpurun_out/multi_n8.log
