# multi-GPU check: bit-identical sharded run + bench at N ranks (N = first argument, default 2)
N=${1:-2}
python -m pytest tests/test_shard_gloo.py -m gpu -q --tb=short 2>&1 | tail -20 > gpurun_out/multi_tests.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/multi_bench_$N.log 2>&1
tail -n 8 gpurun_out/multi_tests.log; tail -n 3 gpurun_out/multi_bench_$N.log | cut -c1-2500
