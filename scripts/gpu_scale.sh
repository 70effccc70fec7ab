# one bench line at N ranks (no tests)
N=${1:-8}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/scale_$N.log 2>&1
tail -n 2 gpurun_out/scale_$N.log | cut -c1-2500
