# quick GPU iteration: parity tests + one bench line (no ncu)
python -m pytest tests -m gpu -q --tb=short 2>&1 | tail -80 > gpurun_out/quick_tests.log
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/quick_bench.log 2>&1
tail -n 30 gpurun_out/quick_tests.log; cat gpurun_out/quick_bench.log | cut -c1-1500
