# bench line + e2e breakdown (no tests)
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/q2_bench.log 2>&1
timeout 600 python scripts/e2e_breakdown.py > gpurun_out/q2_e2e.log 2>&1
tail -c 4000 gpurun_out/q2_bench.log; cat gpurun_out/q2_e2e.log
