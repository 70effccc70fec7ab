python -m pytest tests -m gpu -q 2>&1 | tail -40 > gpurun_out/r1_gputests.log
python __graft_entry__.py smoke > gpurun_out/r1_smoke.log 2>&1
timeout 900 python bench.py > gpurun_out/r1_bench_full.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r1_ncu_launch.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:aper_kernel -s 3 -c 1 -o gpurun_out/r1_aper_full -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --workload aper > gpurun_out/r1_ncu_aper.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:pkfit_kernel -s 3 -c 1 -o gpurun_out/r1_pkfit_full -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --workload pkfit > gpurun_out/r1_ncu_pkfit.log 2>&1
tail -n 5 gpurun_out/r1_gputests.log gpurun_out/r1_smoke.log gpurun_out/r1_bench_full.log
