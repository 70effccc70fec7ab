# profile pass: plain bench (must exit 0) -> ncu launch list of the same command -> ncu --set full of each kernel
TAG=${1:-r1b}
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/${TAG}_ncu_launch.log 2>&1
python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --workload aper > gpurun_out/${TAG}_plain_aper.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:aper_kernel -s 3 -c 1 -o gpurun_out/${TAG}_aper_full -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --workload aper > gpurun_out/${TAG}_ncu_aper.log 2>&1
python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --workload pkfit > gpurun_out/${TAG}_plain_pkfit.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:pkfit_kernel -s 3 -c 1 -o gpurun_out/${TAG}_pkfit_full -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --workload pkfit > gpurun_out/${TAG}_ncu_pkfit.log 2>&1
ls -la gpurun_out | tail -n 12
