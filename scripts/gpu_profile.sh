# profile pass: plain bench (must exit 0) -> ncu launch list of the same command -> ncu --set full of each kernel
TAG=${1:-r02}
KERNELS=${2:-"aper_stage1 aper_mmm aper_sums pkfit_kernel"}
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-parity > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-parity > gpurun_out/${TAG}_ncu_launch.log 2>&1
for KN in $KERNELS; do
  WL=aper; SKIP=50; case $KN in aper_mmm*) SKIP=2;; esac; case $KN in pkfit*) WL=pkfit; SKIP=2;; aper_sums*) SKIP=2;; esac
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-parity --workload $WL > gpurun_out/${TAG}_plain_${KN}.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:$KN -s $SKIP -c 1 -o gpurun_out/${TAG}_${KN} -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-parity --workload $WL > gpurun_out/${TAG}_ncu_${KN}.log 2>&1
  # gpurun brings back at most 64 MiB and one report is ~25 MB: summarise on the box, keep the report only with KEEP_REP=1
  python scripts/profile_summary.py gpurun_out/${TAG}_${KN}.ncu-rep > gpurun_out/${TAG}_${KN}.txt 2> gpurun_out/${TAG}_${KN}.err
  [ "${KEEP_REP:-0}" = "1" ] || rm -f gpurun_out/${TAG}_${KN}.ncu-rep
done
ls -la gpurun_out | tail -n 12
