# A/B of build-knob sets on the config5 pipeline (aper apr=5 -> recentring pkfit_noise) at reduced size, one GPU:
#   gpurun --timeout 1200 -- 'bash scripts/gpu_c5.sh "" "-DPK_MINB_FIT=4"'
set -u -o pipefail
mkdir -p gpurun_out
i=0
for FLAGS in "$@"; do
    make -s -B -C pythonphot_b200/csrc EXTRA="$FLAGS" > gpurun_out/c5_build_$i.log 2>&1 || { echo "BUILD FAILED $FLAGS"; tail -5 gpurun_out/c5_build_$i.log; exit 1; }
    timeout 600 python bench.py --workload config5 --c5-size ${C5_SIZE:-8192} --c5-stars ${C5_STARS:-2500000} --steps 2 --warmup 2 ${C5_ARGS:---no-parity} > gpurun_out/c5_bench_$i.log 2>&1 || echo "bench failed"
    python - "$FLAGS" $i <<'PY'
import json, sys
for l in open("gpurun_out/c5_bench_%s.log" % sys.argv[2]):
    if l.startswith("{"):
        c = json.loads(l)["config5"]
        print("[%s] %-32s | %.2f M stars/s, step %.1f ms, aper %s fit %s gather %s, mean niter %.2f, parity %s" % (
            sys.argv[2], sys.argv[1] or "(default)", c["value"] / 1e6, c["ms_per_step"], c["per_rank"]["aper_ms"], c["per_rank"]["pkfit_ms"],
            c["per_rank"]["exposed_gather_ms"], c["mean_niter"], (c.get("parity") or {}).get("mismatches")))
PY
    i=$((i+1))
done
