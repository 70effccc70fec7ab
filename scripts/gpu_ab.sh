# A/B of build-knob sets on the GPU box: for every argument (a quoted EXTRA string, "" = the defaults) rebuild libbphot.so,
# run the GPU parity suite (unless NOTESTS=1) and one device-level bench line.  Aborts on a failed build.
#   gpurun --timeout 1200 -- 'bash scripts/gpu_ab.sh "" "-DAP_MMM_INDEXED=0"'
set -u -o pipefail
mkdir -p gpurun_out
i=0
for FLAGS in "$@"; do
    if ! make -s -B -C pythonphot_b200/csrc EXTRA="$FLAGS" > gpurun_out/ab_build_$i.log 2>&1; then
        echo "BUILD FAILED with EXTRA='$FLAGS'"; tail -20 gpurun_out/ab_build_$i.log; exit 1
    fi
    python -c "from pythonphot_b200 import _lib; print('[$i] EXTRA=[$FLAGS]', _lib.build_flags())"
    if [ "${NOTESTS:-0}" != "1" ]; then
        timeout 600 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/ab_tests_$i.log 2>&1; echo "[$i] pytest rc=$?"; tail -4 gpurun_out/ab_tests_$i.log
    fi
    timeout 300 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline ${BENCH_ARGS:-} > gpurun_out/ab_bench_$i.log 2>&1 || { echo "[$i] bench failed"; tail -5 gpurun_out/ab_bench_$i.log; }
    i=$((i+1))
done
python - "$@" <<'PY'
import json, sys
for i, tag in enumerate(sys.argv[1:]):
    try:
        for l in open("gpurun_out/ab_bench_%d.log" % i):
            if l.startswith("{"):
                d = json.loads(l)
                print("[%d] %-40s | step ms %.2f" % (i, tag or "(default)", d["ms_per_step"]), {k: round(v["ms"], 2) for k, v in d["stages"].items() if isinstance(v, dict) and "ms" in v},
                      {k: round(v["avg_launch_ms"] * v["launches_per_step"], 2) for k, v in d.get("kernels", {}).items()})
    except Exception as e:
        print(i, tag, "no bench line:", e)
PY
