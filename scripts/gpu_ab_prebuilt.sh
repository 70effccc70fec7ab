# A/B of PREBUILT library variants on the GPU box (no nvcc time on the box's clock):
#   here:  make -B -C pythonphot_b200/csrc EXTRA="-DKNOB=1" && cp pythonphot_b200/csrc/libbphot.so variants/libbphot_<tag>.so   (per variant)
#   gpurun --timeout 900 -- 'bash scripts/gpu_ab_prebuilt.sh tagA tagB ...'
# For every tag: variants/libbphot_<tag>.so is copied over the in-tree library, the GPU parity suite runs for the FIRST tag
# (all tags with TESTS=all, none with TESTS=none) and one device-level bench line (its own parity block included) for each.
set -u -o pipefail
mkdir -p gpurun_out
cp pythonphot_b200/csrc/libbphot.so gpurun_out/.libbphot_keep.so
i=0
for TAG in "$@"; do
    cp "variants/libbphot_$TAG.so" pythonphot_b200/csrc/libbphot.so || exit 1
    python -c "from pythonphot_b200 import _lib; print('[$TAG]', _lib.build_flags())"
    if [ "${TESTS:-first}" = "all" ] || { [ "${TESTS:-first}" = "first" ] && [ $i -eq 0 ]; }; then
        timeout 600 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/abp_tests_$TAG.log 2>&1; echo "[$TAG] pytest rc=$?"; tail -3 gpurun_out/abp_tests_$TAG.log
    fi
    timeout 300 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline ${BENCH_ARGS:-} > gpurun_out/abp_bench_$TAG.log 2>&1 || { echo "[$TAG] bench failed"; tail -5 gpurun_out/abp_bench_$TAG.log; }
    i=$((i+1))
done
cp gpurun_out/.libbphot_keep.so pythonphot_b200/csrc/libbphot.so
python - "$@" <<'PY'
import json, sys
for tag in sys.argv[1:]:
    try:
        for l in open("gpurun_out/abp_bench_%s.log" % tag):
            if l.startswith("{"):
                d = json.loads(l)
                print("[%s] step ms %.2f" % (tag, d["ms_per_step"]), {k: round(v["ms"], 2) for k, v in d["stages"].items() if isinstance(v, dict) and "ms" in v},
                      {k: round(v["avg_launch_ms"] * v["launches_per_step"], 2) for k, v in d.get("kernels", {}).items()},
                      "parity mismatches", d.get("parity", {}).get("mismatches"), "c5", (d.get("config5") or {}).get("ms_per_pass"))
    except Exception as e:
        print(tag, "no bench line:", e)
PY
