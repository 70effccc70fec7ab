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
    if [ "${C5:-0}" = "1" ]; then  # the config5 pipeline at a quarter of its size (aper apr=5 -> recentring pkfit_noise), with its parity block
        timeout 600 python bench.py --workload config5 --c5-size ${C5_SIZE:-8192} --c5-stars ${C5_STARS:-2500000} --steps 2 --warmup 2 > gpurun_out/abp_c5_$TAG.log 2>&1 || { echo "[$TAG] config5 bench failed"; tail -5 gpurun_out/abp_c5_$TAG.log; }
    fi
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
    try:
        for l in open("gpurun_out/abp_c5_%s.log" % tag):
            if l.startswith("{"):
                c = json.loads(l)["config5"]
                print("[%s] config5 %.2f M stars/s, step %.1f ms, aper %s fit %s, mean niter %.2f, parity %s" % (
                    tag, c["value"] / 1e6, c["ms_per_step"], c["per_rank"]["aper_ms"], c["per_rank"]["pkfit_ms"], c["mean_niter"],
                    (c.get("parity") or {}).get("mismatches")))
    except FileNotFoundError:
        pass
    except Exception as e:
        print(tag, "no config5 line:", e)
PY
