# Takes build knobs through the same gate as a default change, on the GPU box:
#   gpurun --timeout 900 -- 'bash scripts/gpu_knob_check.sh "-DAP_ASM_SELECT=1" ["-DOTHER=1" ...]'
# For every knob set: rebuild libbphot.so with EXTRA="<set>", pytest -m gpu (parity), bench.py (device-level line, no e2e / CPU
# legs).  Then the default library is rebuilt and the same bench line is run for the A/B.  Logs under gpurun_out/knob_*.log.
# A failed build aborts the script: nothing is measured against a stale library.
set -u -o pipefail
[ $# -ge 1 ] || { echo "usage: gpu_knob_check.sh \"-DKNOB=VALUE ...\" [more knob sets]"; exit 2; }
mkdir -p gpurun_out
build() {  # $1 = EXTRA flags ("" = default)
    if ! make -s -B -C pythonphot_b200/csrc EXTRA="$1" > gpurun_out/knob_build.log 2>&1; then
        echo "BUILD FAILED with EXTRA='$1'"; tail -20 gpurun_out/knob_build.log; exit 1
    fi
    python -c "from pythonphot_b200 import _lib; print('built EXTRA=[$1] version', _lib.lib().bphot_version(), 'flags', _lib.build_flags())"
}
i=0
for FLAGS in "$@"; do
    i=$((i+1))
    build "$FLAGS"
    timeout 420 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/knob_tests_$i.log 2>&1; echo "[$FLAGS] pytest rc=$?"; tail -3 gpurun_out/knob_tests_$i.log
    timeout 300 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/knob_bench_$i.log 2>&1 || echo "[$FLAGS] bench failed"
done
build ""
timeout 300 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/knob_bench_0.log 2>&1 || echo "default bench failed"
python - "$@" <<'PY'
import json, sys
names = ["default"] + sys.argv[1:]
for i, tag in enumerate(names):
    try:
        for l in open("gpurun_out/knob_bench_%d.log" % i):
            if l.startswith("{"):
                d = json.loads(l)
                print(tag, "| step ms %.2f" % d["ms_per_step"], {k: round(v["ms"], 2) for k, v in d["stages"].items() if isinstance(v, dict) and "ms" in v},
                      {k: round(v["avg_launch_ms"] * v["launches_per_step"], 2) for k, v in d.get("kernels", {}).items()})
    except Exception as e:
        print(tag, "no bench line:", e)
PY
