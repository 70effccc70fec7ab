# Takes a build knob through the same gate as a default change, on the GPU box:
#   gpurun --timeout 600 -- 'bash scripts/gpu_knob_check.sh "-DAP_ASM_SELECT=1"'
# 1. rebuilds libbphot.so with EXTRA="$1"   2. pytest -m gpu (parity)   3. bench.py (device-level line, no e2e / CPU legs)
# 4. rebuilds the default library and runs the same bench line for the A/B.  Logs under gpurun_out/knob_*.log.
set -u
FLAGS="${1:?usage: gpu_knob_check.sh \"-DKNOB=VALUE ...\"}"
make -s -B -C pythonphot_b200/csrc EXTRA="$FLAGS" 2>&1 | grep -i -E "error|warning: v" ; echo "built with $FLAGS"
timeout 300 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/knob_tests.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/knob_tests.log
timeout 300 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/knob_bench_on.log 2>&1
make -s -B -C pythonphot_b200/csrc 2>&1 | grep -i error
timeout 300 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/knob_bench_off.log 2>&1
python - <<'PY'
import json
for tag in ("on", "off"):
    for l in open("gpurun_out/knob_bench_%s.log" % tag):
        if l.startswith("{"):
            d = json.loads(l)
            print(tag, "step ms %.2f" % d["ms_per_step"], {k: round(v["ms"], 2) for k, v in d["stages"].items() if isinstance(v, dict) and "ms" in v},
                  {k: round(v["avg_launch_ms"], 3) for k, v in d.get("kernels", {}).items()})
PY
