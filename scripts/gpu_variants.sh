# build variants on the box and time the aper stage of each: gpu_variants.sh "FLAGS1" "FLAGS2" ...
for V in "$@"; do
  make -s -B -C pythonphot_b200/csrc EXTRA="$V" 2>&1 | grep -i "error" 
  echo "== $V" >> gpurun_out/variants.log
  python bench.py --workload ${WL:-aper} --steps 3 --warmup 3 --no-e2e --no-cpu-baseline 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print({k: round(v['ms'], 3) for k, v in d['stages'].items() if isinstance(v, dict) and 'ms' in v})
" >> gpurun_out/variants.log
done
cat gpurun_out/variants.log
