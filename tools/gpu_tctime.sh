#!/bin/bash
# roofline accounting: event pair per run of consecutive tcgen05 launches (default) vs a pair per launch
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_runs.json 2> gpurun_out/bench_runs.err; echo "rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_runs.json').read().strip().splitlines()[-1])
print("per run   :", {k:d['roofline'][k] for k in ('achieved','peak','frac','avg_launch_us','launches_per_step')}, "value", d['value'], "compute_only", d['config'].get('compute_only_img_s'))
PY
TKB_TC_TIME_PER_LAUNCH=1 timeout 900 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_perlaunch.json 2> gpurun_out/bench_perlaunch.err; echo "rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_perlaunch.json').read().strip().splitlines()[-1])
print("per launch:", {k:d['roofline'][k] for k in ('achieved','peak','frac','avg_launch_us','launches_per_step')})
PY
timeout 300 python tools/conv_probe.py 64 56 64 3 1 | tail -1 | cut -c1-100
timeout 900 python -m pytest tests/test_gpu_float.py tests/test_gpu_parity.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -3
