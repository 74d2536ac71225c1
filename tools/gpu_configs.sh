#!/bin/bash
# bench line of every BASELINE.json config (C1..C5) at N=1 -> gpurun_out/configs.jsonl
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
: > gpurun_out/configs.jsonl
for cfg in "mlp 256" "lenet5 256" "resnet18 64" "mobilenetv2 128" "bert_encoder 8" "bert_base 8"; do
  set -- $cfg
  timeout 600 python bench.py --model $1 --batch $2 --steps 20 --warmup 3 >> gpurun_out/configs.jsonl 2>> gpurun_out/configs.err || echo "{\"model\": \"$1\", \"failed\": true}" >> gpurun_out/configs.jsonl
done
python - <<'PY'
import json
for l in open('gpurun_out/configs.jsonl'):
    d=json.loads(l)
    if d.get('failed'): print(d); continue
    print(d['config']['workload'][:60], '| compute-only %.0f | traced %.0f | e2e %.0f | cpu %.2f | frac %.3f' % (d['config']['compute_only'], d['value'], d['e2e']['value'], d.get('cpu_baseline',{}).get('value',0), d['roofline']['frac'] or 0))
PY
tail -n 3 gpurun_out/configs.err
