#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -5
python bench.py --model mobilenetv2 --batch 128 --steps 10 --warmup 3 --no-cpu-baseline 2>gpurun_out/mb.err | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('mobilenetv2 b128: compute_only %.0f value %.0f e2e %.0f' % (d['config']['compute_only'], d['value'], d['e2e']['value']))"
tail -n 2 gpurun_out/mb.err
TKB_MODEL=mobilenetv2 TKB_BATCH=128 bash tools/gpu_launches_model.sh
