#!/bin/bash
cd "$(dirname "$0")/.."
for cfg in "64 56 64 3 1" "128 28 128 3 1" "256 14 256 3 1" "512 7 512 3 1" "64 56 128 3 2"; do
  for dbg in 0 8; do TKB_DBG=$dbg timeout 120 python tools/conv_probe.py $cfg 2>&1 | tail -1 | cut -c1-90; done
done
python bench.py --steps 20 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('PDL on : compute_only', d['config']['compute_only'], 'value', d['value'])"
TKB_NO_PDL=1 python bench.py --steps 20 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('PDL off: compute_only', d['config']['compute_only'], 'value', d['value'])"
