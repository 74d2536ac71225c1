#!/bin/bash
# parity, device timeline of the layer shapes, bench
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -3
for cfg in "256 14 256 3 1" "512 7 512 3 1"; do
  for dbg in 0 256 6 262; do
  echo "== $cfg dbg=$dbg"; TKB_DBG=$dbg TKB_TIMELINE=1 timeout 120 python tools/conv_probe.py $cfg 2>&1 | grep -E "timeline" | tail -1 | cut -c1-200
  TKB_DBG=$dbg timeout 120 python tools/conv_probe.py $cfg 2>&1 | tail -1 | cut -c1-80
  done
done | tee gpurun_out/timeline.txt
python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cut -c1-330 gpurun_out/bench.json; grep -o '"compute_only": [0-9.]*' gpurun_out/bench.json
