#!/bin/bash
cd "$(dirname "$0")/.."
for cfg in "64 56 64 3 1"; do
  for dbg in 0 1 4 5 32 33 36 37 8; do TKB_DBG=$dbg timeout 120 python tools/conv_probe.py $cfg 2>&1 | tail -1 | cut -c1-90; done
done
