#!/bin/bash
cd "$(dirname "$0")/.."
for cfg in "256 14 256 3 1" "512 7 512 3 1" "128 28 256 3 2" "256 14 512 3 2" "64 56 128 3 2"; do
  for bn in 64 128 256; do TKB_BN=$bn timeout 120 python tools/conv_probe.py $cfg 2>&1 | tail -1 | cut -c1-70 | sed "s/^/bn=$bn /"; done
done
