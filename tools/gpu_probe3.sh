#!/bin/bash
for cfg in "64 56 64 3 1" "256 14 256 3 1"; do
  for dbg in 0 7 8; do TKB_DBG=$dbg timeout 120 python tools/conv_probe.py $cfg 2>&1 | tail -1 | cut -c1-80; done
done
