#!/bin/bash
# where a tile's time goes in the GEMM kernel for MobileNetV2's pointwise layers (TKB_DBG experiment bits)
cd "$(dirname "$0")/.."
for shape in "32 112 16 1 1 128" "16 112 96 1 1 128" "96 56 24 1 1 128"; do
  for dbg in 0 1 2 4 6 7 8; do
    TKB_DBG=$dbg timeout 120 python tools/conv_probe.py $shape 2>&1 | tail -1 | cut -c1-110
  done
done
