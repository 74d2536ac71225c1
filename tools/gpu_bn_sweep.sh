#!/bin/bash
# N-tile width (TKB_BN) for MobileNetV2's expand layers
cd "$(dirname "$0")/.."
for shape in "24 56 144 1 1 128" "32 28 192 1 1 128" "64 14 384 1 1 128" "96 14 576 1 1 128" "160 7 960 1 1 128"; do
  for bn in 0 64 128 256; do
    if [ $bn = 0 ]; then unset TKB_BN; else export TKB_BN=$bn; fi
    echo -n "BN=$bn "; timeout 120 python tools/conv_probe.py $shape 2>&1 | tail -1 | cut -c1-75
  done
done
