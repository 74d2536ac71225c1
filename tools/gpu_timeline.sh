#!/bin/bash
# device timeline (clock64 stamps of CTA 0) of the tcgen05 kernels on the ResNet-18 layer shapes
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for cfg in "256 14 256 3 1" "512 7 512 3 1" "128 28 128 3 1" "64 56 64 3 1"; do
  for dbg in 0 6; do echo "== $cfg dbg=$dbg"; TKB_DBG=$dbg TKB_TIMELINE=1 timeout 120 python tools/conv_probe.py $cfg 2>&1 | grep -E "timeline|plan" | tail -2 | cut -c1-600; done
done | tee gpurun_out/timeline.txt
