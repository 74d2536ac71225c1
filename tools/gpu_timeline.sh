#!/bin/bash
# device timeline (clock64 stamps of CTA 0) of the tcgen05 kernels; needs a library built with TKB_BUILD_TIMELINE=1
# usage: gpu_timeline.sh ["C H O K S" ...]      (TKB_TL_DBG="0 1" selects experiment bits)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
if [ $# -eq 0 ]; then set -- "64 56 64 3 1" "128 28 128 3 1" "256 14 256 3 1" "512 7 512 3 1"; fi
for cfg in "$@"; do
  for dbg in ${TKB_TL_DBG:-0}; do echo "== $cfg dbg=$dbg"; TKB_DBG=$dbg TKB_TIMELINE=1 timeout 120 python tools/conv_probe.py $cfg 2>&1 | grep -E "timeline" | tail -1 | cut -c1-800; done
done | tee gpurun_out/timeline.txt
