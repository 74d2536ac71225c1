#!/bin/bash
# parity + bench sanity, then experiment-bit decomposition of the deep-layer GEMM kernels
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -3
python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/bench.json
for cfg in "256 14 256 3 1" "512 7 512 3 1" "128 28 128 3 1"; do
  for dbg in 0 1 2 4 6 32 33 34 36 38 39 8; do TKB_DBG=$dbg timeout 120 python tools/conv_probe.py $cfg 2>&1 | tail -1 | cut -c1-90; done
done | tee gpurun_out/probe4.txt
