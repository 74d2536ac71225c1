#!/bin/bash
# halo-reuse kernel bring-up: parity with and without the descriptor base offset
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for dbg in 0 16; do
  echo "=== TKB_DBG=$dbg"
  TKB_DBG=$dbg timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "halo or implicit" 2>&1 | tail -15
done
