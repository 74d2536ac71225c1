#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python tests/debug_chain_first_mismatch.py 2>&1 | tail -7
B=4 IMG=224 timeout 300 python tests/debug_chain_first_mismatch.py 2>&1 | tail -4
bash tools/gpu_chain_quick.sh
