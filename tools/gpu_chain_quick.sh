#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -p no:cacheprovider -k "models_bit_exact or two_threads or deferred or b64_224 or qnn_conv2d_profiles or bert_base" > gpurun_out/Tchain.log 2>&1; echo "chain tests rc=$?"
tail -6 gpurun_out/Tchain.log
python tools/chain_probe.py
