#!/bin/bash
# quick A/B of a kernel change: parity of the tcgen05 tests, then layer and step times
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -p no:cacheprovider > gpurun_out/T_ab.log 2>&1; echo "parity rc=$?"
tail -2 gpurun_out/T_ab.log
echo -n "stem "; timeout 120 python tools/stem_probe.py 2>&1 | tail -1 | cut -c1-90
echo -n "stem no-MMA "; TKB_DBG=32 timeout 120 python tools/stem_probe.py 2>&1 | tail -1 | cut -c1-90
echo -n "layer1 64->64 56 3x3 "; timeout 120 python tools/conv_probe.py 64 56 64 3 1 2>&1 | tail -1 | cut -c1-90
echo -n "layer2 128->128 28 3x3 "; timeout 120 python tools/conv_probe.py 128 28 128 3 1 2>&1 | tail -1 | cut -c1-90
echo -n "layer3 256->256 14 3x3 "; timeout 120 python tools/conv_probe.py 256 14 256 3 1 2>&1 | tail -1 | cut -c1-90
echo -n "mbv2 16->96 112 1x1 b128 "; timeout 120 python tools/conv_probe.py 16 112 96 1 1 128 2>&1 | tail -1 | cut -c1-90
echo -n "mbv2 32->16 112 1x1 b128 "; timeout 120 python tools/conv_probe.py 32 112 16 1 1 128 2>&1 | tail -1 | cut -c1-90
timeout 900 python tools/step_probe.py
TKB_MODEL=mobilenetv2 TKB_BATCH=128 timeout 900 python tools/step_probe.py
