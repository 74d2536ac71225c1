#!/bin/bash
# epilogue-group variants (TKB_EG=1: two small CTAs per SM; 2: never; default = heuristic): parity, then layer and step times
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TKB_EG=1 timeout 1500 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -p no:cacheprovider > gpurun_out/T_eg4.log 2>&1; echo "parity EG=1 rc=$?"
tail -4 gpurun_out/T_eg4.log
for eg in 2 1; do
  echo "== TKB_EG=$eg"
  echo -n "stem "; TKB_EG=$eg timeout 120 python tools/stem_probe.py 2>&1 | tail -1
  echo -n "layer1 64->64 56 3x3 "; TKB_EG=$eg timeout 120 python tools/conv_probe.py 64 56 64 3 1 2>&1 | tail -1
  echo -n "mbv2 16->96 112 1x1 b128 "; TKB_EG=$eg timeout 120 python tools/conv_probe.py 16 112 96 1 1 128 2>&1 | tail -1
  echo -n "mbv2 32->16 112 1x1 b128 "; TKB_EG=$eg timeout 120 python tools/conv_probe.py 32 112 16 1 1 128 2>&1 | tail -1
  echo -n "mbv2 96->24 56 1x1 b128 "; TKB_EG=$eg timeout 120 python tools/conv_probe.py 96 56 24 1 1 128 2>&1 | tail -1
  echo -n "mbv2 24->144 56 1x1 b128 "; TKB_EG=$eg timeout 120 python tools/conv_probe.py 24 56 144 1 1 128 2>&1 | tail -1
  echo -n "mbv2 32->192 28 1x1 b128 "; TKB_EG=$eg timeout 120 python tools/conv_probe.py 32 28 192 1 1 128 2>&1 | tail -1
done
timeout 900 python tools/step_probe.py TKB_EG=2 TKB_EG=1
TKB_MODEL=mobilenetv2 TKB_BATCH=128 timeout 900 python tools/step_probe.py TKB_EG=2 TKB_EG=1
