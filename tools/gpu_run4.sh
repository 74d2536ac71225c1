#!/bin/bash
mkdir -p gpurun_out

for cfg in "64 56 64 3 1" "128 28 128 3 1" "256 14 256 3 1" "512 7 512 3 1"; do
  for dbg in 0 7; do TKB_DBG=$dbg timeout 120 python tools/conv_probe.py $cfg 2>&1 | tail -1 | cut -c1-80; done
done
timeout 1500 python -m pytest tests/test_gpu_parity.py -m gpu -q --maxfail=20 -p no:cacheprovider > gpurun_out/T.log 2>&1; echo "tests rc=$?"
tail -5 gpurun_out/T.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.json; tail -3 gpurun_out/bench.err
