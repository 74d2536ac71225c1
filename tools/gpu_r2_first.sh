#!/bin/bash
# round 2, first GPU pass: the new parity tests first (fail fast), then the whole GPU suite, then bench
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -p no:cacheprovider -k "avg_pool or nn_pad or packed_run or two_threads or deferred or debug_executor or b64_224 or mobilenetv2_b or c1_c2" > gpurun_out/Tnew.log 2>&1; echo "new tests rc=$?"
tail -15 gpurun_out/Tnew.log
timeout 1500 python -m pytest tests -m gpu -q --maxfail=10 -p no:cacheprovider > gpurun_out/T.log 2>&1; echo "tests rc=$?"
tail -8 gpurun_out/T.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1; lscpu | head -25 > gpurun_out/lscpu.txt
