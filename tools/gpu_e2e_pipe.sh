#!/bin/bash
# two batches in flight: the new test + trace / thread tests, then the bench line
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -p no:cacheprovider -k "in_flight or two_threads or deferred or trace or b64_224" > gpurun_out/Tpipe.log 2>&1; echo "tests rc=$?"
tail -6 gpurun_out/Tpipe.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_pipe.json 2> gpurun_out/bench_pipe.err; echo "bench rc=$?"
cat gpurun_out/bench_pipe.json; tail -5 gpurun_out/bench_pipe.err
