#!/bin/bash
# full validation: build check, GPU suite (-x like the driver), smoke, bench N=1, reference arm, all configs
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q -p no:cacheprovider -W ignore > gpurun_out/T.log 2>&1; echo "tests rc=$?"; tail -4 gpurun_out/T.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cut -c1-300 gpurun_out/bench.json; tail -3 gpurun_out/bench.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cut -c1-300 gpurun_out/bench_ref.json
bash tools/gpu_configs.sh 2>&1 | tail -8
