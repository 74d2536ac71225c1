#!/bin/bash
# full GPU parity + bench + ncu launch list of the result
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -5
python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cut -c1-700 gpurun_out/bench.json
python tools/profile_step.py tc 0 3 > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv python tools/profile_step.py tc 0 3 > gpurun_out/ncu1.log 2>&1
python tools/parse_launches.py gpurun_out/launches.csv 2>&1 | tail -31 | cut -c1-120
