#!/bin/bash
# ncu launch list (gpu__time_duration) of ResNet-18 b64 steps; only after the same command exited 0 plain
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python tools/profile_step.py tc 0 3 > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv python tools/profile_step.py tc 0 3 > gpurun_out/ncu1.log 2>&1
tail -3 gpurun_out/plain.log gpurun_out/ncu1.log
python tools/parse_launches.py gpurun_out/launches.csv 2>&1 | tail -45
