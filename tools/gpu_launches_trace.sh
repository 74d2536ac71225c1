#!/bin/bash
# ncu launch list of ResNet-18 b64 steps with the device-side trace capture on
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python tools/profile_step.py tc 1 3 > gpurun_out/plain_t.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_trace.csv python tools/profile_step.py tc 1 3 > gpurun_out/ncu_t.log 2>&1
tail -n 2 gpurun_out/plain_t.log gpurun_out/ncu_t.log
python tools/parse_launches.py gpurun_out/launches_trace.csv 2>&1 | grep -v "tc_halo\|tc_gemm" | tail -60 | cut -c1-120
