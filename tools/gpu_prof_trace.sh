#!/bin/bash
mkdir -p gpurun_out
python tools/profile_step.py tc 1 3 > gpurun_out/plain_t.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_trace.csv python tools/profile_step.py tc 1 3 > gpurun_out/ncu_t.log 2>&1
echo rc=$?
