#!/bin/bash
# launch list of one ResNet-18 b64 step + ncu --set full of the tcgen05 kernels (3rd step only)
mkdir -p gpurun_out
python tools/profile_step.py tc 0 3 > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv python tools/profile_step.py tc 0 3 > gpurun_out/ncu1.log 2>&1
python tools/profile_step.py tc 0 2 > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:tc_gemm -s 21 -c 21 -o gpurun_out/prof_tc python tools/profile_step.py tc 0 2 > gpurun_out/ncu2.log 2>&1
tail -3 gpurun_out/plain.log gpurun_out/ncu1.log gpurun_out/ncu2.log
