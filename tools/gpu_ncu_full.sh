#!/bin/bash
# one ncu --set full capture of the tcgen05 kernels of the 2nd ResNet-18 b64 step (after the same command ran plain)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python tools/profile_step.py tc 0 2 > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"tc_halo_kernel|tc_gemm_kernel" -s 21 -c 21 -f -o gpurun_out/prof_tc python tools/profile_step.py tc 0 2 > gpurun_out/ncu2.log 2>&1
tail -n 3 gpurun_out/plain2.log gpurun_out/ncu2.log
ls -la gpurun_out/prof_tc.ncu-rep
