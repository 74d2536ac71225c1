#!/bin/bash
# ncu --set full of the chained launch (one instance, after warm-up steps), report kept in gpurun_out/
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc_chain -s 2 -c 1 -f -o gpurun_out/chain_full3 python tools/profile_step.py tc 0 3 > gpurun_out/ncu_chain.log 2>&1
TKB_DBG=48 timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc_chain -s 2 -c 1 -f -o gpurun_out/chain_full3_dbg48 python tools/profile_step.py tc 0 3 > gpurun_out/ncu_chain48.log 2>&1
tail -2 gpurun_out/ncu_chain.log gpurun_out/ncu_chain48.log; ls -la gpurun_out/*.ncu-rep
