#!/bin/bash
mkdir -p gpurun_out
L=gpurun_out/run3.log
timeout 1500 python -m pytest tests/test_gpu_parity.py -m gpu -q --maxfail=20 -p no:cacheprovider > gpurun_out/T.log 2>&1; echo "tests rc=$?" | tee $L
tail -25 gpurun_out/T.log | tee -a $L
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" | tee -a $L
cat gpurun_out/bench.json | tee -a $L
python tools/profile_step.py tc 0 3 > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv python tools/profile_step.py tc 0 3 > gpurun_out/ncu1.log 2>&1
echo "ncu rc=$?" | tee -a $L
python tools/profile_step.py tc 0 2 > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:tc_gemm -s 21 -c 7 -o gpurun_out/prof_tc python tools/profile_step.py tc 0 2 > gpurun_out/ncu2.log 2>&1
echo "ncu2 rc=$?" | tee -a $L
