#!/bin/bash
# final evidence for profiles/: bench N=1 plain, then the ncu launch list of the SAME bench command (ResNet-18 and MobileNetV2),
# then one ncu --set full capture of the tcgen05 kernels of a ResNet-18 step
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-verify --sustained-s 0 > gpurun_out/ncu_bench.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/launches_mbv2.csv \
  python bench.py --model mobilenetv2 --batch 128 --steps 2 --warmup 3 --no-cpu-baseline --no-verify --sustained-s 0 > gpurun_out/ncu_mbv2.log 2>&1; echo "ncu mbv2 rc=$?"
bash tools/gpu_ncu_full.sh 2>&1 | tail -3
