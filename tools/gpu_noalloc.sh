#!/bin/bash
# the allocation-free entry point alone in a fresh process, then the whole GPU suite (which ends with it), then the
# bench and the launch lists for profiles/
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x -p no:cacheprovider > gpurun_out/Tnoalloc.log 2>&1; echo "noalloc alone rc=$?"
tail -25 gpurun_out/Tnoalloc.log
timeout 2400 python -m pytest tests -m gpu -q -x -p no:cacheprovider > gpurun_out/T.log 2>&1; echo "tests rc=$?"
tail -8 gpurun_out/T.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-verify --sustained-s 0 > gpurun_out/ncu_bench.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/launches_mbv2.csv \
  python bench.py --model mobilenetv2 --batch 128 --steps 2 --warmup 3 --no-cpu-baseline --no-verify --sustained-s 0 > gpurun_out/ncu_mbv2.log 2>&1; echo "ncu mbv2 rc=$?"
