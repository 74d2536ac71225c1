#!/bin/bash
# chained launch: parity first (fail fast), then bench and the ncu launch list of one step
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -p no:cacheprovider -k "models_bit_exact or two_threads or deferred or b64_224 or qnn_conv2d_profiles or bert_base" > gpurun_out/Tchain.log 2>&1; echo "chain tests rc=$?"
tail -15 gpurun_out/Tchain.log
timeout 1200 python -m pytest tests -m gpu -q --maxfail=10 -p no:cacheprovider > gpurun_out/T.log 2>&1; echo "tests rc=$?"
tail -8 gpurun_out/T.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
timeout 300 python tools/profile_step.py tc 0 3 > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv python tools/profile_step.py tc 0 3 > gpurun_out/ncu1.log 2>&1
python tools/parse_launches.py gpurun_out/launches.csv 2>&1 | tail -24 | cut -c1-130
