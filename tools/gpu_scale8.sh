#!/bin/bash
# N-GPU bench on all GPUs of the box; checks that stdout is exactly one JSON line
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
echo "gpus: $N"
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "rc=$?"
echo "stdout lines: $(wc -l < gpurun_out/bench_n$N.json)"
python -c "
import json; d=json.loads(open('gpurun_out/bench_n$N.json').read()); print('N=%d value %.0f compute_only %.0f e2e %.0f ms/step %.3f' % (d['n_gpus'], d['value'], d['config']['compute_only'], d['e2e']['value'], d['ms_per_step']))"
