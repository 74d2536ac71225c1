#!/bin/bash
# v1 kernel bring-up: fast parity subsets first (each under its own timeout), then everything, then bench.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
L=gpurun_out/run2.log
echo "== A: tcgen05 GEMM (plain)" | tee $L
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q --maxfail=5 -k "tcgen05" -p no:cacheprovider > gpurun_out/A.log 2>&1; echo "A rc=$?" | tee -a $L
tail -15 gpurun_out/A.log | tee -a $L
echo "== B: implicit conv + fast epilogue" | tee -a $L
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q --maxfail=30 -k "implicit or fast_epilogue" -p no:cacheprovider > gpurun_out/B.log 2>&1; echo "B rc=$?" | tee -a $L
tail -40 gpurun_out/B.log | tee -a $L
echo "== C: everything else" | tee -a $L
timeout 1500 python -m pytest tests/test_gpu_parity.py -m gpu -q --maxfail=20 -k "not tcgen05 and not implicit and not fast_epilogue" -p no:cacheprovider > gpurun_out/C.log 2>&1; echo "C rc=$?" | tee -a $L
tail -30 gpurun_out/C.log | tee -a $L
echo "== D: smoke" | tee -a $L
timeout 300 python __graft_entry__.py smoke > gpurun_out/D.log 2>&1; echo "D rc=$?" | tee -a $L
tail -5 gpurun_out/D.log | tee -a $L
echo "== E: bench" | tee -a $L
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/E_bench.json 2> gpurun_out/E_bench.err; echo "E rc=$?" | tee -a $L
cat gpurun_out/E_bench.json | tee -a $L
tail -5 gpurun_out/E_bench.err | tee -a $L
