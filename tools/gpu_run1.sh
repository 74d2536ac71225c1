#!/bin/bash
# First GPU bring-up: CUDA-core path first, then the tcgen05 probe, then the tcgen05 GEMM, then everything.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
echo "== A: CUDA-core path" | tee gpurun_out/run1.log
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x --maxfail=8 -k "simt or auto or unsupported" -p no:cacheprovider > gpurun_out/A_simt.log 2>&1; echo "A rc=$?" | tee -a gpurun_out/run1.log
tail -5 gpurun_out/A_simt.log | tee -a gpurun_out/run1.log
echo "== B: int8 peak probe" | tee -a gpurun_out/run1.log
timeout 120 python -c "
from tachikoma_b200 import runtime as RT
for it in (500, 2000, 8000):
    print('iters', it, 'TOPS', RT.probe_int8_peak(it))
" > gpurun_out/B_probe.log 2>&1; echo "B rc=$?" | tee -a gpurun_out/run1.log
cat gpurun_out/B_probe.log | tail -5 | tee -a gpurun_out/run1.log
echo "== C: tcgen05 GEMM" | tee -a gpurun_out/run1.log
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q --maxfail=4 -k "tcgen05" -p no:cacheprovider > gpurun_out/C_tc.log 2>&1; echo "C rc=$?" | tee -a gpurun_out/run1.log
tail -30 gpurun_out/C_tc.log | tee -a gpurun_out/run1.log
echo "== D: tc everything" | tee -a gpurun_out/run1.log
timeout 1200 python -m pytest tests/test_gpu_parity.py -m gpu -q --maxfail=6 -k "tc and not tcgen05" -p no:cacheprovider > gpurun_out/D_all.log 2>&1; echo "D rc=$?" | tee -a gpurun_out/run1.log
tail -30 gpurun_out/D_all.log | tee -a gpurun_out/run1.log
echo "== E: smoke" | tee -a gpurun_out/run1.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/E_smoke.log 2>&1; echo "E rc=$?" | tee -a gpurun_out/run1.log
tail -5 gpurun_out/E_smoke.log | tee -a gpurun_out/run1.log
