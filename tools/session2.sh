#!/bin/bash
mkdir -p gpurun_out
echo "== ubench"; timeout 120 tools/ubench/tmem_bw 2>&1 | tee gpurun_out/ubench_tmem.txt
echo "== mma-only diag"; timeout 300 python tools/perf_sweep.py --what gemm --variants 3,4,7 --shapes 8192x28672x8192,16384x18432x7168 2>&1 | grep '"op": "gemm"' | tee gpurun_out/diag_mma_only.jsonl
echo "== pytest gpu"; timeout 1500 python -m pytest tests -q -m gpu --timeout 900 2>&1 | tail -25
echo "== ncu quant"
python tools/ncu_target.py quant 32768 > gpurun_out/plain_q.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:ptgq_kernel -s 1 -c 1 -o gpurun_out/quant_r1a python tools/ncu_target.py quant 32768 > gpurun_out/ncu_q.log 2>&1
tail -3 gpurun_out/ncu_q.log
echo "== ncu gemm"
python tools/ncu_target.py gemm 8192 8192 8192 4 > gpurun_out/plain_g.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:fp8_blockwise_gemm -s 1 -c 1 -o gpurun_out/gemm_r1a python tools/ncu_target.py gemm 8192 8192 8192 4 > gpurun_out/ncu_g.log 2>&1
tail -3 gpurun_out/ncu_g.log
ls -la gpurun_out
