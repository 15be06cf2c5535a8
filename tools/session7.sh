#!/bin/bash
mkdir -p gpurun_out
python tools/ncu_target.py gemm 64 24576 7168 0 > gpurun_out/plain_d.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:fp8_blockwise_gemm -s 1 -c 1 -o gpurun_out/gemm_decode_r1 python tools/ncu_target.py gemm 64 24576 7168 0 > gpurun_out/ncu_d.log 2>&1
tail -2 gpurun_out/ncu_d.log
