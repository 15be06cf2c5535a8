#!/bin/bash
# session 38: two staging boxes per epilogue warp (variant 29, 5-stage ring) vs the default (13)
mkdir -p gpurun_out; export SCALES=row
timeout 600 python -m pytest tests/test_gemm_gpu.py -m gpu -x -q -k "every_variant or stream_k" 2>&1 | tail -3
L=gpurun_out/s38_2box.log; : > $L
NUM_SMS=32 timeout 300 python tools/ab_gemm.py 4096x4608x7168 13,29 6 >> $L 2>&1
NUM_SMS=32 timeout 300 python tools/ab_gemm.py 4096x4608x2048 13,29 6 >> $L 2>&1
timeout 300 python tools/ab_gemm.py 16384x18432x7168 13,29 8 >> $L 2>&1
timeout 300 python tools/ab_gemm.py 8192x28672x8192 13,29 6 >> $L 2>&1
timeout 300 python tools/ab_gemm.py 4096x7168x2048 13,29 6 >> $L 2>&1
cut -c1-220 $L
