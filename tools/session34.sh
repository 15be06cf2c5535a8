#!/bin/bash
# session 34: early hand-back epilogue on 256x128 pair tiles (variant 27) vs the same tiles (5) and the split-N default (13):
# per-SM efficiency on a 32-SM grid (no L2 / power limit), then the full chip
mkdir -p gpurun_out; export SCALES=row
timeout 600 python -m pytest tests/test_gemm_gpu.py -m gpu -x -q -k "every_variant or stream_k" 2>&1 | tail -3
L=gpurun_out/s34_earlyrel.log; : > $L
NUM_SMS=32 timeout 300 python tools/ab_gemm.py 4096x4608x7168 13,5,27 6 >> $L 2>&1
NUM_SMS=8 timeout 300 python tools/ab_gemm.py 2048x2304x7168 13,5,27 6 >> $L 2>&1
timeout 300 python tools/ab_gemm.py 16384x18432x7168 13,5,27 6 >> $L 2>&1
timeout 300 python tools/ab_gemm.py 8192x28672x8192 13,5,27 6 >> $L 2>&1
cut -c1-220 $L
