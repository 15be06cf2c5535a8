#!/bin/bash
# session 37: decomposition on 32 SMs (full clock, no L2 / power limit) and on the full chip:
# 8 = unscaled GEMM (one hand-off per tile), 12 = same issued as split-N halves, 7 = hand-off per K block without promotion,
# 13 = the product kernel; 2048-deep K to see the per-tile share
mkdir -p gpurun_out; export SCALES=row
L=gpurun_out/s37_decomp.log; : > $L
NUM_SMS=32 timeout 300 python tools/ab_gemm.py 4096x4608x7168 8,12,7,13 6 >> $L 2>&1
NUM_SMS=32 timeout 300 python tools/ab_gemm.py 4096x4608x28672 8,7,13 4 >> $L 2>&1
NUM_SMS=32 timeout 300 python tools/ab_gemm.py 4096x4608x2048 8,7,13 6 >> $L 2>&1
timeout 300 python tools/ab_gemm.py 16384x18432x7168 8,12,7,13 6 >> $L 2>&1
cut -c1-220 $L
