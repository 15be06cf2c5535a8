#!/bin/bash
# session 36: can the issuer sustain 256-clk K blocks?  128-wide pair tiles with immediate hand-back and no promotion (28)
# against the real epilogues (5, 27), the 256-wide no-promotion diagnostic (7) and the default (13), on 32 SMs and the full chip
mkdir -p gpurun_out; export SCALES=row
L=gpurun_out/s36_issuer.log; : > $L
NUM_SMS=32 timeout 300 python tools/ab_gemm.py 4096x4608x7168 13,7,5,27,28 6 >> $L 2>&1
timeout 300 python tools/ab_gemm.py 16384x18432x7168 13,7,5,27,28 6 >> $L 2>&1
cut -c1-220 $L
