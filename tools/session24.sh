#!/bin/bash
mkdir -p gpurun_out; export SCALES=row
L=gpurun_out/s24_decode_m.log
for m in 2 4 8 16 24 32 48; do
  timeout 200 python tools/ab_gemm.py ${m}x2112x7168 0,-1 6 >> $L 2>&1
done
SCALES=col timeout 200 python tools/ab_gemm.py 16x2112x7168 0 6 >> $L 2>&1
cut -c1-200 $L
