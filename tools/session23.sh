#!/bin/bash
# session 23: decode-shape variant sweep (row-major scales = the quant op's default layout)
mkdir -p gpurun_out; export SCALES=row
L=gpurun_out/s23_decode.log
for m in 1 16 64 128; do
  timeout 200 python tools/ab_gemm.py ${m}x2112x7168 0,6 6 >> $L 2>&1
  timeout 200 python tools/ab_gemm.py ${m}x24576x7168 0,6 6 >> $L 2>&1
done
for m in 192 256; do
  timeout 200 python tools/ab_gemm.py ${m}x2112x7168 13,5,0,6 6 >> $L 2>&1
  timeout 200 python tools/ab_gemm.py ${m}x24576x7168 13,5,3,0 6 >> $L 2>&1
done
cut -c1-200 $L
