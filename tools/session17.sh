#!/bin/bash
mkdir -p gpurun_out
for v in 24 25 26; do timeout 180 python tools/variant_eq.py 3 $v 300x512x1024 1000x1280x7168 384x2304x7168 4096x4096x2048 > gpurun_out/s17_eq$v.log 2>&1; echo "variant $v rc=$?"; grep -c " 0 / " gpurun_out/s17_eq$v.log; grep -v " 0 / " gpurun_out/s17_eq$v.log | head -5; done
SCALES=row timeout 300 python tools/ab_gemm.py 8192x8192x8192 3,24,13,25,23,26 4 > gpurun_out/s17_ab.log 2>&1; cat gpurun_out/s17_ab.log
SCALES=row timeout 300 python tools/ab_gemm.py 16384x18432x7168 3,24,13,25,23,26 3 > gpurun_out/s17_ab2.log 2>&1; cat gpurun_out/s17_ab2.log
