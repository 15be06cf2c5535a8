#!/bin/bash
mkdir -p gpurun_out
SCALES=row timeout 300 python tools/ab_gemm.py 8192x8192x8192 3,2,14,13,5 4 > gpurun_out/s13_ab.log 2>&1; cat gpurun_out/s13_ab.log
SCALES=row timeout 300 python tools/ab_gemm.py 16384x18432x7168 3,2,13 3 > gpurun_out/s13_ab2.log 2>&1; cat gpurun_out/s13_ab2.log
