#!/bin/bash
mkdir -p gpurun_out
timeout 180 python tools/variant_eq.py 3 23 > gpurun_out/s15_eq.log 2>&1; echo "rc=$?" >> gpurun_out/s15_eq.log; cat gpurun_out/s15_eq.log
SCALES=row timeout 200 python tools/ab_gemm.py 8192x8192x8192 3,23,2 4 > gpurun_out/s15_ab.log 2>&1; cat gpurun_out/s15_ab.log
SCALES=row timeout 200 python tools/ab_gemm.py 16384x18432x7168 3,23 3 > gpurun_out/s15_ab2.log 2>&1; cat gpurun_out/s15_ab2.log
