#!/bin/bash
mkdir -p gpurun_out
SCALES=row timeout 300 python tools/ab_gemm.py 16384x18432x7168 3,13,25 8 > gpurun_out/s18_a.log 2>&1; cat gpurun_out/s18_a.log
SCALES=row timeout 300 python tools/ab_gemm.py 8192x28672x8192 3,13,25 5 > gpurun_out/s18_b.log 2>&1; cat gpurun_out/s18_b.log
SCALES=col timeout 300 python tools/ab_gemm.py 16384x18432x7168 3,13,25 5 > gpurun_out/s18_c.log 2>&1; cat gpurun_out/s18_c.log
SCALES=row timeout 300 python tools/ab_gemm.py 2048x7168x7168 3,13,25 8 > gpurun_out/s18_d.log 2>&1; cat gpurun_out/s18_d.log
SCALES=row timeout 300 python tools/ab_gemm.py 4096x7168x2048 3,13,25 8 > gpurun_out/s18_e.log 2>&1; cat gpurun_out/s18_e.log
SCALES=row timeout 300 python tools/ab_gemm.py 16384x2304x7168 3,13,25 8 > gpurun_out/s18_f.log 2>&1; cat gpurun_out/s18_f.log
