#!/bin/bash
mkdir -p gpurun_out
SCALES=row timeout 300 python tools/ab_gemm.py 8192x8192x8192 3,15,17,18,19,20,21 3 > gpurun_out/s16_ab.log 2>&1; cat gpurun_out/s16_ab.log
