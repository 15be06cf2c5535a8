#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/quant_exp.py 2,13,14,15,16 16384,32768 > gpurun_out/s10_exp2.log 2>&1; cat gpurun_out/s10_exp2.log
timeout 600 python tools/quant_bw.py > gpurun_out/s10_bw.log 2>&1; cat gpurun_out/s10_bw.log
