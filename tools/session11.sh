#!/bin/bash
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_quant_gpu.py -x -q ) > gpurun_out/s11_pytest.log 2>&1; tail -3 gpurun_out/s11_pytest.log
timeout 600 python tools/quant_exp.py 2,13 16384,32768 > gpurun_out/s11_exp.log 2>&1; cat gpurun_out/s11_exp.log
timeout 600 python tools/quant_bw.py > gpurun_out/s11_bw.log 2>&1; cat gpurun_out/s11_bw.log
