#!/bin/bash
mkdir -p gpurun_out
( timeout 1200 python -m pytest tests/test_gemm_quant_gpu.py -x -q ) > gpurun_out/s14_pytest.log 2>&1; tail -5 gpurun_out/s14_pytest.log
timeout 300 python tools/gemm_quant_bench.py > gpurun_out/s14_bench.log 2>&1; cat gpurun_out/s14_bench.log
