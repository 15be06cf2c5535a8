#!/bin/bash
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_token_quant_gpu.py -x -q ) > gpurun_out/s9_pytest.log 2>&1
tail -15 gpurun_out/s9_pytest.log
timeout 300 python tools/token_quant_bw.py > gpurun_out/s9_bw.log 2>&1; cat gpurun_out/s9_bw.log
