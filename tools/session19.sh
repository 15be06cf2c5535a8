#!/bin/bash
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/s19_pytest.log 2>&1; tail -6 gpurun_out/s19_pytest.log
timeout 600 python bench.py > gpurun_out/s19_bench.log 2>&1; tail -1 gpurun_out/s19_bench.log | cut -c1-1500
