#!/bin/bash
# session 39 (2 GPUs): variant 29 as the prefill default — full GPU suite, smoke, fused all-gather bitwise check, 1-GPU bench
mkdir -p gpurun_out
( timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/s39_pytest.log 2>&1; tail -2 gpurun_out/s39_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1

timeout 600 python bench.py > gpurun_out/s39_bench.log 2>&1; tail -1 gpurun_out/s39_bench.log | cut -c1-1300
