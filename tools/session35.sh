#!/bin/bash
# session 35: final check of round 1 — full GPU suite, smoke, bench
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/s35_pytest.log 2>&1; tail -4 gpurun_out/s35_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/s35_smoke.log 2>&1; tail -1 gpurun_out/s35_smoke.log
timeout 600 python bench.py > gpurun_out/s35_bench.log 2>&1; tail -1 gpurun_out/s35_bench.log | cut -c1-400
