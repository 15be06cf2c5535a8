#!/bin/bash
# session 39: variant 29 as the prefill default (SGLK_PREFILL_VARIANT=29) — whole GPU suite, smoke, bench; then the bench with 13
mkdir -p gpurun_out
export SGLK_PREFILL_VARIANT=29
( timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/s39_pytest.log 2>&1; tail -2 gpurun_out/s39_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/s39_bench29.log 2>&1; tail -1 gpurun_out/s39_bench29.log | cut -c1-900
SGLK_PREFILL_VARIANT=13 timeout 600 python bench.py --no-cpu-baseline --no-e2e > gpurun_out/s39_bench13.log 2>&1; tail -1 gpurun_out/s39_bench13.log | cut -c1-300
