#!/bin/bash
# session 8: full GPU suite + default bench (state check after the grouped-mm commit)
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/s8_pytest.log 2>&1
tail -5 gpurun_out/s8_pytest.log
timeout 600 python bench.py > gpurun_out/s8_bench.log 2>&1
tail -3 gpurun_out/s8_bench.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/s8_smoke.log 2>&1; tail -2 gpurun_out/s8_smoke.log
