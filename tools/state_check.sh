#!/bin/bash
# Round state check on one B200: whole GPU suite, smoke, default bench.  Logs -> gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > gpurun_out/sc_gpu.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -x -q --durations=15 > gpurun_out/sc_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/sc_pytest.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/sc_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/sc_smoke.log
timeout 600 python bench.py > gpurun_out/sc_bench.json 2> gpurun_out/sc_bench.err; echo "bench rc=$?" >> gpurun_out/sc_bench.err
tail -3 gpurun_out/sc_pytest.log; tail -2 gpurun_out/sc_smoke.log; tail -c 600 gpurun_out/sc_bench.json
