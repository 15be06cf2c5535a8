#!/bin/bash
# session 21: state check after container re-creation — full GPU suite, smoke, both bench arms
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/s21_pytest.log 2>&1; tail -6 gpurun_out/s21_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/s21_smoke.log 2>&1; tail -2 gpurun_out/s21_smoke.log
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/s21_ref.log 2>&1; tail -1 gpurun_out/s21_ref.log | cut -c1-600
timeout 600 python bench.py > gpurun_out/s21_bench.log 2>&1; tail -1 gpurun_out/s21_bench.log | cut -c1-3000
