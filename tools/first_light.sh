#!/bin/bash
# First GPU session: smoke, per-variant GEMM checks (each under its own timeout), quant tests, perf sweep.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
echo "== smoke" ; timeout 300 python __graft_entry__.py smoke 2>&1 | tail -5
echo "== quant tests" ; timeout 900 python -m pytest tests/test_quant_gpu.py -q -m gpu -x --timeout 600 2>&1 | tail -15
for v in 0 1 2 3 4 5 6; do
  echo "== variant $v" ; timeout 120 python tools/variant_check.py $v 2>&1 | tail -25
  echo "exit=$?"
done
echo "== perf" ; timeout 600 python tools/perf_sweep.py > gpurun_out/perf_sweep.jsonl 2> gpurun_out/perf_sweep.err ; tail -3 gpurun_out/perf_sweep.err
cat gpurun_out/perf_sweep.jsonl
