#!/bin/bash
# session 32: final state of round 1 — full GPU suite, smoke, both bench arms, ncu evidence (profiles r01d)
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/s32_pytest.log 2>&1; tail -6 gpurun_out/s32_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/s32_smoke.log 2>&1; tail -2 gpurun_out/s32_smoke.log
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/s32_ref.log 2>&1; tail -1 gpurun_out/s32_ref.log | cut -c1-300
timeout 600 python bench.py > gpurun_out/s32_bench.log 2>&1; tail -1 gpurun_out/s32_bench.log | cut -c1-1500
bash tools/profile_round.sh r01d > gpurun_out/s32_profile.log 2>&1; tail -3 gpurun_out/s32_profile.log
