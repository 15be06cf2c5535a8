mkdir -p gpurun_out
timeout 120 python -m pytest tests/test_gated_gpu.py -x -q 2>&1 | tail -2
(timeout 300 python tools/gated_ab.py 8,4,2) > gpurun_out/r18_gated.log 2>&1; echo "rc=$?" >> gpurun_out/r18_gated.log
cat gpurun_out/r18_gated.log
