mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gated_gpu.py -x -q > gpurun_out/r7_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r7_pytest.log
tail -30 gpurun_out/r7_pytest.log
