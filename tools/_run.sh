mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gemm_gpu.py -x -q > gpurun_out/r6_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r6_pytest.log
tail -4 gpurun_out/r6_pytest.log
timeout 600 python tools/decode_ab.py -1,0 1,16,64,128 2112,24576 > gpurun_out/r6_ab.log 2>&1; echo "rc=$?" >> gpurun_out/r6_ab.log
cat gpurun_out/r6_ab.log
