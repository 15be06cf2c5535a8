mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_grouped_gpu.py tests/test_guard_bands_gpu.py tests/test_gemm_quant_gpu.py -x -q > gpurun_out/r23_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r23_pytest.log
tail -4 gpurun_out/r23_pytest.log
timeout 300 python tools/grouped_bench.py 8,32,64 > gpurun_out/r23_grouped.log 2>&1; echo "rc=$?" >> gpurun_out/r23_grouped.log
cat gpurun_out/r23_grouped.log
