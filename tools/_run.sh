mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gemm_gpu.py -x -q -k "32 or decode or rows or ragged" > gpurun_out/r3_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r3_pytest.log
tail -4 gpurun_out/r3_pytest.log
timeout 600 python tools/decode_ab.py 32,0,30 1,16,32,64,128 2112,24576 > gpurun_out/r3_ab.log 2>&1; echo "rc=$?" >> gpurun_out/r3_ab.log
timeout 600 python tools/decode_ab.py 32 16,64 2112 128,1024,4096,7168 >> gpurun_out/r3_ab.log 2>&1; echo "rc=$?" >> gpurun_out/r3_ab.log
timeout 600 python tools/decode_ab.py 32,0,30 8,32,64,128 7168,4096,129280 >> gpurun_out/r3_ab.log 2>&1; echo "rc=$?" >> gpurun_out/r3_ab.log
cat gpurun_out/r3_ab.log
