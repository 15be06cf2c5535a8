#!/bin/bash
# session 30: stream-K fixup with the finisher peek — parity tests, then GPU-side decode time new vs previous build
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gemm_gpu.py tests/test_grouped_gpu.py -m gpu -x -q 2>&1 | tail -3
L=gpurun_out/s30_decode.log; : > $L
LIB=sgl-kernel-xpu_b200/lib/libsglk_sm100a.so
cp $LIB /tmp/new.so
for r in 1; do
  cp /tmp/new.so $LIB; timeout 300 python tools/graph_decode.py new >> $L 2>&1
  cp sgl-kernel-xpu_b200/lib/ab/libsglk_sm100a.so $LIB; timeout 300 python tools/graph_decode.py old >> $L 2>&1
done
cp /tmp/new.so $LIB
cat $L
