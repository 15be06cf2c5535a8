#!/bin/bash
# session 31: generalised finisher peek (all tile widths) — parity tests, decode + stream-K prefill shapes, new vs HEAD build
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gemm_gpu.py tests/test_grouped_gpu.py tests/test_gemm_quant_gpu.py -m gpu -x -q 2>&1 | tail -3
L=gpurun_out/s31_decode.log; : > $L
LIB=sgl-kernel-xpu_b200/lib/libsglk_sm100a.so
cp $LIB /tmp/new.so
for r in 1 2; do
  for w in new old; do
    if [ $w = new ]; then cp /tmp/new.so $LIB; else cp sgl-kernel-xpu_b200/lib/ab/libsglk_sm100a.so $LIB; fi
    timeout 300 python tools/graph_decode.py $w >> $L 2>&1
    for s in 2048x7168x7168 256x24576x7168 1024x24576x1536; do
      echo "$w" >> $L; SCALES=row timeout 200 python tools/ab_gemm.py $s -1 6 >> $L 2>&1
    done
  done
done
cp /tmp/new.so $LIB
cut -c1-160 $L
