#!/bin/bash
# stream-K on/off across wave efficiencies (row-major scales so the scale path is identical)
for s in 4096x7168x2048 4096x4608x7168 2048x7168x7168 8192x4096x7168 1024x7168x2048 512x24576x7168 3072x6144x4096 6144x6144x4096; do
  echo "== $s ws"; SCALES=row python tools/ab_gemm.py $s 3 5 | cut -c1-140
  echo "== $s nows"; SCALES=row NOWS=1 python tools/ab_gemm.py $s 3 5 | cut -c1-140
done
for s in 256x24576x7168 128x24576x7168 64x2112x7168 256x2112x7168 1x24576x7168 1024x4608x7168 768x7168x2048; do
  echo "== $s ws"; SCALES=row python tools/ab_gemm.py $s -1 5 | cut -c1-140
  echo "== $s nows"; SCALES=row NOWS=1 python tools/ab_gemm.py $s -1 5 | cut -c1-140
done
