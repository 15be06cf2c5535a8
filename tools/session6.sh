#!/bin/bash
mkdir -p gpurun_out
echo "== gemm tests"; timeout 1200 python -m pytest tests/test_gemm_gpu.py -q -m gpu --timeout 600 -x 2>&1 | tail -8
echo "== perf decode"; timeout 600 python tools/perf_sweep.py --what gemm --variants 0,2,3 --shapes 1x24576x7168,16x24576x7168,64x24576x7168,128x24576x7168,256x24576x7168,1x2112x7168,64x2112x7168,256x2112x7168,16384x18432x7168 > gpurun_out/perf_s6.jsonl 2>gpurun_out/perf_s6.err; tail -2 gpurun_out/perf_s6.err
python - <<'PY'
import json
for l in open('gpurun_out/perf_s6.jsonl'):
    d=json.loads(l)
    if d['op']=='gemm' and 'error' not in d:
        print(f"gemm {d['M']:6d}x{d['N']:6d}x{d['K']:6d} v{d['variant']:2d} {d['name']:32s} {d['us_med']:10.1f} us {d['TFLOPS_med']:8.1f} TF {d['GBps_med']:8.1f} GB/s (best {d['us_min']:.1f} us)")
    else: print(d)
PY
