#!/bin/bash
mkdir -p gpurun_out
echo "== gemm tests"; timeout 900 python -m pytest tests/test_gemm_gpu.py -q -m gpu --timeout 600 -x 2>&1 | tail -5
echo "== perf gemm"; timeout 600 python tools/perf_sweep.py --what gemm --variants 2,3,7,8 --shapes 8192x28672x8192,16384x18432x7168,8192x8192x8192,4096x7168x2048 > gpurun_out/perf_s5.jsonl 2>gpurun_out/perf_s5.err; tail -2 gpurun_out/perf_s5.err
python - <<'PY'
import json
for l in open('gpurun_out/perf_s5.jsonl'):
    d=json.loads(l)
    if d['op']=='gemm' and 'error' not in d:
        print(f"gemm {d['M']:6d}x{d['N']:6d}x{d['K']:6d} v{d['variant']:2d} {d['name']:32s} {d['us_med']:10.1f} us {d['TFLOPS_med']:8.1f} TF (best {d['TFLOPS_min']:.1f})")
    else: print(d)
PY
