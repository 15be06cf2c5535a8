#!/bin/bash
mkdir -p gpurun_out
echo "== pytest gpu"; timeout 1500 python -m pytest tests -q -m gpu --timeout 900 -x 2>&1 | tail -15
echo "== perf quant"; timeout 300 python tools/perf_sweep.py --what quant > gpurun_out/perf_quant.jsonl 2>gpurun_out/perf_quant.err; tail -2 gpurun_out/perf_quant.err
python - <<'PY'
import json
for l in open('gpurun_out/perf_quant.jsonl'):
    d=json.loads(l)
    print(f"quant M={d['M']:6d} col={d['col_major']!s:5} {d['us_med']:9.1f} us  {d['GBps_med']:8.1f} GB/s (best {d['GBps_min']:.1f})")
PY
echo "== perf gemm"; timeout 600 python tools/perf_sweep.py --what gemm --variants 0,2,3,4,5,7 --shapes 8192x28672x8192,16384x18432x7168,4096x7168x2048,256x24576x7168 > gpurun_out/perf_gemm.jsonl 2>gpurun_out/perf_gemm.err; tail -2 gpurun_out/perf_gemm.err
python - <<'PY'
import json
for l in open('gpurun_out/perf_gemm.jsonl'):
    d=json.loads(l)
    if d['op']=='gemm' and 'error' not in d:
        print(f"gemm {d['M']:6d}x{d['N']:6d}x{d['K']:6d} v{d['variant']:2d} {d['name']:32s} {d['us_med']:10.1f} us {d['TFLOPS_med']:8.1f} TF (best {d['TFLOPS_min']:.1f})")
    else: print(d)
PY
