#!/bin/bash
mkdir -p gpurun_out
echo "== perf gemm diag"; timeout 600 python tools/perf_sweep.py --what gemm --variants 3,7,8,9,10,11 --shapes 8192x28672x8192,16384x18432x7168,8192x8192x8192 > gpurun_out/perf_diag.jsonl 2>gpurun_out/perf_diag.err; tail -2 gpurun_out/perf_diag.err
python - <<'PY'
import json
for l in open('gpurun_out/perf_diag.jsonl'):
    d=json.loads(l)
    if d['op']=='gemm' and 'error' not in d:
        print(f"gemm {d['M']:6d}x{d['N']:6d}x{d['K']:6d} v{d['variant']:2d} {d['name']:32s} {d['us_med']:10.1f} us {d['TFLOPS_med']:8.1f} TF (best {d['TFLOPS_min']:.1f})")
    else: print(d)
PY
for v in 8 3; do
python tools/ncu_target.py gemm 8192 8192 8192 $v > gpurun_out/plain_g$v.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:fp8_blockwise_gemm -s 1 -c 1 -o gpurun_out/gemm_r1b_v$v python tools/ncu_target.py gemm 8192 8192 8192 $v > gpurun_out/ncu_g$v.log 2>&1
tail -2 gpurun_out/ncu_g$v.log
done
