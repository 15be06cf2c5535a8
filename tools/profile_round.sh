#!/bin/bash
# ncu evidence for profiles/: launch list of the bench command + one full capture of each hot kernel.
R=${1:-r01}
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/plain_bench.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$R.csv $CMD > gpurun_out/ncu_launches.log 2>&1
tail -2 gpurun_out/ncu_launches.log
$CMD > gpurun_out/plain_bench2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:fp8_blockwise_gemm -s 3 -c 1 -o gpurun_out/gemm_$R $CMD > gpurun_out/ncu_gemm.log 2>&1
tail -2 gpurun_out/ncu_gemm.log
$CMD > gpurun_out/plain_bench3.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:ptgq_kernel -s 3 -c 1 -o gpurun_out/quant_$R $CMD > gpurun_out/ncu_quant.log 2>&1
tail -2 gpurun_out/ncu_quant.log
ls -la gpurun_out | grep $R
# decode shape (weight streaming, stream-K): one full capture of the 128-wide kernel
DCMD="python tools/ncu_target.py gemm 64 24576 7168"
$DCMD > gpurun_out/plain_decode.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:fp8_blockwise_gemm -s 2 -c 1 -o gpurun_out/decode_$R $DCMD > gpurun_out/ncu_decode.log 2>&1
tail -2 gpurun_out/ncu_decode.log
ls -la gpurun_out | grep $R
# per-token quant (sgl_per_token_quant_fp8): one full capture
TCMD="python tools/ncu_target.py tokenquant 32768"
$TCMD > gpurun_out/plain_tokenquant.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:ptq_kernel -s 2 -c 1 -o gpurun_out/tokenquant_$R $TCMD > gpurun_out/ncu_tokenquant.log 2>&1
tail -2 gpurun_out/ncu_tokenquant.log
ls -la gpurun_out | grep $R
