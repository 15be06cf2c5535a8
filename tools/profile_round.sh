#!/bin/bash
# ncu evidence for profiles/: launch list of the bench command + one full capture of each hot kernel.
R=${1:-r02}
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-configs"
$CMD > gpurun_out/plain_bench.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$R.csv $CMD > gpurun_out/ncu_launches.log 2>&1
tail -2 gpurun_out/ncu_launches.log
$CMD > gpurun_out/plain_bench2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:fp8_blockwise_gemm_kernel -s 3 -c 1 -o gpurun_out/gemm_$R $CMD > gpurun_out/ncu_gemm.log 2>&1
tail -2 gpurun_out/ncu_gemm.log
$CMD > gpurun_out/plain_bench3.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:ptgq_kernel -s 3 -c 1 -o gpurun_out/quant_$R $CMD > gpurun_out/ncu_quant.log 2>&1
tail -2 gpurun_out/ncu_quant.log
ls -la gpurun_out | grep $R
# decode shapes (weight streaming): one full capture each of the row-balanced swap-AB kernel the default dispatch picks —
# N=24576 at 16 and 64 tokens (one 168-row range per CTA), N=2112 at 16 tokens (clusters of 4 over K) — and of the 128-row
# tile kernel at M=128
# (r03: decode128 = what is timed at M=128 now, the one-wave 128x256 split-N tiles of variant 14; decode256 = variant 29
# pair tiles at M=256; decoden128 = the 128-wide stream-K tiles at N=2112 M=128)
for T in "decode 16 24576" "decoden 16 2112" "decode64 64 24576" "decode128 128 24576" "decode256 256 24576" "decoden128 128 2112"; do
  set -- $T
  DCMD="python tools/ncu_target.py gemm $2 $3 7168"
  $DCMD > gpurun_out/plain_$1.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:fp8_blockwise_gemm -s 2 -c 1 -o gpurun_out/$1_$R $DCMD > gpurun_out/ncu_$1.log 2>&1
  tail -2 gpurun_out/ncu_$1.log
done
ls -la gpurun_out | grep $R
# (per-token quant kernel unchanged since r01d: profiles/r01d_tokenquant_*)
# the reports together exceed what gpurun copies back (64 MiB): summarise here, ship the summaries, drop the reports
python tools/summarize_profiles.py $R > gpurun_out/summarize_$R.log 2>&1
mkdir -p gpurun_out/prof_$R && cp profiles/${R}_* gpurun_out/prof_$R/
rm -f gpurun_out/*_$R.ncu-rep
ls gpurun_out/prof_$R
