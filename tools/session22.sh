#!/bin/bash
# session 22: rasterisation group sweep (time + DRAM traffic per setting)
mkdir -p gpurun_out
for s in 16384x18432x7168 8192x28672x8192 16384x2304x7168 4096x7168x2048; do
  timeout 300 python tools/group_m_sweep.py $s 2,4,8,16,32,64 8 >> gpurun_out/s22_sweep.log 2>&1
done
cat gpurun_out/s22_sweep.log
ONCE=1 timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none \
  -k regex:fp8_blockwise --csv --log-file gpurun_out/s22_ncu_traffic.csv python tools/group_m_sweep.py 16384x18432x7168 2,4,8,16,32,64 > gpurun_out/s22_ncu.log 2>&1
grep -E "dram__bytes|gpu__time" gpurun_out/s22_ncu_traffic.csv | awk -F'","' '{print $(NF-2), $(NF-1), $NF}' | tr -d '"'
