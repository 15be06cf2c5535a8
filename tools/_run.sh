mkdir -p gpurun_out
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29519 tools/peer_copy_bw.py > gpurun_out/r21_copybw.log 2>&1; echo "rc=$?" >> gpurun_out/r21_copybw.log
tail -2 gpurun_out/r21_copybw.log
