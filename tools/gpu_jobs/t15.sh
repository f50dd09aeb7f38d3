timeout 900 python tools/tp_timeline.py 64 2048 2 4 > gpurun_out/t15_tp_peer.txt 2>&1
VGATE_B200_TP_ALLREDUCE=nccl timeout 900 python tools/tp_timeline.py 64 2048 2 4 > gpurun_out/t15_tp_nccl.txt 2>&1
tail -3 gpurun_out/t15_tp_peer.txt
