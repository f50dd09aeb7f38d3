nvidia-smi topo -m | head -5
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29521 tools/nccl_probe.py 2>&1 | grep all_reduce
NCCL_DEBUG=INFO timeout 600 python -m pytest tests/test_tp_gpu.py -m gpu -x -q -s 2>&1 | grep -E "NCCL INFO (Channel|Connected|comm|Using|P2P|NVLS|ncclCommInitRank)|passed|failed" | head -30
