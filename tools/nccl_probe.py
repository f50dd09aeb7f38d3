"""What does NCCL deliver between the GPUs of this box at the tensor-parallel engine's message sizes?  Two ranks,
torch.distributed all_reduce (fp32, sum), CUDA-event timed.  usage: torchrun --nproc-per-node 2 tools/nccl_probe.py"""
import os
import torch
import torch.distributed as dist

rank = int(os.environ["RANK"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
for mb in (1.3, 5.2, 42, 168):
    n = int(mb * 1e6 / 4)
    x = torch.ones(n, dtype=torch.float32, device="cuda")
    for _ in range(5):
        dist.all_reduce(x)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 20
    e0.record()
    for _ in range(reps):
        dist.all_reduce(x)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / reps
    if rank == 0:
        print(f"all_reduce fp32 {mb:6.1f} MB: {us:8.1f} us  algbw {mb * 1e6 / us / 1e3:7.1f} GB/s", flush=True)
dist.destroy_process_group()
