#!/usr/bin/env python
"""Time NCCL all-reduce (sum, fp32) at the gradient-bucket sizes of the bench under the current NCCL_* environment.
torchrun --nproc-per-node N scripts/nccl_probe.py [tag]"""
import os
import sys
import torch
import torch.distributed as dist

rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
out = []
for n in (52096, 3222538, 2297130):          # LeNet head, LeNet tail (fc1 + output layer), the C5 net
    t = torch.ones(n, device="cuda")
    for _ in range(10):
        dist.all_reduce(t)
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50):
        dist.all_reduce(t)
    e1.record()
    torch.cuda.synchronize()
    out.append("%d floats: %.1f us" % (n, e0.elapsed_time(e1) / 50 * 1e3))
if rank == 0:
    print("nccl_probe", sys.argv[1] if len(sys.argv) > 1 else "", {k: v for k, v in os.environ.items() if k.startswith("NCCL_")}, "; ".join(out), flush=True)
dist.destroy_process_group()
