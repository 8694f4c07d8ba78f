#!/usr/bin/env python
"""Run one hot-path kernel a few times (for ncu captures): python scripts/run_kernel.py <name> [batch]
names: train_chain, predict_chain, lenet_train, lenet_predict, deep_train"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cloud_server_b200.context import Context

name = sys.argv[1]
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
math = int(sys.argv[3]) if len(sys.argv) > 3 else 0
ctx = Context(0)
g = torch.Generator(device="cuda").manual_seed(0)
if name in ("train_chain", "predict_chain"):
    x = torch.randint(0, 256, (n, 768, 1024, 3), dtype=torch.uint8, device="cuda", generator=g)
    for _ in range(4):
        if name == "train_chain":
            ctx.resize_gray_norm(x)
        else:
            ctx.predict_chain(x, channel_order="BGR")
else:
    from cloud_server_b200.net import Net
    from cloud_server_b200.configs import LENET, DEEPNET
    deep = name.startswith("deep")
    net = Net(DEEPNET, ctx, 32, 32, 3) if deep else Net(LENET, ctx)
    net.set_math(math)
    x = torch.rand((n, 32 * 32 * 3 if deep else 784), device="cuda", generator=g)
    y = torch.nn.functional.one_hot(torch.randint(0, 10, (n,), device="cuda", generator=g), 10).float()
    for _ in range(4):
        if name in ("lenet_train", "deep_train"):
            net.train_step(x, y)
        else:
            net.predict(x)
torch.cuda.synchronize()
print("ok", ctx.launches)
