import sys, os, time; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cloud_server_b200.context import Context
from cloud_server_b200.net import Net
from cloud_server_b200.configs import LENET
ctx = Context(0); net = Net(LENET, ctx); net.set_math(1)
g = torch.Generator(device="cuda").manual_seed(0)
for B in (256, 2048):
    x = torch.rand((B, 784), device="cuda", generator=g)
    y = torch.nn.functional.one_hot(torch.randint(0, 10, (B,), device="cuda", generator=g), 10).float()
    for _ in range(5): net.train_step(x, y)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(20): net.train_step(x, y)
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print("B", B, "host issue time per step %.1f us; total per step %.1f us" % ((t1 - t0) * 5e4, (t2 - t0) * 5e4))
