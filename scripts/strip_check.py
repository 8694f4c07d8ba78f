#!/usr/bin/env python
"""Check the picture-resident input-gradient kernel (conv_strip.cu) against the im2col kernel on LeNet, math mode 3:
gradients of every parameter with NC_CONV_STRIP unset / 1 (plain descriptors) / 2 (descriptors with base offset), and the
stage time of the conv2 input gradient."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from cloud_server_b200.context import Context
from cloud_server_b200.net import Net
from cloud_server_b200.configs import LENET

ctx = Context(0)
for n in (37, 2048, 256):
    g = torch.Generator(device="cuda").manual_seed(0)
    x = torch.rand((n, 784), device="cuda", generator=g)
    y = torch.nn.functional.one_hot(torch.randint(0, 10, (n,), device="cuda", generator=g), 10).float()
    ref = None
    for mode in ("0", "1"):
        os.environ["NC_CONV_STRIP"] = mode
        net = Net(LENET, ctx, seed=3)
        net.backward(x, y, "entropy")
        torch.cuda.synchronize()
        gr = net.grads.cpu().numpy().copy()
        st = net.profile(lambda: net.backward(x, y, "entropy"))
        t = [ms for nm, ms in st if "dgrad" in nm and "conv" in nm]
        if ref is None:
            ref = gr
        # conv1 W/b gradients are the ones downstream of conv2's input gradient: the first 832 values
        d_all = np.abs(gr - ref).max() / np.abs(ref).max()
        d_c1 = np.abs(gr[:832] - ref[:832]).max() / np.abs(ref[:832]).max()
        print("n=%d NC_CONV_STRIP=%s: conv dgrad stage %s ms; |g - g_im2col| / |g| all %.3e, conv1 %.3e, finite %s"
              % (n, mode, ["%.4f" % v for v in t], d_all, d_c1, bool(np.isfinite(gr).all())), flush=True)
        net.close()
