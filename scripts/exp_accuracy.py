#!/usr/bin/env python
"""Accuracy of the tensor-core math modes at the BASELINE sizes against oracle/tfref.py (GPU box):
python scripts/exp_accuracy.py [math ...]   -- prints one line per (net, batch, math, optimizer, step)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from oracle import tfref
from cloud_server_b200.context import Context
from cloud_server_b200.net import Net


def rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def main():
    maths = [int(a) for a in sys.argv[1:]] or [1, 3]
    ctx = Context(0)
    cases = [("lenet", tfref.LENET, (28, 28, 1), 256), ("lenet", tfref.LENET, (28, 28, 1), 2048), ("deep", tfref.DEEPNET, (32, 32, 3), 512)]
    for name, cfg, shp, B in cases:
        rng = np.random.default_rng(11)
        ref = tfref.Net(cfg, *shp)
        flat = ref.flatten(tfref.init_params(ref.param_specs, rng))
        x = rng.random((B, int(np.prod(shp))), dtype=np.float32)
        y = np.eye(10, dtype=np.float32)[rng.integers(0, 10, B)]
        xd, yd = torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda()
        params = ref.unflatten(flat)
        t0 = time.time()
        logits, cache = ref.forward(x, params, keep=True)
        lref, dz = ref.loss_and_grad(logits, y, "entropy")
        gref = ref.flatten(ref.backward(dz, params, cache))
        print("# %s B=%d oracle fwd+bwd %.1fs" % (name, B, time.time() - t0), flush=True)
        for math in maths:
            net = Net(cfg, ctx, *shp)
            net.set_math(math)
            net.set_params(flat)
            got = net.forward(xd).cpu().numpy()
            lg, _ = net.backward(xd, yd, "entropy")
            g = net.grads.cpu().numpy()
            worst = 0.0
            per = []
            gmax = float(np.abs(gref).max())
            for (off, size, shape), spec in zip(net.tensors, ref.param_specs):
                e = float(np.abs(g[off:off + size].astype(np.float64) - gref[off:off + size]).max()) / max(float(np.abs(gref[off:off + size]).max()), 1e-2 * gmax)
                per.append("%s=%.1e" % (spec[0], e))
                worst = max(worst, e)
            print("%s B=%d math=%d logits %.2e loss %.2e grad worst %.2e [%s]" % (name, B, math, rel(got, logits), abs(float(lg.item()) - lref) / max(1, abs(lref)), worst, " ".join(per)), flush=True)
            for opt, lr in (("sgd", 0.05), ("adagrad", 0.05), ("compat", 1e-4)):
                net.set_params(flat)
                net.set_state(0, np.full(flat.size, 0.1, np.float32), 0)
                w = flat.copy()
                o = {"sgd": lambda: tfref.SGD(w.size, lr), "adagrad": lambda: tfref.Adagrad(w.size, lr), "compat": lambda: tfref.Adagrad(w.size, 1e-4)}[opt]()
                for step in range(5):
                    l0, _, _ = tfref.train_step(ref, w, x, y, o)
                    l1, _ = net.train_step(xd, yd, "entropy", opt, lr)
                    got_w = net.get_params()
                    print("   %s B=%d math=%d %s step %d: loss %.6f vs %.6f (%.1e)  weights %.2e  update %.2e" % (
                        name, B, math, opt, step, float(l1.item()), l0, abs(float(l1.item()) - l0) / max(1, abs(l0)), rel(got_w, w), rel(got_w - flat, w - flat)), flush=True)
            net.close()


if __name__ == "__main__":
    main()
