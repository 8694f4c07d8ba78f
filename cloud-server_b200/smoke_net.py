"""smoke(): one tiny LeNet train step + predict on cuda:0, checked against the oracle."""
import numpy as np
import torch


def run(ctx):
    from oracle import tfref
    from .net import Net
    rng = np.random.default_rng(0)
    ref = tfref.Net(tfref.LENET)
    w = ref.flatten(tfref.init_params(ref.param_specs, rng))
    x = rng.random((8, 784), dtype=np.float32)
    y = np.eye(10, dtype=np.float32)[rng.integers(0, 10, 8)]
    net = Net(tfref.LENET, ctx)
    net.set_params(w)
    logits = net.forward(torch.from_numpy(x).cuda()).cpu().numpy()
    want = ref.forward(x, ref.unflatten(w))
    assert np.allclose(logits, want, rtol=1e-3, atol=1e-3), "forward mismatch"
    loss, _ = net.train_step(torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda(), "entropy", "sgd", 0.01)
    w2 = w.copy()
    lref, _, _ = tfref.train_step(ref, w2, x, y, tfref.SGD(w2.size, 0.01))
    assert abs(float(loss.item()) - lref) < 1e-3 * max(1.0, abs(lref)), "loss mismatch"
    got = net.get_params()
    assert np.allclose(got, w2, rtol=1e-3, atol=1e-5), "post-step weights mismatch"
    assert net.predict(torch.from_numpy(x).cuda()).shape[0] == 8
