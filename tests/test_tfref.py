"""CPU tests of the CNN oracle (oracle/tfref.py): TF-1.x semantics restated in numpy, checked
against (a) hand-computed known answers, (b) torch-CPU autograd in float64 with the asymmetric
SAME padding written out explicitly, (c) finite differences.  PARITY UNPINNED vs TensorFlow
itself (not installable here; the reference pins no TF result)."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import tfref
from oracle.torchref import torch_forward, TorchStep


CONFIGS = {
    "api_example": (tfref.API_EXAMPLE, 28, 28, 1),
    "lenet": (tfref.LENET, 28, 28, 1),
    "strided_valid_leaky": ({"middle_layer": [
        {"layer": "conv", "filter": [3, 3, 4], "stride": [2, 2], "padding": "VALID", "isBias": "True"},
        {"layer": "active", "active_func": "leaky_relu", "param": [0.1]},
        {"layer": "conv", "filter": [2, 3, 5], "stride": [1, 2], "padding": "SAME"},
        {"layer": "pool", "kernel": [3, 3], "stride": [2, 2]}, {"layer": "norm", "epsilon": 1e-2},
        {"layer": "active", "active_func": "relu"},
        {"layer": "connect", "hidden": 16}, {"layer": "norm"}, {"layer": "active"}]}, 13, 11, 2),
    "deep_small": (tfref.DEEPNET, 8, 8, 3),
}


def test_param_counts_match_survey():
    assert tfref.Net(tfref.LENET).param_count() == 3274634
    assert tfref.Net(tfref.DEEPNET, 32, 32, 3).param_count() == 2297130


def test_defaults_of_cnn():
    layers, params = tfref.parse_layers(tfref.API_EXAMPLE)
    assert [l["kind"] for l in layers] == ["conv", "conv", "pool", "norm", "active", "connect", "connect", "connect"]
    assert layers[0]["b"] is None and layers[4]["fn"] == "sigmoid" and layers[5]["fout"] == 512
    assert layers[0]["pt"] == 0 and layers[0]["oh"] == 28       # 2x2 SAME: pad 1 total, all on bottom/right
    # isBias is True unless the JSON value is the *string* "False" (construct_distribute.py:228)
    l2, _ = tfref.parse_layers({"middle_layer": [{"layer": "conv", "filter": [1, 1, 1], "isBias": False}]})
    assert l2[0]["b"] is not None
    l3, _ = tfref.parse_layers({"middle_layer": [{"layer": "conv", "filter": [1, 1, 1], "isBias": "False"}]})
    assert l3[0]["b"] is None


def test_known_answers_small():
    # 2x2 SAME conv of ones on a 3x3 ramp: padding goes to bottom/right only
    net = tfref.Net({"middle_layer": [{"layer": "conv", "filter": [2, 2, 1]}]}, 3, 3, 1, classes=1, dtype=np.float64)
    x = np.arange(9, dtype=np.float64).reshape(1, 9)
    y, _ = tfref.conv2d_fwd(x.reshape(1, 3, 3, 1), np.ones((2, 2, 1, 1)), net.layers[0])
    assert np.array_equal(y[0, :, :, 0], [[8, 12, 7], [20, 24, 13], [13, 15, 8]])
    # 2x2/2 SAME max pool on 3x3: windows hang over the bottom/right edge
    L = tfref.parse_layers({"middle_layer": [{"layer": "pool"}]}, 3, 3, 1)[0][0]
    p, arg = tfref.maxpool_fwd(x.reshape(1, 3, 3, 1), L)
    assert np.array_equal(p[0, :, :, 0], [[4, 5], [7, 8]])
    # softmax CE on 2x3
    n2 = tfref.Net({"middle_layer": []}, 1, 1, 3, classes=3, dtype=np.float64)
    loss, dz = n2.loss_and_grad(np.array([[1.0, 2.0, 3.0], [0.0, 0.0, 0.0]]), np.array([[0, 0, 1.0], [1.0, 0, 0]]))
    assert abs(loss - (0.40760596444438046 + np.log(3)) / 2) < 1e-12
    assert np.allclose(dz[1], (np.ones(3) / 3 - [1, 0, 0]) / 2)
    # Adagrad two steps (acc0 = 0.1, no epsilon)
    w = np.array([1.0], np.float64)
    opt = tfref.Adagrad(1, lr=0.5, dtype=np.float64)
    opt.apply(w, np.array([2.0]))
    assert np.allclose(w, 1 - 0.5 * 2 / np.sqrt(4.1))
    opt.apply(w, np.array([1.0]))
    assert np.allclose(w, 1 - 0.5 * 2 / np.sqrt(4.1) - 0.5 / np.sqrt(5.1))
    # Adam first step moves by ~lr
    w = np.array([1.0], np.float64)
    tfref.Adam(1, lr=0.01, dtype=np.float64).apply(w, np.array([3.0]))
    assert abs(w[0] - (1 - 0.01)) < 1e-6


@pytest.mark.parametrize("name", list(CONFIGS))
@pytest.mark.parametrize("loss", ["entropy", "mse"])
def test_forward_backward_vs_torch_autograd(name, loss):
    cfg, h, w, c = CONFIGS[name]
    rng = np.random.default_rng(hash(name) % 1000)
    net = tfref.Net(cfg, h, w, c, dtype=np.float64)
    params = tfref.init_params(net.param_specs, rng, np.float64)
    params = [p + 0.05 * rng.standard_normal(p.shape) for p in params]
    n = 4
    x = rng.random((n, h * w * c))
    y = np.eye(10)[rng.integers(0, 10, n)]
    logits, cache = net.forward(x, params, keep=True)
    lv, dz = net.loss_and_grad(logits, y, loss)
    grads = net.backward(dz, params, cache)

    tp = [torch.tensor(p, dtype=torch.float64, requires_grad=True) for p in params]
    tl = torch_forward(net, torch.tensor(x), tp)
    assert np.allclose(tl.detach().numpy(), logits, rtol=1e-9, atol=1e-10)
    ty = torch.tensor(y)
    tloss = (-(ty * torch.log_softmax(tl, 1)).sum(1)).mean() if loss == "entropy" else ((ty - tl) ** 2).mean()
    assert abs(float(tloss) - lv) < 1e-9
    tloss.backward()
    for g, t, spec in zip(grads, tp, net.param_specs):
        assert np.allclose(g, t.grad.numpy(), rtol=1e-7, atol=1e-10), spec[0]


def test_finite_differences():
    cfg, h, w, c = CONFIGS["strided_valid_leaky"]
    rng = np.random.default_rng(5)
    net = tfref.Net(cfg, h, w, c, dtype=np.float64)
    params = [p + 0.1 * rng.standard_normal(p.shape) for p in tfref.init_params(net.param_specs, rng, np.float64)]
    flat = net.flatten(params)
    x = rng.random((3, h * w * c))
    y = np.eye(10)[rng.integers(0, 10, 3)]

    def f(fl):
        return net.loss_and_grad(net.forward(x, net.unflatten(fl)), y, "entropy")[0]

    logits, cache = net.forward(x, params, keep=True)
    g = net.flatten(net.backward(net.loss_and_grad(logits, y, "entropy")[1], params, cache))
    for i in rng.integers(0, flat.size, 25):
        e = np.zeros_like(flat)
        e[i] = 1e-6
        fd = (f(flat + e) - f(flat - e)) / 2e-6
        assert abs(fd - g[i]) < 1e-6 * max(1, abs(fd)), i


def test_dataset_next_batch_semantics():
    imgs = (np.arange(7 * 4).reshape(7, 2, 2, 1) % 256).astype(np.uint8)
    labs = np.eye(10)[np.arange(7) % 10]
    ds = tfref.DataSet(imgs, labs, rs=np.random.RandomState(0))
    seen = []
    for _ in range(5):
        xb, yb = ds.next_batch(3)
        assert xb.shape == (3, 4) and xb.dtype == np.float32
        seen.append(yb.argmax(1))
    first_epoch = np.concatenate(seen[:2] + [seen[2][:1]])
    assert sorted(first_epoch.tolist()) == list(range(7))      # wrap batch = tail + head of reshuffle
    assert np.allclose(ds._images.max(), np.float32(27) * np.float32(1 / 255.0))


@pytest.mark.parametrize("opt,lr", [("adagrad", 0.05), ("sgd", 0.05)])
def test_torch_cpu_step_equals_tfref_step(opt, lr):
    """oracle.torchref.TorchStep (bench.py's CPU arm for the train step) walks the same trajectory as tfref.train_step."""
    rng = np.random.default_rng(3)
    ref = tfref.Net(tfref.LENET)
    w = ref.flatten(tfref.init_params(ref.param_specs, rng))
    x = rng.random((16, 784), dtype=np.float32)
    y = np.eye(10, dtype=np.float32)[rng.integers(0, 10, 16)]
    ts = TorchStep(tfref.LENET, flat_params=w.copy(), opt=opt, lr=lr, threads=2)
    o = tfref.Adagrad(w.size, lr) if opt == "adagrad" else tfref.SGD(w.size, lr)
    w0 = w.copy()
    for _ in range(3):
        l0, _, _ = tfref.train_step(ref, w, x, y, o)
        l1 = ts.step(torch.from_numpy(x), torch.from_numpy(y))
        assert abs(l0 - l1) < 1e-4 * max(1.0, abs(l0))
    d0, d1 = w - w0, ts.flat() - w0
    assert np.abs(d0 - d1).max() < 1e-3 * np.abs(d0).max()
