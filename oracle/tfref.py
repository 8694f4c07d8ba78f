"""numpy restatement of the CNN half of the hot path (TEST INFRASTRUCTURE ONLY).

PARITY UNPINNED: TensorFlow 1.x (requirment.txt:11, unpinned, API level 1.0-1.4) cannot be
installed here and the reference has no test or golden vector for any TF result.  This file
restates, line by line, what apps/construction/util/construct_distribute.py asks TF to compute
(identical copy in construct_inference.py:53-279), using TF-1.x's documented semantics; it is
cross-checked against torch-CPU autograd in float64 and finite differences (tests/test_tfref.py).

Layout: activations NHWC, conv filters HWIO, dense weights [in, out]; parameters in variable
creation order.  dtype float32 by default (TF's), float64 for shadow runs.
"""
from __future__ import annotations

import json
import math

import numpy as np


# ----------------------------------------------------------------------------- config -> layers
def parse_layers(net_config, in_h=28, in_w=28, in_c=1, classes=10):
    """cnn() construct_distribute.py:208-265: defaults at :222-250, implicit output layer :262-264.
    Returns a list of dict layer specs with resolved shapes, and the list of parameter tensors
    [(name, shape, init_kind, init_arg)] in creation order."""
    if isinstance(net_config, str):
        net_config = json.loads(net_config)
    layers, params = [], []
    shape = (in_h, in_w, in_c)          # per-sample shape: 3-D (H,W,C) or 1-D (F,)

    def conv_out(n, k, s, pad):
        if pad == "SAME":
            o = -(-n // s)
            tot = max((o - 1) * s + k - n, 0)
            return o, tot // 2
        o = -(-(n - k + 1) // s)
        return o, 0

    for li, L in enumerate(net_config["middle_layer"]):
        kind = L["layer"]
        if kind == "conv":
            if len(shape) != 3:
                raise ValueError("conv after a dense layer")
            fh, fw, co = L["filter"]
            sh, sw = L.get("stride", [1, 1])
            pad = L.get("padding", "SAME")
            init = L.get("init", "norm")
            is_bias = False if "isBias" not in L else L["isBias"] != "False"    # :228
            bias_c = L.get("bias_constant", 0.1)
            std = L.get("stddev_norm", 0.1)
            if init not in ("zero", "norm", "xavier"):
                raise ValueError(init + " is not a legal init")                 # :76 assert False
            ih, iw, ci = shape
            oh, pt = conv_out(ih, fh, sh, pad)
            ow, pl = conv_out(iw, fw, sw, pad)
            wi = len(params)
            params.append(("conv%d/w" % li, (fh, fw, ci, co), init, std))
            bi = None
            if is_bias:
                bi = len(params)
                params.append(("conv%d/b" % li, (co,), "const", bias_c))
            layers.append(dict(kind="conv", fh=fh, fw=fw, ci=ci, co=co, sh=sh, sw=sw, pt=pt, pl=pl,
                               ih=ih, iw=iw, oh=oh, ow=ow, w=wi, b=bi))
            shape = (oh, ow, co)
        elif kind == "pool":
            kh, kw = L.get("kernel", [2, 2])
            sh, sw = L.get("stride", [2, 2])
            pad = L.get("padding", "SAME")
            ih, iw, c = shape
            oh, pt = conv_out(ih, kh, sh, pad)
            ow, pl = conv_out(iw, kw, sw, pad)
            layers.append(dict(kind="pool", kh=kh, kw=kw, sh=sh, sw=sw, pt=pt, pl=pl, ih=ih, iw=iw, c=c, oh=oh, ow=ow))
            shape = (oh, ow, c)
        elif kind == "active":
            fn = L.get("active_func", "sigmoid")                                 # :242 default sigmoid
            param = L.get("param", [0.2])
            if fn not in ("relu", "sigmoid", "leaky_relu"):
                raise ValueError(fn + " is not a legal activation")             # :152 assert False
            slope = 0.2 if (param is None or len(param) == 0) else float(param[0])
            layers.append(dict(kind="active", fn=fn, slope=slope, size=int(np.prod(shape))))
        elif kind == "connect":
            hid = L.get("hidden", 512)
            fin = int(np.prod(shape))
            wi = len(params)
            params.append(("connect%d/w" % li, (fin, hid), "norm", 0.1))
            params.append(("connect%d/b" % li, (hid,), "const", 0.1))
            layers.append(dict(kind="connect", fin=fin, fout=hid, w=wi, b=wi + 1))
            shape = (hid,)
        elif kind == "norm":
            eps = L.get("epsilon", 1e-3)
            c = shape[-1]
            wi = len(params)
            params.append(("norm%d/scale" % li, (c,), "const", 1.0))
            params.append(("norm%d/offset" % li, (c,), "const", 0.0))
            layers.append(dict(kind="norm", eps=float(eps), c=c, rows=int(np.prod(shape)) // c, scale=wi, offset=wi + 1))
        # unknown layer names are silently skipped by the reference's if/elif chain (:219-250)
    fin = int(np.prod(shape))
    wi = len(params)
    params.append(("out/w", (fin, classes), "norm", 0.1))
    params.append(("out/b", (classes,), "const", 0.1))
    layers.append(dict(kind="connect", fin=fin, fout=classes, w=wi, b=wi + 1))
    return layers, params


def init_params(params, rng, dtype=np.float32):
    """weight_variable / bias_variable (:57-87).  truncated_normal = resample beyond 2 sigma.
    ("xavier" divides stddev by shape[1]*shape[2]*shape[3], :72-73.)  TF's RNG stream is not
    reproducible; parity tests inject these same arrays into the CUDA side."""
    out = []
    for (_, shape, kind, arg) in params:
        if kind == "const":
            out.append(np.full(shape, arg, dtype))
        elif kind == "zero":
            out.append(np.zeros(shape, dtype))
        else:
            std = arg if kind == "norm" else arg * 1.0 / (shape[1] * shape[2] * shape[3])
            a = rng.standard_normal(shape)
            bad = np.abs(a) > 2
            while bad.any():
                a[bad] = rng.standard_normal(int(bad.sum()))
                bad = np.abs(a) > 2
            out.append((a * std).astype(dtype))
    return out


# ----------------------------------------------------------------------------- primitives
def _im2col(x, fh, fw, sh, sw, pt, pl, oh, ow):
    n, ih, iw, c = x.shape
    pb = max((oh - 1) * sh + fh - ih - pt, 0)
    pr = max((ow - 1) * sw + fw - iw - pl, 0)
    xp = np.pad(x, ((0, 0), (pt, pb), (pl, pr), (0, 0)))
    cols = np.empty((n, oh, ow, fh, fw, c), x.dtype)
    for i in range(fh):
        for j in range(fw):
            cols[:, :, :, i, j, :] = xp[:, i:i + (oh - 1) * sh + 1:sh, j:j + (ow - 1) * sw + 1:sw, :]
    return cols, xp.shape


def conv2d_fwd(x, w, L):
    cols, _ = _im2col(x, L["fh"], L["fw"], L["sh"], L["sw"], L["pt"], L["pl"], L["oh"], L["ow"])
    n = x.shape[0]
    y = cols.reshape(n * L["oh"] * L["ow"], -1) @ w.reshape(-1, L["co"])
    return y.reshape(n, L["oh"], L["ow"], L["co"]), cols


def conv2d_bwd(dy, x, w, L, cols):
    n = x.shape[0]
    fh, fw, sh, sw, pt, pl, oh, ow = (L[k] for k in ("fh", "fw", "sh", "sw", "pt", "pl", "oh", "ow"))
    dy2 = dy.reshape(n * oh * ow, L["co"])
    dw = (cols.reshape(n * oh * ow, -1).T @ dy2).reshape(w.shape)
    dcols = (dy2 @ w.reshape(-1, L["co"]).T).reshape(n, oh, ow, fh, fw, L["ci"])
    pb = max((oh - 1) * sh + fh - L["ih"] - pt, 0)
    pr = max((ow - 1) * sw + fw - L["iw"] - pl, 0)
    dxp = np.zeros((n, L["ih"] + pt + pb, L["iw"] + pl + pr, L["ci"]), x.dtype)
    for i in range(fh):
        for j in range(fw):
            dxp[:, i:i + (oh - 1) * sh + 1:sh, j:j + (ow - 1) * sw + 1:sw, :] += dcols[:, :, :, i, j, :]
    dx = dxp[:, pt:pt + L["ih"], pl:pl + L["iw"], :]
    return dx, dw


def maxpool_fwd(x, L):
    """tf.nn.max_pool; SAME padding never wins (-inf).  argmax = first maximum in row-major
    window order (TF CPU MaxPoolGrad uses strict '>' while scanning)."""
    n, ih, iw, c = x.shape
    kh, kw, sh, sw, pt, pl, oh, ow = (L[k] for k in ("kh", "kw", "sh", "sw", "pt", "pl", "oh", "ow"))
    y = np.full((n, oh, ow, c), -np.inf, x.dtype)
    arg = np.zeros((n, oh, ow, c), np.int32)
    for i in range(kh):
        for j in range(kw):
            ys = np.arange(oh) * sh - pt + i
            xs = np.arange(ow) * sw - pl + j
            vy = (ys >= 0) & (ys < ih)
            vx = (xs >= 0) & (xs < iw)
            v = np.full((n, oh, ow, c), -np.inf, x.dtype)
            sub = x[:, np.clip(ys, 0, ih - 1)][:, :, np.clip(xs, 0, iw - 1)]
            m = vy[:, None] & vx[None, :]
            v[:, m] = sub[:, m]
            better = v > y
            y = np.where(better, v, y)
            arg = np.where(better, i * kw + j, arg)
    return y, arg


def maxpool_bwd(dy, arg, L, n):
    kh, kw, sh, sw, pt, pl, oh, ow = (L[k] for k in ("kh", "kw", "sh", "sw", "pt", "pl", "oh", "ow"))
    dx = np.zeros((n, L["ih"], L["iw"], L["c"]), dy.dtype)
    for i in range(kh):
        for j in range(kw):
            sel = arg == (i * kw + j)
            ys = np.arange(oh) * sh - pt + i
            xs = np.arange(ow) * sw - pl + j
            for a, yy in enumerate(ys):
                if yy < 0 or yy >= L["ih"]:
                    continue
                for b, xx in enumerate(xs):
                    if xx < 0 or xx >= L["iw"]:
                        continue
                    dx[:, yy, xx, :] += np.where(sel[:, a, b, :], dy[:, a, b, :], 0)
    return dx


# ----------------------------------------------------------------------------- network
class Net:
    def __init__(self, net_config, in_h=28, in_w=28, in_c=1, classes=10, dtype=np.float32):
        self.layers, self.param_specs = parse_layers(net_config, in_h, in_w, in_c, classes)
        self.in_shape = (in_h, in_w, in_c)
        self.classes = classes
        self.dtype = dtype

    def param_count(self):
        return int(sum(np.prod(s[1]) for s in self.param_specs))

    def flatten(self, params):
        return np.concatenate([p.reshape(-1) for p in params]).astype(self.dtype)

    def unflatten(self, flat):
        out, o = [], 0
        for (_, shape, _, _) in self.param_specs:
            n = int(np.prod(shape))
            out.append(np.asarray(flat[o:o + n], self.dtype).reshape(shape))
            o += n
        return out

    def forward(self, x, params, keep=False):
        """x [N, H*W*C] -> logits [N, classes]  (reshape at :216, layers :219-250, output :262-264)."""
        dt = self.dtype
        n = x.shape[0]
        a = np.asarray(x, dt).reshape((n,) + self.in_shape)
        cache = []
        for L in self.layers:
            k = L["kind"]
            if k == "conv":
                y, cols = conv2d_fwd(a, params[L["w"]], L)
                if L["b"] is not None:
                    y = y + params[L["b"]]
                cache.append((a, cols))
                a = y.astype(dt)
            elif k == "pool":
                y, arg = maxpool_fwd(a, L)
                cache.append(arg)
                a = y
            elif k == "active":
                if L["fn"] == "relu":
                    y = np.maximum(a, 0)
                elif L["fn"] == "sigmoid":
                    y = (1.0 / (1.0 + np.exp(-a.astype(np.float64)))).astype(dt) if dt == np.float32 else 1.0 / (1.0 + np.exp(-a))
                else:
                    y = np.maximum(a, dt(L["slope"]) * a)
                cache.append((a, y))
                a = y
            elif k == "norm":
                # normalize() :155-165: batch statistics always, biased variance
                axes = tuple(range(a.ndim - 1))
                mean = a.mean(axis=axes, dtype=np.float64).astype(dt) if dt == np.float32 else a.mean(axis=axes)
                var = ((a - mean) ** 2).mean(axis=axes, dtype=np.float64).astype(dt) if dt == np.float32 else ((a - mean) ** 2).mean(axis=axes)
                inv = 1.0 / np.sqrt(var + dt(L["eps"]))
                xhat = (a - mean) * inv
                y = xhat * params[L["scale"]] + params[L["offset"]]
                cache.append((xhat, inv))
                a = y.astype(dt)
            elif k == "connect":
                a2 = a.reshape(n, -1)                      # NHWC flatten, (h,w,c) order (:175)
                y = a2 @ params[L["w"]] + params[L["b"]]
                cache.append((a2, a.shape))
                a = y.astype(dt)
        return (a, cache) if keep else a

    def loss_and_grad(self, logits, y_onehot, loss_name="entropy"):
        """loss_function :285-298 and d loss / d logits."""
        n = logits.shape[0]
        z = logits.astype(np.float64)
        y = np.asarray(y_onehot, np.float64)
        if loss_name == "entropy":
            m = z.max(axis=1, keepdims=True)
            lse = m + np.log(np.exp(z - m).sum(axis=1, keepdims=True))
            logp = z - lse
            loss = float((-(y * logp).sum(axis=1)).mean())
            dz = (np.exp(logp) - y) / n
        elif loss_name == "mse":
            loss = float(((y - z) ** 2).mean())
            dz = 2.0 * (z - y) / z.size
        else:
            raise ValueError("loss_function returns None for %r (reference crashes in minimize)" % loss_name)
        return loss, dz.astype(self.dtype)

    def backward(self, dz, params, cache):
        dt = self.dtype
        grads = [None] * len(params)
        d = dz
        for L, c in zip(reversed(self.layers), reversed(cache)):
            k = L["kind"]
            if k == "connect":
                a2, shp = c
                grads[L["w"]] = (a2.T @ d).astype(dt)
                grads[L["b"]] = d.sum(axis=0).astype(dt)
                d = (d @ params[L["w"]].T).reshape(shp).astype(dt)
            elif k == "active":
                a, y = c
                if L["fn"] == "relu":
                    d = np.where(y > 0, d, 0).astype(dt)
                elif L["fn"] == "sigmoid":
                    d = (d * y * (1 - y)).astype(dt)
                else:
                    s = dt(L["slope"])
                    d = np.where(a >= s * a, d, s * d).astype(dt)
            elif k == "pool":
                d = maxpool_bwd(d, c, L, d.shape[0])
            elif k == "norm":
                xhat, inv = c
                axes = tuple(range(d.ndim - 1))
                m = float(np.prod([d.shape[i] for i in axes]))
                grads[L["offset"]] = d.sum(axis=axes).astype(dt)
                grads[L["scale"]] = (d * xhat).sum(axis=axes).astype(dt)
                dxhat = d * params[L["scale"]]
                d = (inv / m * (m * dxhat - dxhat.sum(axis=axes) - xhat * (dxhat * xhat).sum(axis=axes))).astype(dt)
            elif k == "conv":
                a, cols = c
                if L["b"] is not None:
                    grads[L["b"]] = d.sum(axis=(0, 1, 2)).astype(dt)
                d, dw = conv2d_bwd(d, a, params[L["w"]], L, cols)
                grads[L["w"]] = dw.astype(dt)
                d = d.astype(dt)
        return grads

    def accuracy(self, logits, y_onehot):
        """:381-382."""
        return float((logits.argmax(1) == np.asarray(y_onehot).argmax(1)).mean())


# ----------------------------------------------------------------------------- optimizers
class Adagrad:
    """tf.train.AdagradOptimizer(lr) -- the optimizer that actually runs (:372-373, lr 1e-4):
    accumulator starts at 0.1; acc += g*g; w -= lr * g / sqrt(acc)   (no epsilon)."""

    def __init__(self, n, lr=1e-4, init_acc=0.1, dtype=np.float32):
        self.lr, self.acc, self.dt = dtype(lr), np.full(n, init_acc, dtype), dtype

    def apply(self, w, g):
        self.acc += g * g
        w -= self.lr * g / np.sqrt(self.acc)


class SGD:
    """tf.train.GradientDescentOptimizer (optimizer_function :309): w -= lr * g."""

    def __init__(self, n, lr, dtype=np.float32):
        self.lr = dtype(lr)

    def apply(self, w, g):
        w -= self.lr * g


class Adam:
    """tf.train.AdamOptimizer TF-1.x (options.py:28): lr_t = lr*sqrt(1-b2^t)/(1-b1^t);
    m += (g-m)(1-b1); v += (g*g-v)(1-b2); w -= lr_t*m/(sqrt(v)+eps)."""

    def __init__(self, n, lr=1e-3, b1=0.9, b2=0.999, eps=1e-8, dtype=np.float32):
        self.lr, self.b1, self.b2, self.eps, self.dt = lr, b1, b2, eps, dtype
        self.m, self.v, self.t = np.zeros(n, dtype), np.zeros(n, dtype), 0

    def apply(self, w, g):
        self.t += 1
        lr_t = self.dt(self.lr * math.sqrt(1 - self.b2 ** self.t) / (1 - self.b1 ** self.t))
        # TF's ApplyAdam evaluates (1 - beta) in the variable dtype: float32(1) - float32(0.999)
        self.m += (g - self.m) * (self.dt(1) - self.dt(self.b1))
        self.v += (g * g - self.v) * (self.dt(1) - self.dt(self.b2))
        w -= lr_t * self.m / (np.sqrt(self.v) + self.dt(self.eps))


def train_step(net, flat_w, x, y, opt, loss_name="entropy"):
    """One sess.run(train_op) (:413): returns (loss, logits); updates flat_w and opt in place."""
    params = net.unflatten(flat_w)
    logits, cache = net.forward(x, params, keep=True)
    loss, dz = net.loss_and_grad(logits, y, loss_name)
    grads = net.backward(dz, params, cache)
    g = net.flatten(grads)
    opt.apply(flat_w, g)
    return loss, logits, g


# ----------------------------------------------------------------------------- data
class DataSet:
    """DataSet :466-569 (the shuffle RNG is numpy's global generator, unseeded in the reference;
    pass a RandomState to make it testable)."""

    def __init__(self, images, labels, rs=None):
        assert images.shape[0] == labels.shape[0]
        self._num_examples = images.shape[0]
        images = images.reshape(images.shape[0], -1)
        images = images.astype(np.float32)
        images = np.multiply(images, 1.0 / 255.0)
        self._images, self._labels = images, labels
        self._epochs_completed, self._index_in_epoch = 0, 0
        self._rs = rs if rs is not None else np.random

    def next_batch(self, batch_size, shuffle=True):
        start = self._index_in_epoch
        if self._epochs_completed == 0 and start == 0 and shuffle:
            perm0 = np.arange(self._num_examples)
            self._rs.shuffle(perm0)
            self._images, self._labels = self._images[perm0], self._labels[perm0]
        if start + batch_size > self._num_examples:
            self._epochs_completed += 1
            rest = self._num_examples - start
            ir, lr_ = self._images[start:self._num_examples], self._labels[start:self._num_examples]
            if shuffle:
                perm = np.arange(self._num_examples)
                self._rs.shuffle(perm)
                self._images, self._labels = self._images[perm], self._labels[perm]
            start = 0
            self._index_in_epoch = batch_size - rest
            end = self._index_in_epoch
            return np.concatenate((ir, self._images[start:end]), 0), np.concatenate((lr_, self._labels[start:end]), 0)
        self._index_in_epoch += batch_size
        end = self._index_in_epoch
        return self._images[start:end], self._labels[start:end]


LENET = {"middle_layer": [
    {"layer": "conv", "filter": [5, 5, 32], "isBias": "True"}, {"layer": "active", "active_func": "relu"}, {"layer": "pool"},
    {"layer": "conv", "filter": [5, 5, 64], "isBias": "True"}, {"layer": "active", "active_func": "relu"}, {"layer": "pool"},
    {"layer": "connect", "hidden": 1024}, {"layer": "active", "active_func": "relu"}]}

DEEPNET = {"middle_layer": [
    {"layer": "conv", "filter": [3, 3, 32], "isBias": "True"}, {"layer": "active", "active_func": "relu"},
    {"layer": "conv", "filter": [3, 3, 32], "isBias": "True"}, {"layer": "active", "active_func": "relu"}, {"layer": "pool"},
    {"layer": "conv", "filter": [3, 3, 64], "isBias": "True"}, {"layer": "active", "active_func": "relu"},
    {"layer": "conv", "filter": [3, 3, 64], "isBias": "True"}, {"layer": "active", "active_func": "relu"}, {"layer": "pool"},
    {"layer": "connect", "hidden": 512}, {"layer": "active", "active_func": "relu"},
    {"layer": "connect", "hidden": 256}, {"layer": "active", "active_func": "relu"}]}

API_EXAMPLE = {"middle_layer": [  # API.md:306-332
    {"layer": "conv", "filter": [2, 2, 10]}, {"layer": "conv", "filter": [2, 2, 20]}, {"layer": "pool"},
    {"layer": "norm"}, {"layer": "active"}, {"layer": "connect"}, {"layer": "connect"}]}
