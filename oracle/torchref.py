"""torch-CPU restatement of the CNN half (TEST INFRASTRUCTURE / CPU BASELINE ONLY -- never imported by the product).

Second, independent derivation of what oracle/tfref.py restates from
apps/construction/util/construct_distribute.py:57-311 (+ AdagradOptimizer(1e-4) :372-373): the same graph through
torch's own conv2d / max_pool2d / autograd on the host cores, with TF's asymmetric SAME padding written out
explicitly (extra row / column on the bottom / right).  Two uses:
  * tests/test_tfref.py checks tfref's forward and hand-written backward against it in float64;
  * bench.py's `cpu_baseline` and `--impl reference` legs time `TorchStep` (float32, all host threads, MKL/oneDNN)
    as the CPU arm of the train-step benchmark: TensorFlow 1.x itself cannot be installed here, so this is the
    closest multi-threaded CPU implementation of the same step (`kind: "port"`).
PARITY UNPINNED against TensorFlow, like tfref.
"""
import numpy as np
import torch
import torch.nn.functional as F

from . import tfref


def torch_forward(net, x, params):
    """net: tfref.Net (layer specs); x [N, H*W*C]; params: list of tensors in tfref's variable order (HWIO / [in,out])."""
    n = x.shape[0]
    a = x.reshape((n,) + net.in_shape).permute(0, 3, 1, 2)
    for L in net.layers:
        k = L["kind"]
        if k == "conv":
            w = params[L["w"]].permute(3, 2, 0, 1)     # HWIO -> OIHW
            pb = max((L["oh"] - 1) * L["sh"] + L["fh"] - L["ih"] - L["pt"], 0)
            pr = max((L["ow"] - 1) * L["sw"] + L["fw"] - L["iw"] - L["pl"], 0)
            a = F.pad(a, (L["pl"], pr, L["pt"], pb))
            a = F.conv2d(a, w, params[L["b"]] if L["b"] is not None else None, stride=(L["sh"], L["sw"]))
            a = a[:, :, :L["oh"], :L["ow"]]
        elif k == "pool":
            pb = max((L["oh"] - 1) * L["sh"] + L["kh"] - L["ih"] - L["pt"], 0)
            pr = max((L["ow"] - 1) * L["sw"] + L["kw"] - L["iw"] - L["pl"], 0)
            a = F.pad(a, (L["pl"], pr, L["pt"], pb), value=float("-inf"))
            a = F.max_pool2d(a, (L["kh"], L["kw"]), (L["sh"], L["sw"]))[:, :, :L["oh"], :L["ow"]]
        elif k == "active":
            a = torch.relu(a) if L["fn"] == "relu" else torch.sigmoid(a) if L["fn"] == "sigmoid" else torch.maximum(a, L["slope"] * a)
        elif k == "norm":
            axes = [0, 2, 3] if a.dim() == 4 else [0]
            mean = a.mean(dim=axes, keepdim=True)
            var = ((a - mean) ** 2).mean(dim=axes, keepdim=True)
            shp = (1, -1, 1, 1) if a.dim() == 4 else (1, -1)
            a = (a - mean) / torch.sqrt(var + L["eps"]) * params[L["scale"]].reshape(shp) + params[L["offset"]].reshape(shp)
        elif k == "connect":
            if a.dim() == 4:
                a = a.permute(0, 2, 3, 1).reshape(n, -1)
            a = a @ params[L["w"]] + params[L["b"]]
    return a


class TorchStep:
    """One sess.run(train_op) (construct_distribute.py:413) on the host cores: forward, softmax-CE (or mse) batch mean,
    autograd backward, TF-1.x Adagrad (accumulator 0.1, no epsilon) or SGD on every variable."""

    def __init__(self, net_config, in_h=28, in_w=28, in_c=1, classes=10, flat_params=None, seed=0, threads=None,
                 opt="adagrad", lr=1e-4, loss_name="entropy"):
        if threads:
            torch.set_num_threads(int(threads))
        self.net = tfref.Net(net_config, in_h, in_w, in_c, classes)
        if flat_params is None:
            flat_params = self.net.flatten(tfref.init_params(self.net.param_specs, np.random.default_rng(seed)))
        self.params = [torch.tensor(p, dtype=torch.float32, requires_grad=True) for p in self.net.unflatten(flat_params)]
        self.acc = [torch.full_like(p, 0.1) for p in self.params]
        self.opt, self.lr, self.loss_name = opt, float(lr), loss_name

    def flat(self):
        return np.concatenate([p.detach().numpy().reshape(-1) for p in self.params])

    def step(self, x, y):
        """x [N, features] float32 tensor, y [N, classes] one-hot float32 tensor -> loss (python float)"""
        logits = torch_forward(self.net, x, self.params)
        if self.loss_name == "entropy":
            loss = (-(y * torch.log_softmax(logits, 1)).sum(1)).mean()
        else:
            loss = ((y - logits) ** 2).mean()
        grads = torch.autograd.grad(loss, self.params)
        with torch.no_grad():
            for p, g, a in zip(self.params, grads, self.acc):
                if self.opt == "adagrad":
                    a.addcmul_(g, g)
                    p.addcdiv_(g, a.sqrt(), value=-self.lr)
                else:
                    p.add_(g, alpha=-self.lr)
        return float(loss.detach())
