"""Data-parallel training.  CPU: the host-side recipe of net.DataParallel (broadcast,
dLoss scaled by 1/W, all-reduce SUM of the flat gradient buffer, identical optimizer step) on a
world_size-2 gloo group equals the single-process step at the global batch.  GPU (needs >= 2 GPUs):
tests/dp_check.py under torchrun with NCCL."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT


class _CpuNet:
    """Stand-in with the interface DataParallel uses (params, grads, backward, apply): a linear
    softmax classifier in torch-CPU.  Only the distributed plumbing is under test here."""

    def __init__(self, seed):
        g = torch.Generator().manual_seed(seed)
        self.params = torch.randn(784 * 10 + 10, generator=g) * 0.1
        self.grads = torch.zeros_like(self.params)

    def backward(self, x, y, loss_name="entropy", grad_scale=1.0):
        w = self.params[:7840].reshape(784, 10).clone().requires_grad_(True)
        b = self.params[7840:].clone().requires_grad_(True)
        loss = (-(y * torch.log_softmax(x @ w + b, 1)).sum(1)).mean()
        (loss * grad_scale).backward()
        self.grads.copy_(torch.cat([w.grad.reshape(-1), b.grad]))
        return loss.detach().reshape(1), torch.zeros(1, dtype=torch.int32)

    def apply(self, opt="sgd", lr=0.1):
        self.params -= lr * self.grads


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from cloud_server_b200.net import DataParallel
    g = torch.Generator().manual_seed(3)
    x = torch.rand((2, 32, 784), generator=g)
    y = torch.nn.functional.one_hot(torch.randint(0, 10, (2, 32), generator=g), 10).float()
    net = _CpuNet(seed=10 + rank)                 # ranks start different; DataParallel must broadcast rank 0's
    dp = DataParallel(net)
    n = 32 // world
    for s in range(2):
        dp.train_step(x[s, rank * n:(rank + 1) * n], y[s, rank * n:(rank + 1) * n], "entropy", "sgd", 0.5)
    if rank == 0:
        single = _CpuNet(seed=10)
        for s in range(2):
            single.backward(x[s], y[s])
            single.apply("sgd", 0.5)
        out.put((net.params.numpy().copy(), single.params.numpy().copy()))
    gathered = [torch.zeros_like(net.params) for _ in range(world)]
    dist.all_gather(gathered, net.params)
    assert all(torch.equal(gathered[0], t) for t in gathered)
    dist.barrier()
    dist.destroy_process_group()


def test_data_parallel_recipe_gloo_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29400 + os.getpid() % 500
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    a, b = q.get(timeout=120)
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert np.abs(a - b).max() < 1e-5 * max(1.0, np.abs(b).max())


@pytest.mark.gpu
def test_data_parallel_nccl_matches_single_device():
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs (run through `gpurun --gpus 2`)")
    w = 2 if n < 4 else 4
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(w), "--master-addr", "127.0.0.1",
           "--master-port", "29517", os.path.join(ROOT, "tests", "dp_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
