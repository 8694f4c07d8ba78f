"""Data-parallel parity check, run under torchrun on W GPUs (one rank per GPU):
W ranks x (B/W) rows, NCCL all-reduce of the flat gradient buffer  ==  one device at batch B
(post-step weights within 1e-3 relative of the update).  Exit code 0 = pass.
  torchrun --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tests/dp_check.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist


def main():
    from cloud_server_b200.context import Context
    from cloud_server_b200.net import Net, DataParallel
    from cloud_server_b200.configs import LENET
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx = Context(local)
    B = 64 * world
    g = torch.Generator(device="cuda").manual_seed(7)            # same data on every rank
    x = torch.rand((3, B, 784), device="cuda", generator=g)
    y = torch.nn.functional.one_hot(torch.randint(0, 10, (3, B), device="cuda", generator=g), 10).float()
    ok = True
    for opt, lr in (("compat", 1e-4), ("sgd", 0.05)):
        net = Net(LENET, ctx, seed=100 + rank)                   # different init per rank: broadcast must fix it
        dp = DataParallel(net)
        w0 = net.get_params().copy()
        single = Net(LENET, ctx, seed=5)
        single.set_params(w0)
        n = B // world
        for s in range(3):
            loss, _ = dp.train_step(x[s, rank * n:(rank + 1) * n], y[s, rank * n:(rank + 1) * n], "entropy", opt, lr)
            single.train_step(x[s], y[s], "entropy", opt, lr)
        a, b = net.get_params(), single.get_params()
        upd = np.abs(b - w0).max()
        err = np.abs(a - b).max() / upd
        # every rank must hold identical weights
        t = net.params.clone()
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        same = bool(torch.equal(t, net.params))
        if rank == 0:
            print("dp_check world=%d opt=%s: |dp - single| / |update| = %.3e, ranks identical = %s" % (world, opt, err, same), flush=True)
        ok = ok and err < 1e-3 and same
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
