"""GPU test of nc_dense_forward_noalloc, the one entry point for callers that may not allocate
device memory at request time (a server whose pool owns all of HBM, a latency-critical thread,
an audited tenant).  It must give what the ordinary path gives: the reference here is the existing
entry point nc_net_forward on a net that consists of the implicit output layer only (so its
forward IS one dense layer, and at this shape it runs the split-K kernel with the library's
malloc'ed workspace).  Two calls in a row, the second with other data and a smaller batch, show
that what the first call left in the shared scratch array is never observed.

Collected last (file name), so a run that stops at the first failure has run everything else."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

K, N = 28 * 28 * 4, 1024          # 3136 -> 1024: the shape of LeNet's first dense layer


def _call(ctx, x, w, b, y):
    from cloud_server_b200 import _lib
    rc = _lib.lib().nc_dense_forward_noalloc(ctx.handle, C.c_void_p(x.data_ptr()), C.c_void_p(w.data_ptr()),
                                             C.c_void_p(b.data_ptr()), x.shape[0], K, N, C.c_void_p(y.data_ptr()),
                                             ctx._stream())
    _lib.check(ctx.handle, rc)
    return y


def test_noalloc_dense_matches_net_forward_twice(ctx):
    from cloud_server_b200.net import Net
    net = Net({"middle_layer": []}, ctx, 28, 28, 4, classes=N, seed=11)
    (w_off, w_size, w_shape), (b_off, b_size, _) = net.tensors
    assert w_shape == (K, N)
    w = net.params[w_off:w_off + w_size]
    b = net.params[b_off:b_off + b_size]
    g = torch.Generator(device="cuda").manual_seed(3)
    outs = [torch.empty((64, N), dtype=torch.float32, device="cuda") for _ in range(2)]   # caller-owned outputs, allocated up front
    # inputs and the reference results of both calls first: torch's caching allocator and nc_net_forward (malloc'ed workspace)
    # may take device memory, and neither is what this test is about
    xs = [torch.randn((m, K), device="cuda", generator=g) for m in (64, 37)]          # two calls in a row; the second is smaller
    wants = [net.forward(x).clone() for x in xs]                                        # existing entry point
    torch.cuda.synchronize()
    _call(ctx, xs[0], w, b, outs[0])                     # new entry point (fixed scratch); the first call on a device does the
    torch.cuda.synchronize()                             # one-time set-up (and CUDA loads the kernels lazily)
    free_after_first = torch.cuda.mem_get_info()[0]
    _call(ctx, xs[1], w, b, outs[1])
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free_after_first, "a later call changed free device memory"
    for x, want, out in zip(xs, wants, outs):
        m = x.shape[0]
        got = out[:m]
        err = float((got - want).abs().max() / want.abs().max())
        assert err < 2e-5, (m, err)                      # tolerance of the fp32 SIMT path (tests/test_net_gpu.py)
    # against float64 on the host as well
    ref = x.double().cpu() @ w.reshape(K, N).double().cpu() + b.double().cpu()
    assert float((got.double().cpu() - ref).abs().max() / ref.abs().max()) < 2e-5


def test_noalloc_dense_rejects_oversize_and_bad_arguments(ctx):
    from cloud_server_b200 import _lib
    x = torch.zeros((2048, 16), device="cuda")
    w = torch.zeros((16, 1024), device="cuda")
    y = torch.zeros((2048, 1024), device="cuda")
    L = _lib.lib()
    rc = L.nc_dense_forward_noalloc(ctx.handle, C.c_void_p(x.data_ptr()), C.c_void_p(w.data_ptr()), None,
                                    2048, 16, 1024, C.c_void_p(y.data_ptr()), ctx._stream())
    assert rc == _lib.NC_EINVAL and b"exceed the fixed scratch" in L.nc_last_error(ctx.handle)
    rc = L.nc_dense_forward_noalloc(ctx.handle, None, C.c_void_p(w.data_ptr()), None, 4, 16, 1024,
                                    C.c_void_p(y.data_ptr()), ctx._stream())
    assert rc == _lib.NC_EINVAL
    # largest legal size, no bias
    x1 = torch.randn((1024, 16), device="cuda")
    w1 = torch.randn((16, 1024), device="cuda")
    rc = L.nc_dense_forward_noalloc(ctx.handle, C.c_void_p(x1.data_ptr()), C.c_void_p(w1.data_ptr()), None,
                                    1024, 16, 1024, C.c_void_p(y.data_ptr()), ctx._stream())
    assert rc == 0
    assert torch.allclose(y[:1024], x1 @ w1, rtol=1e-4, atol=1e-4)
