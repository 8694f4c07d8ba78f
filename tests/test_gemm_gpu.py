"""GPU tests of the tcgen05 TF32 GEMM (nc_matmul) against float64 on the host.  TF32 keeps 10
mantissa bits per operand: tolerance 2e-3 of the output scale (north_star allows 1e-3 relative on
logits / loss / weights; the net-level tests hold that)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _check(ctx, m, k, n, ta, tb, bias, seed, precision=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    a = torch.randn((k, m) if ta else (m, k), device="cuda", generator=g)
    b = torch.randn((n, k) if tb else (k, n), device="cuda", generator=g)
    bv = torch.randn((n,), device="cuda", generator=g) if bias else None
    got = ctx.matmul(a, b, bv, ta, tb, precision).double().cpu()
    A = a.double().cpu().T if ta else a.double().cpu()
    B = b.double().cpu().T if tb else b.double().cpu()
    want = A @ B + (bv.double().cpu() if bias else 0)
    err = float((got - want).abs().max() / want.abs().max())
    # 3x: hi = the operand as the tensor core truncates it, lo = the exact remainder (its own truncation and the dropped
    # lo.lo term are one-sided, ~2^-21 each); 1x: truncated TF32 operands
    assert err < (5e-6 if precision == 0 else 4e-3), (m, k, n, ta, tb, precision, err)
    return err


@pytest.mark.parametrize("ta,tb", [(False, True), (False, False), (True, False), (True, True)])
def test_matmul_all_operand_layouts(ctx, ta, tb):
    # one full tile, several k blocks; both precisions
    _check(ctx, 128, 128, 128, ta, tb, False, 1)
    _check(ctx, 128, 128, 128, ta, tb, False, 1, precision=1)
    _check(ctx, 200, 104, 72, ta, tb, True, 2, precision=1)
    # tails in every dimension, bias, N <= 64 path
    _check(ctx, 200, 104, 72, ta, tb, True, 2)
    _check(ctx, 60, 40, 64, ta, tb, True, 3)
    _check(ctx, 300, 36, 24, ta, tb, False, 4)


def test_matmul_lenet_dense_shapes_and_split_k(ctx):
    # fc1 forward (x.w + b), dgrad (dy.w^T), wgrad (x^T.dy) at batch 256: split-K paths
    _check(ctx, 256, 3136, 1024, False, False, True, 5)      # forward: b is [k][n]
    _check(ctx, 256, 1024, 3136, False, True, False, 6)      # dgrad: b given as [n][k] = W[in][out]
    _check(ctx, 3136, 256, 1024, True, False, False, 7)      # wgrad: a given as [k][m] = X[batch][in]
    _check(ctx, 2048, 1024, 16, False, False, True, 8)       # narrow output, N = 16 (< one tile)
    # exact small integers survive TF32: result is exact
    a = torch.randint(-8, 9, (128, 256), device="cuda").float()
    b = torch.randint(-8, 9, (256, 128), device="cuda").float()
    assert torch.equal(ctx.matmul(a, b), a @ b)


def test_matmul_rejects_unaligned(ctx):
    from cloud_server_b200._lib import NcError
    a = torch.zeros((16, 30), device="cuda")       # row pitch 30 floats: not 16-byte aligned
    b = torch.zeros((30, 16), device="cuda")
    with pytest.raises(NcError):
        ctx.matmul(a, b)
