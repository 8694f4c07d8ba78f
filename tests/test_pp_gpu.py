"""GPU parity tests of the preprocess half: CUDA kernels (through the C ABI) against the oracle
restatement, the committed golden vectors, and -- when importable on the box -- the live cv2 /
Pillow calls.  Bar: bit-exact (uint8, float32 outputs are exact products of exact bytes)."""
import numpy as np
import pytest
import torch

from conftest import golden
from oracle import pp_restate as R
from oracle import cvref

pytestmark = pytest.mark.gpu


def synth_upload(seed, h, w, natural):
    rng = np.random.default_rng(seed)
    a = rng.integers(0, 256, (h, w, 3), np.uint8)
    if natural:
        import cv2
        a = cv2.GaussianBlur(a, (0, 0), 3)
    return a


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


# ------------------------------------------------------------------ train-side chain (P2-P4)
def test_train_chain_golden(ctx):
    g = golden("train_chain.npz")
    for seed, h, w, nat in g["cases"]:
        if nat and not cvref.have_cv2():
            continue
        img = synth_upload(int(seed), int(h), int(w), bool(nat))
        u8, f32 = ctx.resize_gray_norm(dev(img[None]))
        assert np.array_equal(u8[0].cpu().numpy(), g[f"u8_{seed}"]), (seed, h, w)
        assert np.array_equal(f32[0].cpu().numpy(), g[f"f32_{seed}"])


def test_train_chain_random_shapes_vs_oracle(ctx):
    rng = np.random.default_rng(21)
    for it in range(40):
        h, w = int(rng.integers(29, 420)), int(rng.integers(29, 420))
        n = int(rng.integers(1, 4))
        imgs = rng.integers(0, 256, (n, h, w, 3), np.uint8)
        u8, f32 = ctx.resize_gray_norm(dev(imgs))
        u8, f32 = u8.cpu().numpy(), f32.cpu().numpy()
        for i in range(n):
            e8, ef = R.train_chain(imgs[i])
            assert np.array_equal(u8[i], e8), (h, w, i)
            assert np.array_equal(f32[i], ef)
            if cvref.have_cv2():
                c8, cf = cvref.train_chain(imgs[i])
                assert np.array_equal(u8[i], c8), ("cv2", h, w)


def test_train_chain_gray_single_channel_and_identity(ctx):
    rng = np.random.default_rng(22)
    for h, w in [(28, 28), (64, 40), (300, 301)]:
        g = rng.integers(0, 256, (2, h, w), np.uint8)
        u8, f32 = ctx.resize_gray_norm(dev(g))
        for i in range(2):
            e = R.cv_resize_cubic_u8(g[i], 28, 28)
            assert np.array_equal(u8[i].cpu().numpy(), e), (h, w)
            assert np.array_equal(f32[i].cpu().numpy(), R.dataset_normalise(e[None])[0])
    # 28x28x3: OpenCV copies, then gray
    c = rng.integers(0, 256, (3, 28, 28, 3), np.uint8)
    u8, _ = ctx.resize_gray_norm(dev(c))
    for i in range(3):
        assert np.array_equal(u8[i].cpu().numpy(), R.cv_bgr2gray(c[i]))


def test_train_chain_bundled_pictures(ctx):
    """Config 1: the 99 bundled 28x28 test pictures, decoded as cv2.imread(p) does (3 equal
    channels): resize is the identity and gray of (v,v,v) is v."""
    p = golden("pics.npz")
    bgr = np.repeat(p["gray_cv"][..., None], 3, axis=3)
    u8, f32 = ctx.resize_gray_norm(dev(bgr))
    assert np.array_equal(u8.cpu().numpy(), p["gray_cv"])
    assert np.array_equal(f32.cpu().numpy(), R.dataset_normalise(p["gray_cv"]))


def test_train_chain_full_size_batch_properties(ctx):
    """BASELINE config 2 shape (1024x768x3), a batch larger than the oracle checks one by one:
    (1) 3 pictures spot-checked bit-exact, (2) batch result independent of batch position and
    stride, (3) constant pictures map to the constant, (4) empty batch is a no-op."""
    rng = np.random.default_rng(23)
    n = 64
    imgs = torch.from_numpy(rng.integers(0, 256, (n, 768, 1024, 3), np.uint8)).cuda()
    u8, f32 = ctx.resize_gray_norm(imgs)
    for i in (0, 31, 63):
        e8, ef = R.train_chain(imgs[i].cpu().numpy())
        assert np.array_equal(u8[i].cpu().numpy(), e8)
        assert np.array_equal(f32[i].cpu().numpy(), ef)
    perm = torch.randperm(n, device="cuda")
    u8p, _ = ctx.resize_gray_norm(imgs[perm].contiguous())
    assert torch.equal(u8p, u8[perm])
    # strided view (every second picture) goes through img_stride
    u8s, _ = ctx.resize_gray_norm(imgs[::2])
    assert torch.equal(u8s, u8[::2])
    const = torch.full((2, 768, 1024, 3), 77, dtype=torch.uint8, device="cuda")
    u8c, f32c = ctx.resize_gray_norm(const)
    assert int(u8c.min()) == 77 and int(u8c.max()) == 77
    assert torch.equal(f32.to(torch.float64), (u8.reshape(n, 784).to(torch.float32) * np.float32(1 / 255.0)).to(torch.float64))
    e8, ef = ctx.resize_gray_norm(imgs[:0])
    assert e8.shape[0] == 0 and ef.shape[0] == 0


def test_train_chain_both_kernels_agree(ctx, monkeypatch):
    """Row-streaming (TMA) kernel and tap-gather kernel: identical bytes; both equal the oracle."""
    rng = np.random.default_rng(25)
    for (h, w) in [(32, 32), (29, 48), (100, 160), (768, 1024), (1080, 1920), (480, 640), (5, 1776)]:
        imgs = dev(rng.integers(0, 256, (3, h, w, 3), np.uint8))
        monkeypatch.delenv("NC_TRAIN_NO_ROWS", raising=False)
        a8, af = ctx.resize_gray_norm(imgs)
        monkeypatch.setenv("NC_TRAIN_NO_ROWS", "1")
        b8, bf = ctx.resize_gray_norm(imgs)
        monkeypatch.delenv("NC_TRAIN_NO_ROWS", raising=False)
        assert torch.equal(a8, b8) and torch.equal(af, bf), (h, w)
        e8, ef = R.train_chain(imgs[2].cpu().numpy())
        assert np.array_equal(a8[2].cpu().numpy(), e8), (h, w)


def test_train_chain_host_buffers(ctx):
    rng = np.random.default_rng(24)
    imgs = rng.integers(0, 256, (5, 200, 304, 3), np.uint8)      # 16-byte rows: row-streaming kernel over PCIe
    want = np.stack([R.train_chain(i)[0] for i in imgs])
    pinned = torch.from_numpy(imgs).pin_memory()
    u8, _ = ctx.resize_gray_norm(pinned)
    torch.cuda.synchronize()
    assert np.array_equal(u8.cpu().numpy(), want)
    pageable = torch.from_numpy(imgs.copy())
    u8b, _ = ctx.resize_gray_norm(pageable)
    torch.cuda.synchronize()
    assert np.array_equal(u8b.cpu().numpy(), want)
    odd = rng.integers(0, 256, (3, 77, 301, 3), np.uint8)        # unaligned rows: tap-gather kernel over PCIe
    u8c, _ = ctx.resize_gray_norm(torch.from_numpy(odd).pin_memory())
    torch.cuda.synchronize()
    assert np.array_equal(u8c.cpu().numpy(), np.stack([R.train_chain(i)[0] for i in odd]))


def test_train_chain_bad_arguments(ctx):
    with pytest.raises(ValueError):
        ctx.resize_gray_norm(torch.zeros((1, 30, 30, 2), dtype=torch.uint8, device="cuda"))
    with pytest.raises(ValueError):
        ctx.resize_gray_norm(torch.zeros((1, 30, 30, 3), dtype=torch.float32, device="cuda"))


def test_train_chain_ipp_arithmetic_mode1(ctx):
    """P2 mode 1 (the wheel's default arithmetic, Intel IPP): the kernel equals the oracle's float model bit for bit, the
    committed golden outputs of the IPP-enabled wheel to >= 99.9 % (measured 100 %), and -- on a box whose cv2 has IPP --
    the live call; pageable host input and the colour output take the same mode."""
    g = golden("train_chain_ipp.npz")
    same = tot = 0
    for seed, h, w, nat in g["cases"]:
        if nat and not cvref.have_cv2():
            continue
        img = synth_upload(int(seed), int(h), int(w), bool(nat))
        u8, f32 = ctx.resize_gray_norm(dev(img[None]), mode=1)
        e8, ef = R.train_chain(img, mode=1)
        assert np.array_equal(u8[0].cpu().numpy(), e8) and np.array_equal(f32[0].cpu().numpy(), ef), (seed, h, w)
        col = ctx.resize_cubic(dev(img[None]), mode=1)[0].cpu().numpy()
        assert np.array_equal(col, R.cv_resize_cubic_ipp(img))
        same += int((col == g[f"bgr_{seed}"]).sum())
        tot += col.size
    assert same >= 0.999 * tot, (same, tot)
    rng = np.random.default_rng(31)
    for it in range(12):
        h, w = int(rng.integers(29, 500)), int(rng.integers(29, 500))
        imgs = rng.integers(0, 256, (2, h, w, 3), np.uint8)
        u8, _ = ctx.resize_gray_norm(dev(imgs), mode=1)
        for i in range(2):
            assert np.array_equal(u8[i].cpu().numpy(), R.train_chain(imgs[i], mode=1)[0]), (h, w)
    gray = rng.integers(0, 256, (2, 90, 70), np.uint8)
    u8, _ = ctx.resize_gray_norm(dev(gray), mode=1)
    assert np.array_equal(u8[0].cpu().numpy(), R.cv_resize_cubic_ipp(gray[0]))
    host = torch.from_numpy(imgs)                                  # pageable host memory: staged, same arithmetic
    u8h, _ = ctx.resize_gray_norm(host, mode=1)
    assert torch.equal(u8h, u8.new_tensor(np.stack([R.train_chain(imgs[i], mode=1)[0] for i in range(2)])))
    if cvref.have_cv2():
        import cv2
        try:
            cv2.ipp.setUseIPP(True)
            if cv2.ipp.useIPP():
                same = tot = 0
                for h, w in ((768, 1024), (480, 640), (255, 257)):
                    img = rng.integers(0, 256, (h, w, 3), np.uint8)
                    want = cv2.resize(img, (28, 28), interpolation=cv2.INTER_CUBIC)
                    got = ctx.resize_cubic(dev(img[None]), mode=1)[0].cpu().numpy()
                    same += int((got == want).sum())
                    tot += want.size
                assert same >= 0.999 * tot, (same, tot)
        finally:
            cv2.ipp.setUseIPP(False)
    with pytest.raises(ValueError):
        ctx.resize_gray_norm(dev(imgs), mode=2)


# ------------------------------------------------------------------ predict-side chain (P5)
def test_predict_chain_bundled_pictures(ctx):
    p = golden("pics.npz")
    g = golden("predict_chain.npz")
    src = dev(p["gray_pil"])
    out, status, small = ctx.predict_chain(src, compat=False, want_small20=True)
    assert np.array_equal(small.cpu().numpy(), g["small20"])
    assert np.array_equal(out.cpu().numpy(), g["out_intended"])
    out_c, status_c, _ = ctx.predict_chain(src, compat=True)
    assert np.array_equal(status_c.cpu().numpy(), g["compat_indexerror"])
    ok = g["compat_indexerror"] == 0
    assert ok.sum() == 17
    assert np.array_equal(out_c.cpu().numpy()[ok], g["out_compat"][ok])
    assert not out_c.cpu().numpy()[~ok].any()


def test_predict_chain_golden_rgb(ctx):
    if not cvref.have_cv2():
        pytest.skip("cv2 needed to regenerate the synthetic inputs")
    g = golden("predict_chain.npz")
    for seed, h, w in g["rgb_cases"]:
        img = synth_upload(int(seed), int(h), int(w), True)
        out, _, small = ctx.predict_chain(dev(img[None]), channel_order="RGB", want_small20=True)
        assert np.array_equal(small[0].cpu().numpy(), g[f"small20_{seed}"]), (seed, h, w)
        assert np.array_equal(out[0].cpu().numpy(), g[f"out_{seed}"])
        # BGR buffer of the same picture
        out2, _, small2 = ctx.predict_chain(dev(img[None, :, :, ::-1]), channel_order="BGR", want_small20=True)
        assert torch.equal(small2, small)


def test_predict_chain_random_shapes_vs_oracle(ctx):
    rng = np.random.default_rng(31)
    shapes = [(20, 20), (21, 23), (22, 64), (100, 40), (128, 128), (255, 513), (600, 36), (37, 1000)]
    for (h, w) in shapes:
        for c in (1, 3):
            n = 2
            imgs = rng.integers(0, 256, (n, h, w, c), np.uint8)
            if c == 1:
                imgs = imgs[..., 0]
            out, status, small = ctx.predict_chain(dev(imgs), want_small20=True, compat=True)
            for i in range(n):
                gray = imgs[i] if c == 1 else R.pil_rgb2l(imgs[i])
                es = R.pil_resize_bicubic_l(gray)
                assert np.array_equal(small[i].cpu().numpy(), es), (h, w, c)
                try:
                    eo = R.predict_canvas(es, compat=True)
                    assert int(status[i]) == 0
                    assert np.array_equal(out[i].cpu().numpy(), eo)
                except IndexError:
                    assert int(status[i]) == 1


def test_predict_chain_all_kernels_agree(ctx, monkeypatch):
    """The three kernels behind nc_pp_predict_chain give identical bytes: TMA-staged rows + tensor-core
    (u8 x s8 IMMA) horizontal pass, TMA-staged rows + dp4a horizontal pass, and the generic kernel.
    Shapes cover every instantiation of the IMMA kernel (<= 5, <= 9, <= 16 chunks per warp), upscaling
    (w < 20), partial last stages and widths past the IMMA plan (falls back to dp4a)."""
    rng = np.random.default_rng(33)
    for (h, w) in [(16, 16), (40, 64), (97, 160), (768, 1024), (333, 1008), (50, 1936), (17, 2048), (200, 320), (31, 48)]:
        imgs = dev(rng.integers(0, 256, (3, h, w, 3), np.uint8))
        monkeypatch.delenv("NC_PREDICT_NO_TMA", raising=False)
        monkeypatch.delenv("NC_PREDICT_NO_MMA", raising=False)
        o1, s1, m1 = ctx.predict_chain(imgs, compat=True, want_small20=True)
        monkeypatch.setenv("NC_PREDICT_NO_MMA", "1")
        o3, s3, m3 = ctx.predict_chain(imgs, compat=True, want_small20=True)
        monkeypatch.setenv("NC_PREDICT_NO_TMA", "1")
        o2, s2, m2 = ctx.predict_chain(imgs, compat=True, want_small20=True)
        monkeypatch.delenv("NC_PREDICT_NO_TMA", raising=False)
        monkeypatch.delenv("NC_PREDICT_NO_MMA", raising=False)
        assert torch.equal(m1, m2) and torch.equal(m1, m3), (h, w)
        assert torch.equal(o1, o2) and torch.equal(s1, s2) and torch.equal(o1, o3) and torch.equal(s1, s3)
        es = R.pil_resize_bicubic_l(R.pil_rgb2l(imgs[1].cpu().numpy()))
        assert np.array_equal(m1[1].cpu().numpy(), es), (h, w)
    # extreme pixel values: the signed-digit split must stay exact at the ends of the u8 range
    for fill in (0, 255):
        imgs = torch.full((2, 64, 1024, 3), fill, dtype=torch.uint8, device="cuda")
        imgs[1, :, ::2] = 255 - fill
        _, _, m = ctx.predict_chain(imgs, want_small20=True)
        for i in range(2):
            assert np.array_equal(m[i].cpu().numpy(), R.pil_resize_bicubic_l(R.pil_rgb2l(imgs[i].cpu().numpy())))


def test_both_chains_single_pass_equals_the_two_kernels(ctx, monkeypatch):
    """nc_pp_both_chains: the fused single-pass kernel (where the shape allows it) and the two-launch fallback produce
    the bytes of nc_pp_resize_gray_norm and nc_pp_predict_chain; the train side is also checked against the oracle."""
    rng = np.random.default_rng(35)
    seen = set()
    for (h, w, expect_fused) in [(768, 1024, True), (1080, 1920, True), (600, 800, True), (333, 1008, True), (480, 640, False),
                                 (97, 160, False), (28, 32, False), (768, 1000, False)]:
        imgs = dev(rng.integers(0, 256, (3, h, w, 3), np.uint8))
        monkeypatch.setenv("NC_FUSED_CHAINS", "1")
        u8, f32, pred, status, small, fused = ctx.both_chains(imgs, channel_order="BGR", compat=True, want_small20=True)
        monkeypatch.delenv("NC_FUSED_CHAINS")
        assert fused == expect_fused, (h, w, fused)
        seen.add(fused)
        e8, ef = ctx.resize_gray_norm(imgs)
        ep, es, em = ctx.predict_chain(imgs, channel_order="BGR", compat=True, want_small20=True)
        assert torch.equal(u8, e8) and torch.equal(f32, ef), (h, w)
        assert torch.equal(pred, ep) and torch.equal(status, es) and torch.equal(small, em), (h, w)
        o8, _ = R.train_chain(imgs[2].cpu().numpy())
        assert np.array_equal(u8[2].cpu().numpy(), o8), (h, w)
        # default (two launches): same bytes
        u8b, f32b, predb, _, _, fb = ctx.both_chains(imgs, channel_order="BGR", compat=True)
        assert not fb and torch.equal(u8b, u8) and torch.equal(f32b, f32) and torch.equal(predb, pred)
    assert seen == {True, False}
    # only one of the train-side outputs requested; empty batch; argument errors
    imgs = dev(rng.integers(0, 256, (2, 768, 1024, 3), np.uint8))
    full = ctx.both_chains(imgs)
    import ctypes as C
    from cloud_server_b200 import _lib
    f32 = torch.zeros((2, 784), dtype=torch.float32, device="cuda")
    pred = torch.zeros((2, 784), dtype=torch.float32, device="cuda")
    rc = _lib.lib().nc_pp_both_chains(ctx.handle, C.c_void_p(imgs.data_ptr()), 2, 768, 1024, 3, 768 * 1024 * 3, 1, 0, None,
                                      C.c_void_p(f32.data_ptr()), C.c_void_p(pred.data_ptr()), None, None, None, None)
    assert rc == 0 and torch.equal(f32, full[1]) and torch.equal(pred, full[2])
    rc = _lib.lib().nc_pp_both_chains(ctx.handle, C.c_void_p(imgs.data_ptr()), 2, 768, 1024, 3, 768 * 1024 * 3, 1, 0, None,
                                      None, C.c_void_p(pred.data_ptr()), None, None, None, None)
    assert rc == -1
    empty = ctx.both_chains(imgs[:0])
    assert empty[0].shape[0] == 0 and empty[5] is False


def test_predict_chain_full_size_properties(ctx):
    rng = np.random.default_rng(32)
    n = 48
    imgs = torch.from_numpy(rng.integers(0, 256, (n, 768, 1024, 3), np.uint8)).cuda()
    # smooth the noise on device so the threshold produces structure: block-constant pictures
    imgs = imgs[:, ::64, ::64].repeat_interleave(64, 1).repeat_interleave(64, 2).contiguous()
    out, _, small = ctx.predict_chain(imgs, want_small20=True)
    for i in (0, 47):
        a = imgs[i].cpu().numpy()
        es = R.pil_resize_bicubic_l(R.pil_rgb2l(a))
        assert np.array_equal(small[i].cpu().numpy(), es)
        assert np.array_equal(out[i].cpu().numpy(), R.predict_canvas(es, compat=False))
    vals = torch.unique(out)
    assert set(np.round(vals.cpu().numpy(), 6)).issubset({0.0, np.round(np.float32(254 / 255.0), 6)})
    # a constant picture resizes to the constant (weights are normalised to 2^22)
    const = torch.full((1, 768, 1024, 3), 200, dtype=torch.uint8, device="cuda")
    _, _, s = ctx.predict_chain(const, want_small20=True)
    assert int(s.min()) == 200 and int(s.max()) == 200


# ------------------------------------------------------------------ the 14 ops (P6-P19)
OP_IDS = {"flip_up_down": 0, "flip_left_right": 1, "transpose_image": 2, "adjust_brightness_contrast": 3,
          "mean_filter": 4, "gaussian_blur": 5, "median_filter": 6, "nl_denoise_gray": 7,
          "add_salt_pepper_noise": 8, "equalize_hist": 9, "clahe": 10, "erode": 11, "dilate": 12}


def test_ops_golden(ctx):
    g = golden("ops.npz")
    imgs = g["inputs"]
    d_imgs = dev(imgs)
    for ci in range(int(g["n_cases"][0])):
        name = str(g[f"case{ci}_name"])
        v1 = g[f"case{ci}_v1"].item()
        v2 = g[f"case{ci}_v2"].item()
        v1 = None if np.isnan(v1) else (int(v1) if bool(g[f"case{ci}_v1_is_int"]) else v1)
        v2 = None if np.isnan(v2) else (int(v2) if bool(g[f"case{ci}_v2_is_int"]) else v2)
        draws = g[f"case{ci}_draws"] if name == "add_salt_pepper_noise" else None
        out = ctx.op(OP_IDS[name], d_imgs, v1, v2, draws).cpu().numpy()
        bad = int((out != g[f"case{ci}_out"]).sum())
        assert bad == 0, (name, v1, v2, bad)


def test_ops_vs_oracle_more_parameters(ctx):
    rng = np.random.default_rng(41)
    imgs = rng.integers(0, 256, (6, 28, 28), np.uint8)
    imgs[4] = 9          # constant picture (equalize special case)
    imgs[5] = np.clip(rng.normal(128, 12, (28, 28)), 0, 255).astype(np.uint8)
    d = dev(imgs)
    cases = [("mean_filter", k) for k in (1, 4, 6, 7, 9, 16, 17, 21)] + \
            [("gaussian_blur", k) for k in (1, 7, 13, 15, 21, 27)] + \
            [("median_filter", k) for k in (1, 7, 9)] + \
            [("erode", k) for k in (1, 4, 5, 6, 7)] + [("dilate", k) for k in (1, 4, 6, 7)] + \
            [("equalize_hist", None), ("clahe", None), ("nl_denoise_gray", 5), ("nl_denoise_gray", 30)]
    for name, v in cases:
        out = ctx.op(OP_IDS[name], d, v).cpu().numpy()
        for i in range(len(imgs)):
            fn = getattr(R, name)
            want = fn(imgs[i]) if v is None else fn(imgs[i], v)
            assert np.array_equal(out[i], want), (name, v, i)
    for a, b in [(2, 3), (2, 3.5), (1.5, 3), (1.5, -20.25), (0.0, 0), (255, 255), (3, -7.0), (0.7, 300)]:
        out = ctx.op(3, d, a, b).cpu().numpy()
        for i in range(len(imgs)):
            assert np.array_equal(out[i], R.adjust_brightness_contrast(imgs[i], a, b)), (a, b)


def test_ops_errors_match_reference(ctx):
    d = torch.zeros((1, 28, 28), dtype=torch.uint8, device="cuda")
    from cloud_server_b200._lib import NcError
    with pytest.raises(NcError):
        ctx.op(5, d, 4)        # even Gaussian kernel: cv2.error in the reference
    with pytest.raises(NcError):
        ctx.op(6, d, 2)        # even median kernel
    with pytest.raises(ValueError):
        ctx.op(3, d, 300, 0)   # OverflowError in numpy 2


def test_ops_involutions_large_batch(ctx):
    rng = np.random.default_rng(42)
    d = torch.from_numpy(rng.integers(0, 256, (20000, 28, 28), np.uint8)).cuda()
    for op in (0, 1, 2):
        assert torch.equal(ctx.op(op, ctx.op(op, d)), d)
    assert torch.equal(ctx.op(2, d), ctx.op(0, ctx.op(1, d)))
    e = ctx.op(11, d, 3)
    assert bool((e <= d).all()) and bool((ctx.op(12, d, 3) >= d).all())
    assert torch.equal(ctx.normalise(d), (d.reshape(20000, 784).float() * np.float32(1 / 255.0)))
