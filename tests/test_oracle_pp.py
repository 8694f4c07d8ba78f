"""CPU tests: the numpy restatement (oracle/pp_restate.py) against the committed golden vectors
(generated from the reference's own module + cv2/Pillow by tests/golden/make_golden.py) and,
where the libraries are importable, against the live calls."""
import numpy as np
import pytest

from conftest import golden
from oracle import pp_restate as R
from oracle import cvref


def synth_upload(seed, h, w, natural):
    rng = np.random.default_rng(seed)
    a = rng.integers(0, 256, (h, w, 3), np.uint8)
    if natural:
        import cv2
        a = cv2.GaussianBlur(a, (0, 0), 3)
    return a


def test_train_chain_golden():
    g = golden("train_chain.npz")
    for seed, h, w, nat in g["cases"]:
        if nat and not cvref.have_cv2():
            continue
        img = synth_upload(int(seed), int(h), int(w), bool(nat))
        assert int(img.astype(np.uint64).sum()) == int(g[f"insum_{seed}"][0])
        u8, f32 = R.train_chain(img)
        assert np.array_equal(u8, g[f"u8_{seed}"]), (seed, h, w)
        assert np.array_equal(f32, g[f"f32_{seed}"])


def test_predict_chain_golden_pics():
    p = golden("pics.npz")
    g = golden("predict_chain.npz")
    n_err = 0
    for i, gray in enumerate(p["gray_pil"]):
        assert np.array_equal(R.pil_resize_bicubic_l(gray), g["small20"][i])
        assert np.array_equal(R.predict_chain(gray, compat=False), g["out_intended"][i])
        if g["compat_indexerror"][i]:
            n_err += 1
            with pytest.raises(IndexError):
                R.predict_chain(gray, compat=True)
        else:
            assert np.array_equal(R.predict_chain(gray, compat=True), g["out_compat"][i])
    assert n_err == 82


def test_predict_chain_golden_rgb():
    if not cvref.have_cv2():
        pytest.skip("cv2 needed to regenerate the synthetic inputs")
    g = golden("predict_chain.npz")
    for seed, h, w in g["rgb_cases"]:
        img = synth_upload(int(seed), int(h), int(w), True)
        small = R.pil_resize_bicubic_l(R.pil_rgb2l(img))
        assert np.array_equal(small, g[f"small20_{seed}"]), (seed, h, w)
        assert np.array_equal(R.predict_chain(img, compat=False), g[f"out_{seed}"])


def _restate_op(name, img, v1, v2, draws):
    if name == "adjust_brightness_contrast":
        return R.adjust_brightness_contrast(img, v1, v2)
    if name == "add_salt_pepper_noise":
        return R.add_salt_pepper_noise(img, v1, draws)
    fn = getattr(R, name)
    return fn(img) if v1 is None else fn(img, v1)


def test_ops_golden():
    g = golden("ops.npz")
    imgs = g["inputs"]
    for ci in range(int(g["n_cases"][0])):
        name = str(g[f"case{ci}_name"])
        v1 = g[f"case{ci}_v1"].item()
        v2 = g[f"case{ci}_v2"].item()
        v1 = None if np.isnan(v1) else (int(v1) if bool(g[f"case{ci}_v1_is_int"]) else v1)
        v2 = None if np.isnan(v2) else (int(v2) if bool(g[f"case{ci}_v2_is_int"]) else v2)
        sel = range(len(imgs)) if name != "nl_denoise_gray" else range(0, len(imgs), 6)
        for k in sel:
            draws = g[f"case{ci}_draws"][k] if name == "add_salt_pepper_noise" else None
            got = _restate_op(name, imgs[k], v1, v2, draws)
            assert np.array_equal(got, g[f"case{ci}_out"][k]), (name, v1, v2, k)


def test_brightness_mixed_int_float_matches_numpy():
    rng = np.random.default_rng(7)
    img = rng.integers(0, 256, (28, 28), np.uint8)
    for a, b in [(2, 3), (2, 3.5), (1.5, 3), (1.5, -20.25), (0, 0), (255, 255), (3, -7.0)]:
        want = cvref.op("adjust_brightness_contrast", img, a, b) if cvref.have_cv2() else None
        got = R.adjust_brightness_contrast(img, a, b)
        if want is not None:
            assert np.array_equal(got, want), (a, b)
    with pytest.raises(OverflowError):
        R.adjust_brightness_contrast(img, 300, 0)
    with pytest.raises(OverflowError):
        R.adjust_brightness_contrast(img, 2, -1)


def test_ref_files_tier_b():
    """The reference's file-based functions (JPEG round trips) stay within JPEG noise of the
    in-memory chain; identity resize of a 28x28 picture keeps it within +-8 grey levels."""
    p = golden("pics.npz")
    t = golden("ref_files.npz")
    n = len(t["names"])
    d = np.abs(t["resize_then_gray"].astype(int) - p["gray_cv"][:n].astype(int))
    assert d.max() <= 12 and d.mean() < 2.0
    d2 = np.abs(t["big_decoded_train_u8"].astype(int) - t["big_file_gray"].astype(int))
    assert d2.max() <= 12


@pytest.mark.skipif(not cvref.have_cv2(), reason="cv2 not importable")
def test_live_cv2_random_shapes():
    rng = np.random.default_rng(11)
    for _ in range(25):
        h, w = int(rng.integers(29, 300)), int(rng.integers(29, 300))
        img = rng.integers(0, 256, (h, w, 3), np.uint8)
        a, af = cvref.train_chain(img)
        b, bf = R.train_chain(img)
        assert np.array_equal(a, b) and np.array_equal(af, bf)
    for _ in range(8):
        h, w = int(rng.integers(21, 300)), int(rng.integers(21, 300))
        img = rng.integers(0, 256, (h, w, 3), np.uint8)
        assert np.array_equal(cvref.predict_small20(img), R.pil_resize_bicubic_l(R.pil_rgb2l(img)))
    imgs = [rng.integers(0, 256, (28, 28), np.uint8) for _ in range(6)]
    for im in imgs:
        for k in (13, 15, 21, 27):
            assert np.array_equal(cvref.op("gaussian_blur", im, k), R.gaussian_blur(im, k)), k
        for k in (6, 9, 16, 17, 20):
            assert np.array_equal(cvref.op("mean_filter", im, k), R.mean_filter(im, k)), k
        for k in (7, 9, 13):
            assert np.array_equal(cvref.op("median_filter", im, k), R.median_filter(im, k)), k
        for k in (1, 6, 7):
            assert np.array_equal(cvref.op("erode", im, k), R.erode(im, k)), k
            assert np.array_equal(cvref.op("dilate", im, k), R.dilate(im, k)), k


def test_ipp_arithmetic_model_golden_and_live():
    """P2 mode 1: the float model of cv2.resize's DEFAULT arithmetic (Intel IPP) -- oracle.cv_resize_cubic_ipp -- against the
    committed golden decodes (tests/golden/make_golden_ipp.py ran the wheel with IPP on) and, where cv2 is importable, the live
    call: >= 99.9 % of the outputs equal (measured here: 100 %).  Mode 0 (IPP off, the canonical target) differs from the
    default wheel on ~3-5 % of the outputs by +-1: stated, not hidden."""
    g = golden("train_chain_ipp.npz")
    same = tot = 0
    for seed, h, w, nat in g["cases"]:
        if nat and not cvref.have_cv2():
            continue
        img = synth_upload(int(seed), int(h), int(w), bool(nat))
        assert int(img.astype(np.uint64).sum()) == int(g[f"insum_{seed}"][0])
        got = R.cv_resize_cubic_ipp(img)
        same += int((got == g[f"bgr_{seed}"]).sum())
        tot += got.size
        u8, _ = R.train_chain(img, mode=1)
        assert (u8 != g[f"u8_{seed}"]).mean() <= 1e-3
    assert same >= 0.999 * tot, (same, tot)
    d, t = g["mode0_vs_ipp"]
    assert 0.01 < d / t < 0.08                        # the two arithmetics of the same wheel really differ
    if cvref.have_cv2():
        import cv2
        rng = np.random.default_rng(77)
        try:
            cv2.ipp.setUseIPP(True)
            if not cv2.ipp.useIPP():
                pytest.skip("this cv2 build has no IPP")
            same = tot = 0
            for h, w in ((768, 1024), (480, 640), (100, 80), (255, 257), (31, 400), (333, 30)):
                img = rng.integers(0, 256, (h, w, 3), np.uint8)
                want = cv2.resize(img, (28, 28), interpolation=cv2.INTER_CUBIC)
                got = R.cv_resize_cubic_ipp(img)
                same += int((got == want).sum())
                tot += want.size
            assert same >= 0.999 * tot, (same, tot)
        finally:
            cv2.ipp.setUseIPP(False)


def test_ipp_host_tables_equal_oracle(built_lib):
    import ctypes as C
    for src in (29, 30, 100, 257, 768, 1024, 1920):
        ofs = np.zeros(112, np.int32)
        coef = np.zeros(112, np.float32)
        assert built_lib.nc_pp_cubic_table_f32(src, ofs.ctypes.data_as(C.c_void_p), coef.ctypes.data_as(C.c_void_p)) == 0
        o, c = R.ipp_cubic_tables(src, 28)
        assert np.array_equal(ofs.reshape(28, 4), o) and np.array_equal(coef.reshape(28, 4), c), src
