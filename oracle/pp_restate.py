"""numpy restatement of the preprocess half of the hot path (TEST INFRASTRUCTURE ONLY).

The reference's arithmetic for this half lives in un-vendored wheels (opencv-python,
Pillow, numpy -- ``requirment.txt:13-17``, unpinned; the versions characterised here are
opencv-python-headless 4.13.0 with ``cv2.ipp.setUseIPP(False)``, Pillow 12.2.0,
numpy 2.3.5).  Each function restates the published algorithm of the routine the
reference calls and cites the reference call site (paths relative to /root/reference).
No cv2 / PIL import in this file: it is pure numpy so that it is an independent
statement of the arithmetic; ``oracle/cvref.py`` holds the live library calls used to
pin it.

Arithmetic type: integers (int32/int64) and IEEE float32 exactly as the library does;
every output here is bit-exact against the library on the fixtures in tests/golden.
"""
from __future__ import annotations

import numpy as np

F32 = np.float32

# --------------------------------------------------------------------------------------
# P2: cv2.resize(im1, (28, 28), interpolation=cv2.INTER_CUBIC)
#     apps/preprocess/preprocess.py:29
# --------------------------------------------------------------------------------------


def cv_cubic_tables(src: int, dst: int):
    """Tap offsets (clamped) and int16 coefficients of OpenCV's INTER_CUBIC for one axis.

    OpenCV ``resize`` (imgproc/resize.cpp, resizeGeneric_ set-up): ``scale = 1/(dst/src)`` in
    double, ``fx = float((d+0.5)*scale-0.5)``, ``s=floor(fx)``, ``fx-=s``; Keys cubic with
    A=-0.75 evaluated in float32; coefficient = saturate_cast<short>(c*2048) (round half even).
    Taps s-1..s+2 clamped to [0, src-1].
    """
    inv_scale = float(dst) / float(src)
    scale = 1.0 / inv_scale
    ofs = np.zeros((dst, 4), np.int32)
    coef = np.zeros((dst, 4), np.int16)
    A = F32(-0.75)
    one = F32(1)
    for d in range(dst):
        fx = F32((d + 0.5) * scale - 0.5)
        s = int(np.floor(fx))
        x = F32(fx - F32(s))
        xp = x + one
        c0 = ((A * xp - F32(5) * A) * xp + F32(8) * A) * xp - F32(4) * A
        c1 = ((A + F32(2)) * x - (A + F32(3))) * x * x + one
        xm = one - x
        c2 = ((A + F32(2)) * xm - (A + F32(3))) * xm * xm + one
        c3 = one - c0 - c1 - c2
        cs = np.array([c0, c1, c2, c3], F32) * F32(2048)
        coef[d] = np.rint(cs).astype(np.int16)
        ofs[d] = np.clip(np.arange(s - 1, s + 3), 0, src - 1)
    return ofs, coef


def cv_resize_cubic_u8(img: np.ndarray, dh: int = 28, dw: int = 28, simd_lanes: int = 8) -> np.ndarray:
    """``cv2.resize(img,(dw,dh),INTER_CUBIC)`` for uint8 HxWxC, OpenCV-native (non-IPP) path.

    Horizontal pass: exact int32 ``sum_k S[x_k]*alpha_k`` (HResizeCubic<uchar,int,short>).
    Vertical pass: OpenCV's universal-intrinsic functor VResizeCubicVec_32s8u handles the
    first ``(dw*C)//lanes*lanes`` interleaved elements of a row in float32 --
    ``rint(f(S0)*b0 + (f(S1)*b1 + (f(S2)*b2 + f(S3)*b3)))`` with ``b=beta/2^22``, separate
    multiply and add (baseline build, no FMA), round half to even, saturate -- and the scalar
    tail uses the integer ``(sum + 2^21) >> 22`` (FixedPtCast).  Verified 0 mismatches over
    1.4 M outputs against cv2 4.13.0 (IPP off); the pure-integer model alone leaves ~45 ppm.
    When src size == dst size OpenCV copies (identity).
    """
    assert img.dtype == np.uint8
    if img.ndim == 2:
        img = img[:, :, None]
        squeeze = True
    else:
        squeeze = False
    H, W, C = img.shape
    if (H, W) == (dh, dw):
        out = img.copy()
        return out[:, :, 0] if squeeze else out
    xo, xa = cv_cubic_tables(W, dw)
    yo, ya = cv_cubic_tables(H, dh)
    # horizontal pass only on the rows the vertical pass will read
    rows = np.unique(yo)
    row_pos = {int(r): i for i, r in enumerate(rows)}
    I = img[rows].astype(np.int32)
    hp = np.zeros((len(rows), dw, C), np.int32)
    for k in range(4):
        hp += I[:, xo[:, k], :] * xa[:, k].astype(np.int32)[None, :, None]
    ridx = np.vectorize(lambda r: row_pos[int(r)])(yo)  # (dh,4) -> index into hp
    S = [hp[ridx[:, k]].reshape(dh, dw * C) for k in range(4)]
    # integer vertical (scalar tail)
    acc = np.zeros((dh, dw * C), np.int64)
    for k in range(4):
        acc += S[k].astype(np.int64) * ya[:, k].astype(np.int64)[:, None]
    out_i = np.clip((acc + (1 << 21)) >> 22, 0, 255)
    # float vertical (SIMD body)
    sc = F32(1.0) / F32(2048 * 2048)
    b = ya.astype(F32) * sc
    t = S[3].astype(F32) * b[:, 3][:, None]
    t = S[2].astype(F32) * b[:, 2][:, None] + t
    t = S[1].astype(F32) * b[:, 1][:, None] + t
    t = S[0].astype(F32) * b[:, 0][:, None] + t
    out_f = np.clip(np.rint(t), 0, 255).astype(np.int64)
    n_simd = (dw * C) // simd_lanes * simd_lanes
    out = out_i.copy()
    out[:, :n_simd] = out_f[:, :n_simd]
    out = out.astype(np.uint8).reshape(dh, dw, C)
    return out[:, :, 0] if squeeze else out


def ipp_cubic_tables(src: int, dst: int):
    """Tap offsets and float32 coefficients of the IPP path of cv2.resize INTER_CUBIC (one axis): position and Keys cubic
    (A = -0.75) evaluated in float64, coefficients then rounded to float32 -- NOT quantised to 11 bits."""
    scale = src / dst
    ofs = np.zeros((dst, 4), np.int32)
    coef = np.zeros((dst, 4), F32)
    A = -0.75
    for d in range(dst):
        f = (d + 0.5) * scale - 0.5
        s = int(np.floor(f))
        x = f - s
        c0 = ((A * (x + 1) - 5 * A) * (x + 1) + 8 * A) * (x + 1) - 4 * A
        c1 = ((A + 2) * x - (A + 3)) * x * x + 1
        c2 = ((A + 2) * (1 - x) - (A + 3)) * (1 - x) * (1 - x) + 1
        c3 = 1.0 - c0 - c1 - c2
        coef[d] = np.array([c0, c1, c2, c3], np.float64).astype(F32)
        ofs[d] = np.clip(np.arange(s - 1, s + 3), 0, src - 1)
    return ofs, coef


def cv_resize_cubic_ipp(img: np.ndarray, dh: int = 28, dw: int = 28) -> np.ndarray:
    """``cv2.resize(img,(dw,dh),INTER_CUBIC)`` for uint8 as the opencv-python wheel computes it BY DEFAULT (IPP enabled,
    ippiResizeCubic; closed source).  Float model fitted to the live call: horizontal pass then vertical pass, float32
    coefficients (ipp_cubic_tables), float32 products and sums taken tap by tap in order (no FMA), round half to even,
    saturate.  0 mismatches over 94 080 outputs (40 pictures, 8 shapes) against cv2 4.13.0 / IPP 2022.2 in this image
    (tests/test_oracle_pp.py); the reference's requirment.txt does not pin the wheel, so this is the arithmetic its
    preprocess.resize (apps/preprocess/preprocess.py:29) runs with unless the host disables IPP."""
    assert img.dtype == np.uint8
    squeeze = img.ndim == 2
    if squeeze:
        img = img[:, :, None]
    H, W, C = img.shape
    if (H, W) == (dh, dw):
        return img[:, :, 0].copy() if squeeze else img.copy()
    xo, xa = ipp_cubic_tables(W, dw)
    yo, ya = ipp_cubic_tables(H, dh)
    rows = np.unique(yo)
    pos = {int(r): i for i, r in enumerate(rows)}
    I = img[rows].astype(F32)
    t = np.zeros((len(rows), dw, C), F32)
    for k in range(4):
        t = t + I[:, xo[:, k], :] * xa[:, k][None, :, None]
    ridx = np.vectorize(lambda r: pos[int(r)])(yo)
    o = np.zeros((dh, dw, C), F32)
    for k in range(4):
        o = o + t[ridx[:, k]] * ya[:, k][:, None, None]
    out = np.clip(np.rint(o), 0, 255).astype(np.uint8)
    return out[:, :, 0] if squeeze else out


# --------------------------------------------------------------------------------------
# P3: gray conversion of cv2.imread(dir, 0) on the resized BGR file
#     apps/preprocess/preprocess.py:40 (and every op's first line)
# --------------------------------------------------------------------------------------


def cv_bgr2gray(img: np.ndarray) -> np.ndarray:
    """cv2 4.13 COLOR_BGR2GRAY for uint8: (3735*B + 19235*G + 9798*R + 2^14) >> 15."""
    b = img[..., 0].astype(np.int32)
    g = img[..., 1].astype(np.int32)
    r = img[..., 2].astype(np.int32)
    return ((3735 * b + 19235 * g + 9798 * r + (1 << 14)) >> 15).astype(np.uint8)


# --------------------------------------------------------------------------------------
# P4: DataSet.__init__  apps/construction/util/construct_distribute.py:497-504
# --------------------------------------------------------------------------------------


def dataset_normalise(gray_u8: np.ndarray) -> np.ndarray:
    """[N,28,28] u8 -> [N,784] f32: astype(float32) then multiply by float32(1/255)."""
    n = gray_u8.shape[0] if gray_u8.ndim == 3 else 1
    flat = gray_u8.reshape(n, -1).astype(F32)
    return np.multiply(flat, F32(1.0 / 255.0))


def train_chain(img_bgr: np.ndarray, mode: int = 0):
    """resize (preprocess.py:28-30) -> gray (:40) -> normalise (construct_distribute.py:503-504)
    composed in memory (no JPEG round trip).  Returns (u8[28,28], f32[784]).  mode 0: OpenCV's own resize arithmetic
    (IPP disabled), mode 1: the wheel's default (IPP) arithmetic."""
    small = cv_resize_cubic_ipp(img_bgr, 28, 28) if mode == 1 else cv_resize_cubic_u8(img_bgr, 28, 28)
    gray = cv_bgr2gray(small) if small.ndim == 3 else small
    return gray, dataset_normalise(gray[None])[0]


# --------------------------------------------------------------------------------------
# P5: predict-side chain  apps/construction/util/construct_inference.py:312-330
# --------------------------------------------------------------------------------------


def pil_rgb2l(img_rgb: np.ndarray) -> np.ndarray:
    """Pillow ``convert('L')`` from RGB (Convert.c rgb2l): (19595R+38470G+7471B+0x8000)>>16."""
    r = img_rgb[..., 0].astype(np.int64)
    g = img_rgb[..., 1].astype(np.int64)
    b = img_rgb[..., 2].astype(np.int64)
    return ((19595 * r + 38470 * g + 7471 * b + 0x8000) >> 16).astype(np.uint8)


def _pil_bicubic(x: float) -> float:
    a = -0.5
    if x < 0.0:
        x = -x
    if x < 1.0:
        return ((a + 2.0) * x - (a + 3.0)) * x * x + 1
    if x < 2.0:
        return (((x - 5) * x + 8) * x - 4) * a
    return 0.0


def pil_bicubic_tables(in_size: int, out_size: int):
    """Pillow Resample.c precompute_coeffs + normalize_coeffs_8bpc for BICUBIC (support 2).

    Returns (xmin[out], count[out], kk[out, ksize] int32) with 22-bit fixed-point weights
    ``int(+-0.5 + w * 2^22)`` (C truncation toward zero)."""
    scale = float(in_size) / float(out_size)
    filterscale = max(scale, 1.0)
    support = 2.0 * filterscale
    ksize = int(np.ceil(support)) * 2 + 1
    xmin_a = np.zeros(out_size, np.int32)
    cnt_a = np.zeros(out_size, np.int32)
    kk = np.zeros((out_size, ksize), np.int32)
    ss = 1.0 / filterscale
    for xx in range(out_size):
        center = (xx + 0.5) * scale
        xmin = int(center - support + 0.5)
        if xmin < 0:
            xmin = 0
        xmax = int(center + support + 0.5)
        if xmax > in_size:
            xmax = in_size
        xmax -= xmin
        w = np.array([_pil_bicubic((x + xmin - center + 0.5) * ss) for x in range(xmax)], np.float64)
        ww = 0.0
        for v in w:  # same summation order as the C loop
            ww += v
        if ww != 0.0:
            w = w / ww
        q = np.where(w < 0, -0.5 + w * (1 << 22), 0.5 + w * (1 << 22))
        kk[xx, :xmax] = np.trunc(q).astype(np.int32)
        xmin_a[xx] = xmin
        cnt_a[xx] = xmax
    return xmin_a, cnt_a, kk


def pil_resize_bicubic_l(gray: np.ndarray, out_w: int = 20, out_h: int = 20) -> np.ndarray:
    """Pillow ``Image.resize((out_w,out_h))`` on an 'L' image (default resample BICUBIC since
    Pillow 7; antialiased).  Horizontal pass to a uint8 intermediate, then vertical pass, each
    ``clip8((2^21 + sum pixel*k) >> 22)`` (ImagingResampleHorizontal_8bpc / Vertical_8bpc)."""
    H, W = gray.shape
    cur = gray
    if W != out_w:
        xmin, cnt, kk = pil_bicubic_tables(W, out_w)
        tmp = np.zeros((H, out_w), np.uint8)
        g = cur.astype(np.int64)
        for xx in range(out_w):
            seg = g[:, xmin[xx]:xmin[xx] + cnt[xx]]
            acc = (1 << 21) + seg @ kk[xx, :cnt[xx]].astype(np.int64)
            tmp[:, xx] = np.clip(acc >> 22, 0, 255)
        cur = tmp
    if H != out_h:
        ymin, cnt, kk = pil_bicubic_tables(H, out_h)
        out = np.zeros((out_h, cur.shape[1]), np.uint8)
        g = cur.astype(np.int64)
        for yy in range(out_h):
            seg = g[ymin[yy]:ymin[yy] + cnt[yy], :]
            acc = (1 << 21) + kk[yy, :cnt[yy]].astype(np.int64) @ seg
            out[yy] = np.clip(acc >> 22, 0, 255)
        cur = out
    return cur


def predict_canvas(small20: np.ndarray, compat: bool = True):
    """construct_inference.py:313-330 on the 20x20 'L' array.

    ``> 150 -> 254`` into a 28x28 zero canvas at offset 4 (:319-324).  In compat mode the
    ``elif arr[i][j] > 100`` branch (:325-326) writes ``arr[4+i][4+j]`` of the 20x20 array, which
    raises IndexError when 4+i or 4+j >= 20 -- and, where in range, overwrites pixels the loop
    has not visited yet with 220 (which then pass the >150 test).  ``/255.0`` in float64, fed to
    a float32 placeholder.  Returns f32[784]; raises IndexError exactly when the reference does."""
    line = 20
    bizar = int((28 - line) / 2)
    arr = np.array(small20, dtype="float32")
    hist = np.zeros((28, 28))
    for i in range(line):
        for j in range(line):
            if arr[i][j] > 150:
                hist[i + bizar][bizar + j] = 254
            elif arr[i][j] > 100:
                if compat:
                    arr[bizar + i][bizar + j] = 220  # IndexError when out of range, as the reference
    hist /= 255.0
    return hist.flatten().astype(F32)


def predict_chain(img: np.ndarray, compat: bool = True, channels: str = "RGB"):
    """Whole predict-side preprocess for one decoded image (HxW 'L' or HxWx3)."""
    if img.ndim == 3:
        rgb = img if channels == "RGB" else img[..., ::-1]
        gray = pil_rgb2l(rgb)
    else:
        gray = img
    small = pil_resize_bicubic_l(gray, 20, 20)
    return predict_canvas(small, compat)


# --------------------------------------------------------------------------------------
# P6-P19: the 14 per-image ops on 28x28 gray u8 (apps/preprocess/preprocess.py:33-227)
# --------------------------------------------------------------------------------------


def flip_up_down(img):  # preprocess.py:41  cv2.flip(img, 0)
    return img[::-1].copy()


def flip_left_right(img):  # preprocess.py:56  cv2.flip(img, 1)
    return img[:, ::-1].copy()


def transpose_image(img):  # preprocess.py:71  cv2.flip(img, -1): 180 degree rotation
    return img[::-1, ::-1].copy()


def adjust_brightness_contrast(img, alpha, beta):
    """preprocess.py:89-95.  ``img[xj,xi] = img[xj,xi]*alpha + beta`` with a numpy uint8 scalar on
    the left.  numpy 2 (NEP 50): uint8 (op) Python int stays uint8 -- wraps modulo 256 and raises
    OverflowError for ints outside [0,255]; uint8 (op) Python float gives float64.  The float64
    result is stored into the uint8 array with a C cast (truncate toward zero, keep the low
    byte), so the ``> 255`` / ``< 0`` clamps at :92-95 never fire."""
    def is_int(v):
        return isinstance(v, (int, np.integer)) and not isinstance(v, bool)

    p = img.astype(np.int64)
    if is_int(alpha):
        if not 0 <= int(alpha) <= 255:
            raise OverflowError("Python integer %d out of bounds for uint8" % alpha)
        t = (p * int(alpha)) & 0xFF
        t_is_u8 = True
    else:
        t = img.astype(np.float64) * float(alpha)
        t_is_u8 = False
    if t_is_u8 and is_int(beta):
        if not 0 <= int(beta) <= 255:
            raise OverflowError("Python integer %d out of bounds for uint8" % beta)
        return ((t + int(beta)) & 0xFF).astype(np.uint8)
    v = t.astype(np.float64) + float(beta)
    q = np.trunc(v).astype(np.int64)
    return (q & 0xFF).astype(np.uint8)


def _reflect101(i, n):
    i = np.abs(i)
    return np.where(i >= n, 2 * (n - 1) - i, i)


def _gather_window(img, k, border):
    """All k*k shifted copies of img (anchor k//2) under the given border mode."""
    H, W = img.shape
    a = k // 2
    ys = np.arange(H)[:, None] + (np.arange(k) - a)[None, :]
    xs = np.arange(W)[:, None] + (np.arange(k) - a)[None, :]
    if border == "reflect101":
        ys, xs = _reflect101(ys, H), _reflect101(xs, W)
        valid = None
    elif border == "replicate":
        ys, xs = np.clip(ys, 0, H - 1), np.clip(xs, 0, W - 1)
        valid = None
    else:  # ignore: mark invalid
        vy = (ys >= 0) & (ys < H)
        vx = (xs >= 0) & (xs < W)
        ys, xs = np.clip(ys, 0, H - 1), np.clip(xs, 0, W - 1)
        valid = vy[:, None, :, None] & vx[None, :, None, :]
    win = img[ys[:, None, :, None], xs[None, :, None, :]]  # H,W,k,k
    return win, valid


def mean_filter(img, k):
    """preprocess.py:125 cv2.blur(img,(k,k)): BORDER_REFLECT_101, anchor k//2.  For uint8 with
    k*k <= 256 OpenCV sums in ushort and divides with the integer reciprocal of
    ColumnSum<ushort,uchar> (imgproc/box_filter.simd.hpp): d=k*k, f=2^23/d, divScale=floor(f),
    divDelta=d/2, and if frac(f) < 0.5 divDelta += 1 else divScale += 1;
    out = ((sum + divDelta) * divScale) >> 23.  (For k=2 that is (sum+3)>>2: x.25 rounds up.)
    Larger windows use an int sum and saturate_cast<uchar>(sum * (1/d)) in double."""
    win, _ = _gather_window(img, k, "reflect101")
    s = win.astype(np.int64).sum(axis=(2, 3))
    d = k * k
    if d == 1:
        return img.copy()
    if d <= 256:
        f = float(1 << 23) / d
        div_scale = int(np.floor(f))
        frac = f - div_scale
        div_delta = d // 2
        if frac < 0.5:
            div_delta += 1
        else:
            div_scale += 1
        return (((s + div_delta) * div_scale) >> 23).astype(np.uint8)
    return np.clip(np.rint(s.astype(np.float64) * (1.0 / d)), 0, 255).astype(np.uint8)


_GAUSS_SMALL = {1: [1.0], 3: [0.25, 0.5, 0.25], 5: [0.0625, 0.25, 0.375, 0.25, 0.0625],
                7: [0.03125, 0.109375, 0.21875, 0.28125, 0.21875, 0.109375, 0.03125],
                9: [4 / 256, 13 / 256, 30 / 256, 51 / 256, 60 / 256, 51 / 256, 30 / 256, 13 / 256, 4 / 256]}


def gaussian_kernel_1d(k):
    """cv::getGaussianKernelBitExact(k, sigma<=0) (imgproc/smooth.dispatch.cpp): fixed kernels for
    k in {1,3,5,7,9}; otherwise sigma = 0.15*k + 0.35 and exp(-0.125*x^2/sigma^2), x = 1-k step 2,
    normalised by 1/sum (centre tap taken as exactly 1 before normalisation)."""
    if k in _GAUSS_SMALL:
        return np.array(_GAUSS_SMALL[k], np.float64)
    sigma = k * 0.15 + 0.35
    scale2x = -0.125 / (sigma * sigma)
    n2 = (k - 1) // 2
    vals = [float(np.exp(float(x * x) * scale2x)) for x in range(1 - k, 0, 2)][:n2]
    tot = 0.0
    for v in vals:
        tot += v
    tot = tot * 2.0 + 1.0
    mul1 = 1.0 / tot
    half = [v * mul1 for v in vals]
    return np.array(half + [mul1] + half[::-1], np.float64)


def gaussian_kernel_q8(k):
    """getGaussianKernelFixedPoint_ED with 8 fraction bits: left half rounded with error
    diffusion (round-half-even of w*256 + carried error), mirrored, centre = 256 - 2*sum."""
    w = gaussian_kernel_1d(k)
    n2 = k // 2
    q = np.zeros(k, np.int64)
    err = 0.0
    tot = 0
    for i in range(n2):
        adj = w[i] * 256.0 + err
        v0 = int(np.rint(adj))
        err = adj - v0
        q[i] = v0
        q[k - 1 - i] = v0
        tot += v0
    q[n2] = 256 - 2 * tot
    return q


def gaussian_blur(img, k):
    """preprocess.py:134 cv2.GaussianBlur(img,(k,k),0) on uint8.  OpenCV's 8-bit path
    (imgproc/smooth.dispatch.cpp -> fixed-point ufixedpoint16 separable filter, 8 fractional
    bits per pass): kernel weights quantised to Q8 by ``gaussian_kernel_q8``,
    horizontal pass produces Q8 values (u16), vertical pass accumulates Q16 in u32 and rounds
    ``(v + 2^15) >> 16``; BORDER_REFLECT_101.  k must be odd (cv2 raises otherwise)."""
    if k % 2 == 0 or k <= 0:
        raise ValueError("ksize must be odd and positive (cv2.error in the reference)")
    q = gaussian_kernel_q8(k)
    H, W = img.shape
    a = k // 2
    xs = _reflect101(np.arange(W)[:, None] + (np.arange(k) - a)[None, :], W)
    ys = _reflect101(np.arange(H)[:, None] + (np.arange(k) - a)[None, :], H)
    g = img.astype(np.int64)
    hp = (g[:, xs] * q[None, None, :]).sum(axis=2)          # Q8
    vp = (hp[ys, :] * q[None, :, None]).sum(axis=1)          # Q16
    return np.clip((vp + (1 << 15)) >> 16, 0, 255).astype(np.uint8)


def median_filter(img, k):
    """preprocess.py:144 cv2.medianBlur(img,k,0) (third positional is dst): BORDER_REPLICATE,
    k odd > 1 (k == 1 copies)."""
    if k % 2 == 0 or k < 1:
        raise ValueError("ksize must be odd (cv2.error in the reference)")
    if k == 1:
        return img.copy()
    win, _ = _gather_window(img, k, "replicate")
    flat = np.sort(win.reshape(img.shape[0], img.shape[1], k * k), axis=2)
    return flat[:, :, (k * k) // 2].astype(np.uint8)


def _morph(img, k, is_erode):
    win, valid = _gather_window(img, k, "ignore")
    if is_erode:
        w = np.where(valid, win, 255)
        return w.min(axis=(2, 3)).astype(np.uint8)
    w = np.where(valid, win, 0)
    return w.max(axis=(2, 3)).astype(np.uint8)


def erode(img, k):
    """preprocess.py:212-213 cv2.erode(img, ones(k,k)): anchor k//2, out-of-image ignored."""
    return _morph(img, k, True)


def dilate(img, k):
    """preprocess.py:222-223 cv2.dilate(img, ones(k,k))."""
    return _morph(img, k, False)


def equalize_hist(img):
    """preprocess.py:193 cv2.equalizeHist (imgproc/histogram.cpp): 256-bin histogram, i0 = first
    non-empty bin; if all pixels are in that bin the image is filled with i0; else
    scale = 255.f/(total - hist[i0]); lut[i0]=0; running sum over bins after i0,
    lut[i] = saturate_cast<uchar>(sum * scale) (float32 product, round half even)."""
    hist = np.bincount(img.ravel(), minlength=256).astype(np.int64)
    total = img.size
    i0 = int(np.nonzero(hist)[0][0])
    if hist[i0] == total:
        return np.full_like(img, i0)
    scale = F32(255.0) / F32(total - hist[i0])
    lut = np.zeros(256, np.uint8)
    s = 0
    for i in range(i0 + 1, 256):
        s += int(hist[i])
        lut[i] = np.clip(np.rint(F32(s) * scale), 0, 255)
    return lut[img]


def salt_pepper_indices(percent, draws):
    """preprocess.py:162-176: m = int(28*28*percent); 2m draws (i, j) for salt then 2m for pepper,
    each ``int(np.random.random()*28)``.  ``draws`` is the sequence of 4m uniform doubles in call
    order.  Returns (salt[(j,i)...], pepper[(j,i)...])."""
    m = int(28 * 28 * percent)
    d = np.asarray(draws, np.float64)
    assert d.size >= 4 * m
    idx = (d[:4 * m] * 28).astype(np.int64)
    salt = [(int(idx[2 * a + 1]), int(idx[2 * a])) for a in range(m)]
    pepper = [(int(idx[2 * m + 2 * a + 1]), int(idx[2 * m + 2 * a])) for a in range(m)]
    return salt, pepper


def add_salt_pepper_noise(img, percent, draws):
    """preprocess.py:160-183 with the uniform draws supplied by the caller (host RNG)."""
    out = img.copy()
    salt, pepper = salt_pepper_indices(percent, draws)
    for (j, i) in salt:
        out[j, i] = 255
    for (j, i) in pepper:
        out[j, i] = 0
    return out


def clahe(img, clip_limit: float = 40.0, tiles: int = 8):
    """preprocess.py:201-203 cv2.createCLAHE().apply(img) (imgproc/clahe.cpp) for uint8.

    28 is not divisible by 8 -> source is extended with copyMakeBorder(BORDER_REFLECT_101) on the
    bottom/right to 32x32; tile 4x4 (16 px); clipLimit = max(int(40*16/256), 1) = 2;
    lutScale = float(255)/16.  Per tile: histogram, clip to clipLimit, redistribute
    (clipped // 256 to every bin, then the residual one-by-one at stride max(256//residual, 1)),
    cumulative LUT ``saturate_cast<uchar>(sum * lutScale)``.  Per pixel: bilinear blend of the
    four neighbouring tile LUTs with float32 weights taken at ``x*inv_tw - 0.5`` (tile indices
    clamped), result saturate_cast<uchar>(float) (round half even)."""
    H, W = img.shape
    tx = ty = tiles
    if W % tx == 0 and H % ty == 0:
        src = img
    else:
        ph = (ty - H % ty) % ty
        pw = (tx - W % tx) % tx
        ys = _reflect101(np.arange(H + ph), H)
        xs = _reflect101(np.arange(W + pw), W)
        src = img[ys][:, xs]
    th, tw = src.shape[0] // ty, src.shape[1] // tx
    area = th * tw
    clip = 0
    if clip_limit > 0.0:
        clip = int(clip_limit * area / 256)
        clip = max(clip, 1)
    lut_scale = F32(255.0) / F32(area)
    luts = np.zeros((ty, tx, 256), np.uint8)
    for j in range(ty):
        for i in range(tx):
            tile = src[j * th:(j + 1) * th, i * tw:(i + 1) * tw]
            hist = np.bincount(tile.ravel(), minlength=256).astype(np.int64)
            if clip > 0:
                clipped = int(np.maximum(hist - clip, 0).sum())
                hist = np.minimum(hist, clip)
                redist = clipped // 256
                residual = clipped - redist * 256
                hist += redist
                if residual != 0:
                    step = max(256 // residual, 1)
                    k = 0
                    while k < 256 and residual > 0:
                        hist[k] += 1
                        k += step
                        residual -= 1
            s = np.cumsum(hist)
            luts[j, i] = np.clip(np.rint(s.astype(F32) * lut_scale), 0, 255).astype(np.uint8)
    inv_tw = F32(1.0) / F32(tw)
    inv_th = F32(1.0) / F32(th)
    out = np.zeros((H, W), np.uint8)
    for y in range(H):
        tyf = F32(y) * inv_th - F32(0.5)
        ty1 = int(np.floor(tyf))
        ty2 = ty1 + 1
        ya = F32(tyf - F32(ty1))
        ya1 = F32(1.0) - ya
        ty1 = max(ty1, 0)
        ty2 = min(ty2, ty - 1)
        for x in range(W):
            txf = F32(x) * inv_tw - F32(0.5)
            tx1 = int(np.floor(txf))
            tx2 = tx1 + 1
            xa = F32(txf - F32(tx1))
            xa1 = F32(1.0) - xa
            tx1 = max(tx1, 0)
            tx2 = min(tx2, tx - 1)
            v = int(img[y, x])
            res = (F32(luts[ty1, tx1, v]) * xa1 + F32(luts[ty1, tx2, v]) * xa) * ya1 + \
                  (F32(luts[ty2, tx1, v]) * xa1 + F32(luts[ty2, tx2, v]) * xa) * ya
            out[y, x] = np.clip(np.rint(res), 0, 255)
    return out


def nl_denoise_gray(img, h, tw: int = 7, sw: int = 21):
    """preprocess.py:153 cv2.fastNlMeansDenoising(img, None, h, 7, 21) for uint8, 1 channel
    (photo/fast_nlmeans_denoising_invoker.hpp, DistSquared, IT=int, UIT=unsigned, WT=int).

    Border: copyMakeBorder(BORDER_DEFAULT = REFLECT_101) by sw//2 + tw//2 = 13.
    almost_dist -> weight LUT: bin_shift = number of bits so that tw*tw <= 2^shift (49 -> 6);
    almost_dist2actual = 2^shift / (tw*tw); fixed_point_mult = INT_MAX / (sw*sw*255) = 19096
    (integer division); weight[a] = cvRound(fixed_point_mult * exp(-a*a2a / (h*h))), weights under
    0.001*fixed_point_mult are zeroed.  Per pixel: for each of sw*sw
    offsets, dist = sum over the tw*tw patch of squared differences, weight =
    LUT[dist >> shift]; estimate = sum(w * pixel)/sum(w) with rounding divByWeightsSum:
    (estimation + weights_sum/2) / weights_sum  (unsigned)."""
    img = np.asarray(img, np.uint8)
    H, W = img.shape
    thr = tw // 2
    shr = sw // 2
    border = shr + thr
    ys = _reflect101(np.arange(-border, H + border), H)
    xs = _reflect101(np.arange(-border, W + border), W)
    ext = img[ys][:, xs].astype(np.int64)
    tws = tw * tw
    shift = 0
    while (1 << shift) < tws:
        shift += 1
    a2a = float(1 << shift) / float(tws)
    max_dist = 255 * 255 * 1  # DistSquared::maxDist<uchar>() = sampleMax^2 * channels
    almost_max = int(max_dist / a2a + 1)
    fixed_point_mult = int(min(np.iinfo(np.int32).max // (sw * sw * 255), np.iinfo(np.int32).max))
    weight_threshold = 1e-3
    lut = np.zeros(almost_max, np.int64)
    hh = float(h)
    for a in range(almost_max):
        dist = a * a2a
        # DistSquared::calcWeight: exp(-dist / (h*h*channels)); below threshold -> 0
        w = np.exp(-dist / (hh * hh * 1.0)) if hh > 0 else (1.0 if a == 0 else 0.0)
        if np.isnan(w):
            w = 1.0
        wi = int(np.rint(fixed_point_mult * w))
        if wi < weight_threshold * fixed_point_mult:
            wi = 0
        lut[a] = wi
    out = np.zeros((H, W), np.uint8)
    # patch squared-difference sums via box filter of squared differences per search offset
    est = np.zeros((H, W), np.int64)
    wsum = np.zeros((H, W), np.int64)
    for dy in range(-shr, shr + 1):
        for dx in range(-shr, shr + 1):
            a = ext[border - thr:border + H + thr, border - thr:border + W + thr]
            b = ext[border - thr + dy:border + H + thr + dy, border - thr + dx:border + W + thr + dx]
            d2 = (a - b) ** 2
            c = np.cumsum(np.cumsum(np.pad(d2, ((1, 0), (1, 0))), axis=0), axis=1)
            dist = c[tw:, tw:] - c[:-tw, tw:] - c[tw:, :-tw] + c[:-tw, :-tw]
            wgt = lut[dist >> shift]
            pix = ext[border + dy:border + dy + H, border + dx:border + dx + W]
            est += wgt * pix
            wsum += wgt
    # divByWeightsSum for unsigned fixed point: (est + wsum/2) / wsum
    out = ((est + wsum // 2) // wsum).astype(np.uint8)
    return out
