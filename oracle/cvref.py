"""Live third-party calls the reference makes, composed in memory (TEST INFRASTRUCTURE ONLY).

These are the *same library calls* as apps/preprocess/preprocess.py and
apps/construction/util/construct_inference.py:312-330, minus the JPEG file round trip:
they pin ``oracle/pp_restate.py`` and serve as the timed CPU baseline in ``bench.py``.
cv2 runs with IPP disabled (``cv2.ipp.setUseIPP(False)``): that selects OpenCV's own
open-source arithmetic, the canonical target named in DESIGN.md.
"""
from __future__ import annotations

import numpy as np

try:  # the GPU box has the same image; gate anyway
    import cv2

    cv2.ipp.setUseIPP(False)
    # OpenCV's useIPP flag is thread-local and its row-stripe parallel_for_ is not bit-reproducible
    # across thread counts (observed: GaussianBlur k=5 on a 28x28 picture differs in 13 pixels with a
    # 24-thread pool, never with 1 thread).  The reference ran on a 1-vCPU node (API.md:428-431), so
    # the canonical, deterministic target is single-threaded OpenCV.
    cv2.setNumThreads(1)
except Exception:  # pragma: no cover
    cv2 = None
try:
    from PIL import Image
except Exception:  # pragma: no cover
    Image = None


def have_cv2():
    return cv2 is not None


def have_pil():
    return Image is not None


def train_chain(img_bgr: np.ndarray):
    """preprocess.py:28-29 resize, :40 gray, construct_distribute.py:503-504 normalise."""
    small = cv2.resize(img_bgr, (28, 28), interpolation=cv2.INTER_CUBIC)
    gray = cv2.cvtColor(small, cv2.COLOR_BGR2GRAY) if small.ndim == 3 else small
    f = np.multiply(gray.reshape(-1).astype(np.float32), np.float32(1.0 / 255.0))
    return gray, f


def predict_small20(img: np.ndarray) -> np.ndarray:
    """construct_inference.py:312-316: convert('L') then resize((20,20)) (real Pillow calls)."""
    im = Image.fromarray(img).convert("L")
    im = im.resize((20, 20))
    return np.asarray(im)


def predict_chain(img: np.ndarray, compat: bool = True) -> np.ndarray:
    """construct_inference.py:312-330 (loop body kept statement-for-statement)."""
    line = 20
    bizar = int((28 - line) / 2)
    arr = np.array(predict_small20(img), dtype="float32")
    hist = np.zeros((28, 28))
    for i in range(0, line):
        for j in range(0, line):
            if arr[i][j] > 150:
                hist[i + bizar][bizar + j] = 254
            elif arr[i][j] > 100 and compat:
                arr[bizar + i][bizar + j] = 220
    arr = hist
    arr /= 255.0
    return arr.flatten().astype(np.float32)


def op(name: str, img: np.ndarray, v1=None, v2=None, draws=None) -> np.ndarray:
    """The cv2/numpy call each of the 14 ops makes (preprocess.py:33-227), in memory."""
    img = np.ascontiguousarray(img)
    if name == "flip_up_down":
        return cv2.flip(img, 0)
    if name == "flip_left_right":
        return cv2.flip(img, 1)
    if name == "transpose_image":
        return cv2.flip(img, -1)
    if name == "adjust_brightness_contrast":
        out = img.copy()
        h, w = out.shape
        with np.errstate(all="ignore"):
            for xi in range(w):
                for xj in range(h):
                    out[xj, xi] = out[xj, xi] * v1 + v2
                    if out[xj, xi] > 255:
                        out[xj, xi] = 255
                    elif out[xj, xi] < 0:
                        out[xj, xi] = 0
        return out
    if name == "mean_filter":
        return cv2.blur(img, (v1, v1))
    if name == "gaussian_blur":
        return cv2.GaussianBlur(img, (v1, v1), 0)
    if name == "median_filter":
        return cv2.medianBlur(img, v1, 0)
    if name == "nl_denoise_gray":
        return cv2.fastNlMeansDenoising(img, None, v1, 7, 21)
    if name == "add_salt_pepper_noise":
        out = img.copy()
        m = int(28 * 28 * v1)
        it = iter(draws)
        for _ in range(m):
            i = int(next(it) * 28)
            j = int(next(it) * 28)
            out[j, i] = 255
        for _ in range(m):
            i = int(next(it) * 28)
            j = int(next(it) * 28)
            out[j, i] = 0
        return out
    if name == "equalize_hist":
        return cv2.equalizeHist(img)
    if name == "clahe":
        return cv2.createCLAHE().apply(img)
    if name == "erode":
        return cv2.erode(img, np.ones((v1, v1), np.uint8))
    if name == "dilate":
        return cv2.dilate(img, np.ones((v1, v1), np.uint8))
    raise KeyError(name)
