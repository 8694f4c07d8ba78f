"""Drop-in for the reference module apps/preprocess/preprocess.py: the same 16 names, the same
positional signatures ``f(dir)`` / ``f(dir, overlap, value1, value2)``, the same file side effects
(``dir`` overwritten, or ``copied_name(dir)`` written when ``overlap`` is true).  JPEG decode and
encode stay on the host with cv2 exactly as in the reference; every pixel operation runs in the
CUDA library (nc_pp_resize_cubic / nc_pp_op).  PreprocessView.execute (apps/preprocess/views.py:
95-137) can import this module unchanged.

For whole datasets use the batched surface instead (cloud_server_b200.context.Context): one launch
per operation for all pictures, no per-file codec round trips."""
import os
import random

import cv2
import numpy as np
import torch

from cloud_server_b200.context import default_context
from cloud_server_b200._lib import NcError

(_FLIP_UD, _FLIP_LR, _TRANSPOSE, _BRIGHT, _MEAN, _GAUSS, _MEDIAN, _NLM, _SALT, _EQ, _CLAHE, _ERODE, _DILATE) = range(13)


def copied_name(name):                                           # preprocess.py:8-15
    parts = os.path.splitext(name)
    return parts[0] + '_copy' + parts[1]


def _raise_like_reference(e):
    msg = str(e)
    for cls in (OverflowError, IndexError):
        if cls.__name__ + ":" in msg:
            raise cls(msg.split(cls.__name__ + ":", 1)[1].strip()) from None
    raise e


def _run(op_id, dir, overlap, v1=None, v2=None, draws=None):
    img = cv2.imread(dir, 0)                                      # e.g. preprocess.py:40
    if img is None:
        raise AttributeError("'NoneType' object has no attribute 'shape'")   # what the reference hits on a bad path
    if img.shape != (28, 28):
        raise NotImplementedError("the B200 ops take 28x28 pictures: call resize(dir) first, as PreprocessView.execute does")
    ctx = default_context()
    d = torch.from_numpy(img[None]).to(ctx.tdev)
    try:
        out = ctx.op(op_id, d, v1, v2, draws)
    except (ValueError, NcError) as e:
        _raise_like_reference(e)
    new_img = out[0].cpu().numpy()
    if not overlap:
        cv2.imwrite(dir, new_img)
    else:
        cv2.imwrite(copied_name(dir), new_img)


def resize(dir):                                                  # preprocess.py:18-30
    print('resizing image: ' + dir)
    im1 = cv2.imread(dir)
    ctx = default_context()
    img = ctx.resize_cubic(torch.from_numpy(im1[None]).to(ctx.tdev))[0].cpu().numpy()
    cv2.imwrite(dir, img)


def flip_up_down(dir, overlap, value1, value2):                   # preprocess.py:33-45
    _run(_FLIP_UD, dir, overlap)


def flip_left_right(dir, overlap, value1, value2):                # preprocess.py:48-60
    _run(_FLIP_LR, dir, overlap)


def transpose_image(dir, overlap, value1, value2):                # preprocess.py:63-75 (a 180 degree rotation)
    _run(_TRANSPOSE, dir, overlap)


def adjust_brightness_contrast(dir, overlap, alpha, beta):        # preprocess.py:78-99
    _run(_BRIGHT, dir, overlap, alpha, beta)


def random_brightness_contrast(dir, overlap, max_alpha, max_beta):  # preprocess.py:102-120
    alpha = random.uniform(0, max_alpha)                            # host RNG, same calls, same order
    beta = random.randint(0 - max_beta, max_beta)
    _run(_BRIGHT, dir, overlap, alpha, beta)


def mean_filter(dir, overlap, height, value2):                    # preprocess.py:123-129
    _run(_MEAN, dir, overlap, height)


def gaussian_blur(dir, overlap, height, value2):                  # preprocess.py:132-138
    _run(_GAUSS, dir, overlap, height)


def median_filter(dir, overlap, height, value2):                  # preprocess.py:141-148
    print(overlap)
    _run(_MEDIAN, dir, overlap, height)


def nl_denoise_gray(dir, overlap, h, value2):                     # preprocess.py:151-157
    _run(_NLM, dir, overlap, h)


def add_salt_pepper_noise(dir, overlap, percent, value2):         # preprocess.py:160-188
    m = int(28 * 28 * percent)
    draws = np.array([np.random.random() for _ in range(4 * max(m, 0))], np.float64)   # global MT19937, call order kept
    _run(_SALT, dir, overlap, percent, None, draws if m > 0 else None)


def equalize_hist(dir, overlap, value1, value2):                  # preprocess.py:191-197
    _run(_EQ, dir, overlap)


def clahe(dir, overlap, value1, value2):                          # preprocess.py:200-207
    _run(_CLAHE, dir, overlap)


def erode(dir, overlap, kernel_value, value2):                    # preprocess.py:210-217
    _run(_ERODE, dir, overlap, kernel_value)


def dilate(dir, overlap, kernel_value, value2):                   # preprocess.py:220-227
    _run(_DILATE, dir, overlap, kernel_value)
