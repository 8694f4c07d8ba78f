"""Generate the committed golden vectors for the preprocess half.

Run HERE (the build container), never on the GPU box:  python tests/golden/make_golden.py
It imports the reference's own module from /root/reference (read-only; every file the
reference functions overwrite is first copied to a temp dir) and the third-party routines the
reference calls (cv2 4.13.0 with IPP off, Pillow 12.2.0), and writes small .npz files next to
this script.  The GPU box has no /root/reference: tests only read the .npz files.

Files written
  pics.npz          the 99 bundled test pictures decoded as the reference decodes them
                    (cv2.imread(p,0) and PIL 'L'), + their tag.json labels
  train_chain.npz   cv2.resize(INTER_CUBIC)->BGR2GRAY->*1/255 on seeded synthetic uploads
  predict_chain.npz PIL convert('L')->resize((20,20)) + construct_inference.py:317-330 outcome
  ops.npz           the 14 ops (cv2 calls of preprocess.py:33-227) on 24 pictures
  ref_files.npz     Tier B: the reference's *file-based* functions run unchanged on temp copies
                    (JPEG round trips included) -- compared with a tolerance, not bit-exact
"""
import json
import os
import shutil
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = "/root/reference"

import cv2  # noqa: E402
from PIL import Image  # noqa: E402

cv2.ipp.setUseIPP(False)
from oracle import cvref  # noqa: E402


def synth_upload(seed, h, w, natural):
    rng = np.random.default_rng(seed)
    a = rng.integers(0, 256, (h, w, 3), np.uint8)
    if natural:
        a = cv2.GaussianBlur(a, (0, 0), 3)
    return a


TRAIN_CASES = [  # (seed, h, w, natural)
    (100, 29, 29, False), (101, 37, 53, False), (102, 100, 80, True), (103, 255, 257, False),
    (104, 480, 640, True), (105, 768, 1024, False), (106, 768, 1024, True), (107, 31, 400, False),
    (108, 333, 30, True), (109, 28, 28, False), (110, 56, 56, False), (111, 1080, 1920, True),
]

OP_CASES = [
    ("flip_up_down", None, None), ("flip_left_right", None, None), ("transpose_image", None, None),
    ("adjust_brightness_contrast", 1.5, 30), ("adjust_brightness_contrast", 0.5, -20),
    ("adjust_brightness_contrast", 1.2, 10.5), ("adjust_brightness_contrast", 2, 3),
    ("mean_filter", 2, None), ("mean_filter", 3, None), ("mean_filter", 5, None),
    ("gaussian_blur", 3, None), ("gaussian_blur", 5, None), ("gaussian_blur", 9, None), ("gaussian_blur", 11, None),
    ("median_filter", 3, None), ("median_filter", 5, None),
    ("nl_denoise_gray", 3, None), ("nl_denoise_gray", 10, None),
    ("add_salt_pepper_noise", 0.05, None), ("add_salt_pepper_noise", 0.3, None),
    ("equalize_hist", None, None), ("clahe", None, None),
    ("erode", 2, None), ("erode", 3, None), ("dilate", 2, None), ("dilate", 3, None), ("dilate", 5, None),
]


def main():
    pic_dir = os.path.join(REF, "test-data", "test-pics")
    tags = json.load(open(os.path.join(REF, "test-data", "tag.json")))
    names = sorted(tags.keys(), key=lambda s: int(s[4:-4]))
    gray_cv = np.stack([cv2.imread(os.path.join(pic_dir, n), 0) for n in names])
    gray_pil = np.stack([np.asarray(Image.open(os.path.join(pic_dir, n)).convert("L")) for n in names])
    labels = np.array([int(tags[n]) for n in names], np.int32)
    np.savez_compressed(os.path.join(HERE, "pics.npz"), names=np.array(names), gray_cv=gray_cv,
                        gray_pil=gray_pil, labels=labels)

    # ---- train-side chain
    out = {"cases": np.array(TRAIN_CASES, np.int64)}
    for (seed, h, w, nat) in TRAIN_CASES:
        img = synth_upload(seed, h, w, nat)
        g, f = cvref.train_chain(img)
        out[f"u8_{seed}"] = g
        out[f"f32_{seed}"] = f
        out[f"insum_{seed}"] = np.array([int(img.astype(np.uint64).sum())], np.uint64)
    np.savez_compressed(os.path.join(HERE, "train_chain.npz"), **out)

    # ---- predict-side chain: bundled pictures + synthetic RGB uploads
    small = np.stack([cvref.predict_small20(g) for g in gray_pil])
    status = np.zeros(len(names), np.int32)
    outs_c = np.zeros((len(names), 784), np.float32)
    outs_n = np.zeros((len(names), 784), np.float32)
    for i, g in enumerate(gray_pil):
        try:
            outs_c[i] = cvref.predict_chain(g, True)
        except IndexError:
            status[i] = 1
        outs_n[i] = cvref.predict_chain(g, False)
    pc = {"small20": small, "compat_indexerror": status, "out_compat": outs_c, "out_intended": outs_n}
    rgb_cases = [(200, 64, 48), (201, 300, 200), (202, 768, 1024), (203, 21, 977), (204, 500, 23)]
    pc["rgb_cases"] = np.array(rgb_cases, np.int64)
    for (seed, h, w) in rgb_cases:
        img = synth_upload(seed, h, w, True)
        pc[f"small20_{seed}"] = cvref.predict_small20(img)
        pc[f"out_{seed}"] = cvref.predict_chain(img, False)
    np.savez_compressed(os.path.join(HERE, "predict_chain.npz"), **pc)

    # ---- the 14 ops on 24 pictures (+ 4 synthetic noise images)
    rng = np.random.default_rng(300)
    imgs = np.concatenate([gray_cv[:24], rng.integers(0, 256, (4, 28, 28), np.uint8)])
    ops = {"inputs": imgs, "n_cases": np.array([len(OP_CASES)])}
    for ci, (name, v1, v2) in enumerate(OP_CASES):
        res = []
        draws_all = []
        for k, im in enumerate(imgs):
            draws = None
            if name == "add_salt_pepper_noise":
                m = int(28 * 28 * v1)
                draws = np.random.default_rng(1000 + ci * 100 + k).random(4 * m)
                draws_all.append(draws)
            res.append(cvref.op(name, im, v1, v2, draws))
        ops[f"case{ci}_name"] = np.array(name)
        ops[f"case{ci}_v1"] = np.array(np.nan if v1 is None else v1, np.float64)
        ops[f"case{ci}_v2"] = np.array(np.nan if v2 is None else v2, np.float64)
        ops[f"case{ci}_v1_is_int"] = np.array(isinstance(v1, int))
        ops[f"case{ci}_v2_is_int"] = np.array(isinstance(v2, int))
        ops[f"case{ci}_out"] = np.stack(res)
        if draws_all:
            ops[f"case{ci}_draws"] = np.stack(draws_all)
    np.savez_compressed(os.path.join(HERE, "ops.npz"), **ops)

    # ---- Tier B: the reference's file-based functions, unchanged, on temp copies
    sys.path.insert(0, REF)
    import apps.preprocess.preprocess as refpp  # the reference module itself

    tmp = tempfile.mkdtemp(prefix="golden_ref_")
    try:
        tb = {}
        sel = names[:12]
        for n in sel:
            shutil.copy(os.path.join(pic_dir, n), os.path.join(tmp, n))
        resized = []
        for n in sel:
            p = os.path.join(tmp, n)
            refpp.resize(p)                       # preprocess.py:18-30
            resized.append(cv2.imread(p, 0))
        tb["names"] = np.array(sel)
        tb["resize_then_gray"] = np.stack(resized)
        for fname, v1, v2 in [("flip_up_down", None, None), ("mean_filter", 3, None), ("erode", 3, None),
                              ("equalize_hist", None, None), ("adjust_brightness_contrast", 1.5, 30)]:
            outs = []
            for n in sel:
                p = os.path.join(tmp, n)
                getattr(refpp, fname)(p, True, v1, v2)      # overlap=True -> writes *_copy.jpg
                outs.append(cv2.imread(refpp.copied_name(p), 0))
            tb[f"file_{fname}"] = np.stack(outs)
        # one large synthetic upload through the real resize() (JPEG in, JPEG out)
        big = synth_upload(106, 768, 1024, True)
        pbig = os.path.join(tmp, "big.jpg")
        cv2.imwrite(pbig, big)
        decoded = cv2.imread(pbig)
        refpp.resize(pbig)
        tb["big_decoded_train_u8"] = cvref.train_chain(decoded)[0]   # in-memory chain on the decoded pixels
        tb["big_file_gray"] = cv2.imread(pbig, 0)                     # what the reference leaves on disk
        np.savez_compressed(os.path.join(HERE, "ref_files.npz"), **tb)
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)))


if __name__ == "__main__":
    main()
