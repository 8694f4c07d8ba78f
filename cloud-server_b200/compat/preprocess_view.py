"""In-process stand-in for PreprocessView.execute (apps/preprocess/views.py:95-137), the caller of the
preprocess path: same inputs (one operation dict, the model's data directory, the
tag.json location), same effects on disk (every picture resized to 28x28 and re-encoded, the op's result
written over the picture or to `<name>_copy<ext>`, tag.json extended with the copies) -- but the pictures
go through the GPU in ONE batch per step instead of one file at a time.

Byte-compatible with looping the per-file functions: the JPEG write + re-read that the reference performs
between `resize` and the op (preprocess.py:30 then :40) is reproduced with cv2.imencode / imdecode."""
import json
import os
import random

import cv2
import numpy as np
import torch

from cloud_server_b200.context import default_context
from cloud_server_b200.compat.apps.preprocess import preprocess as pp

op_map = {                                                       # views.py:15-40
    "上下翻转": "flip_up_down", "左右翻转": "flip_left_right", "对角线翻转": "transpose_image",
    "对比度亮度调整": "adjust_brightness_contrast", "随机对比度亮度调整": "random_brightness_contrast",
    "直方图均衡化": "equalize_hist", "CLAHE均衡化": "clahe", "腐蚀": "erode", "膨胀": "dilate",
    "均值滤波": "mean_filter", "高斯模糊": "gaussian_blur", "中值滤波": "median_filter",
    "灰度非局部平均值去噪": "nl_denoise_gray", "添加椒盐噪声": "add_salt_pepper_noise",
}
_OP_ID = {"flip_up_down": 0, "flip_left_right": 1, "transpose_image": 2, "adjust_brightness_contrast": 3,
          "random_brightness_contrast": 3, "mean_filter": 4, "gaussian_blur": 5, "median_filter": 6, "nl_denoise_gray": 7,
          "add_salt_pepper_noise": 8, "equalize_hist": 9, "clahe": 10, "erode": 11, "dilate": 12}


def execute(operation, data_dir, tag_location, ctx=None):
    """operation: {"operationName": <UI name>, "overlap": bool, "value1": .., "value2": ..} (views.py:97-104)."""
    name = op_map.get(operation['operationName'])
    if name is None or not hasattr(pp, name):
        return                                                    # views.py:109: unknown names are skipped
    ctx = ctx or default_context()
    with open(tag_location, 'r') as f:
        tags = json.load(f)
    tags_edit = tags.copy()
    overlap = operation.get('overlap') is True or operation.get('overlap') == 'true'
    v1, v2 = operation.get('value1'), operation.get('value2')
    rels = list(tags.keys())
    paths = [data_dir + r for r in rels]
    # ---- resize every picture (preprocess.resize): decode on the host, group by shape, one launch per group
    decoded = [cv2.imread(p) for p in paths]
    groups = {}
    for i, im in enumerate(decoded):
        groups.setdefault(im.shape, []).append(i)
    resized = [None] * len(paths)
    for shape, idx in groups.items():
        batch = torch.from_numpy(np.stack([decoded[i] for i in idx])).to(ctx.tdev)
        out = ctx.resize_cubic(batch).cpu().numpy()
        for k, i in enumerate(idx):
            resized[i] = out[k]
    gray = []
    for p, im in zip(paths, resized):
        ext = os.path.splitext(p)[1]
        ok, buf = cv2.imencode(ext, im)                           # what cv2.imwrite(dir, img) puts on disk (:30)
        buf.tofile(p)
        gray.append(cv2.imdecode(buf, 0))                          # what cv2.imread(dir, 0) then reads (:40)
    # ---- the op, one launch for the whole data set; host-side random draws in the reference's order
    g = torch.from_numpy(np.stack(gray)).to(ctx.tdev)
    n = len(paths)
    if name == "random_brightness_contrast":
        outs = []
        for i in range(n):                                        # alpha/beta are drawn per picture (:106-107)
            alpha = random.uniform(0, v1)
            beta = random.randint(0 - v2, v2)
            outs.append(ctx.op(3, g[i:i + 1], alpha, beta))
        res = torch.cat(outs)
    elif name == "add_salt_pepper_noise":
        m = max(int(28 * 28 * v1), 0)
        draws = np.array([[np.random.random() for _ in range(4 * m)] for _ in range(n)], np.float64).reshape(n, 4 * m)
        res = ctx.op(8, g, v1, None, draws if m > 0 else None)
    else:
        res = ctx.op(_OP_ID[name], g, v1, v2)
    res = res.cpu().numpy()
    for rel, p, im in zip(rels, paths, res):
        cv2.imwrite(pp.copied_name(p) if overlap else p, im)
        if overlap:
            parts = os.path.splitext(rel)
            tags_edit[parts[0] + '_copy' + parts[1]] = tags[rel]   # views.py:129-132
    with open(tag_location, 'w') as f:
        f.write(json.dumps(tags_edit))
