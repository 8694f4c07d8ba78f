#!/usr/bin/env python3
"""Drop-in for apps/construction/util/construct_inference.py: same command line (cmd.py:84:
--config --filename, plus --model_name --save_path --data_dir --result, :17-23), same result.json
(:343-347).  PIL decodes on the host as there; the predict-side preprocess chain (:312-330) and the
forward pass + argmax (:332-342) run in the CUDA library.

compat (default): the reference's `arr[bizar+i][bizar+j] = 220` statement (:325-326) raises
IndexError for most pictures; so does this script.  --intended skips that statement."""
import argparse
import json
import os
import sys

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(_HERE)))

from cloud_server_b200.compat.construct_distribute import get_model, is_already_save, load_checkpoint  # noqa: E402
from cloud_server_b200.compat.graph import get_net, placeholder, compile_net  # noqa: E402


def build_flags(argv=None):
    ap = argparse.ArgumentParser(allow_abbrev=False)
    ap.add_argument("--model_name", default="model")
    ap.add_argument("--save_path", default="./train_model/")
    ap.add_argument("--data_dir", default="data")
    ap.add_argument("--filename", default="")
    ap.add_argument("--result", default="result.json")
    ap.add_argument("--config", default="")
    ap.add_argument("--intended", action="store_true", help="(extension) skip the out-of-range write of :325-326")
    return ap.parse_args(argv)


def main(argv=None):
    FLAGS = build_flags(argv)
    data = {"result": "fail", "message": "no model please train a model first"}
    config = json.loads(FLAGS.config)
    net_type = config["net_type"]
    net_config = config["net_config"]
    save_path = FLAGS.save_path + FLAGS.model_name + ".ckpt-" + str(get_model(FLAGS.save_path, FLAGS.model_name))
    if not is_already_save(save_path):
        # the reference opens result_path with mode 'file' here (:308), which raises before anything is written
        raise ValueError("invalid mode: 'file'")
    import torch
    from PIL import Image
    from cloud_server_b200.context import Context

    ctx = Context(int(os.environ.get("LOCAL_RANK", "0")))
    im = np.asarray(Image.open(FLAGS.filename))                   # :312 (convert('L') happens on the GPU)
    if im.ndim == 3 and im.shape[2] == 4:
        im = im[:, :, :3]
    src = torch.from_numpy(np.array(im)[None]).to(ctx.tdev)
    arr, status, small = ctx.predict_chain(src, channel_order="RGB", compat=not FLAGS.intended, want_small20=True)
    print(tuple(small[0].shape))                                  # print(arr.shape), :317
    if int(status[0].item()) != 0:
        raise IndexError("index %d is out of bounds for axis 0 with size 20" % 20)    # :326
    net = compile_net(get_net(net_type, net_config, placeholder(28, 28, 1), None), ctx)
    load_checkpoint(net, save_path)
    ans = int(net.predict(arr)[0].item())
    data["result"] = "success"
    data["message"] = str(ans)
    print(data)
    with open(FLAGS.result, 'w') as file:
        json.dump(data, file)
    return 0


if __name__ == "__main__":
    sys.exit(main())
