"""Drop-in surface of the reference: names, signatures, flags, file formats.  CPU part:
host logic only.  GPU part (marked): the shims run end to end on temp copies of bundled pictures."""
import inspect
import json
import os
import re

import numpy as np
import pytest

from conftest import golden
from oracle import tfref

REF_NAMES = ["copied_name", "resize", "flip_up_down", "flip_left_right", "transpose_image", "adjust_brightness_contrast",
             "random_brightness_contrast", "mean_filter", "gaussian_blur", "median_filter", "nl_denoise_gray",
             "add_salt_pepper_noise", "equalize_hist", "clahe", "erode", "dilate"]


def test_preprocess_module_surface():
    from cloud_server_b200.compat.apps.preprocess import preprocess as pp
    for n in REF_NAMES:
        assert hasattr(pp, n), n
    assert list(inspect.signature(pp.resize).parameters) == ["dir"]
    for n in REF_NAMES[2:]:
        assert len(inspect.signature(getattr(pp, n)).parameters) == 4, n
    assert pp.copied_name("/a/b/test1.jpg") == "/a/b/test1_copy.jpg"


def test_train_script_flags_match_reference_command_line():
    from cloud_server_b200.compat import construct_distribute as cd
    # the exact string cmd.py:63-69 builds
    cfg = '{"iter":"1000","learning_rate":"0.01","loss_name":"entropy","optimizer_name":"GradientDescentOptimizer",' \
          '"net_type":"CNN","net_config":{"middle_layer":[{"layer":"conv","filter":[2,2,10]}],"output_layer":{}}}'
    argv = ["--mode=train", "--ps_hosts=127.0.0.1:2222", "--worker_hosts=127.0.0.1:2223", "--job_name=ps", "--task_index=0",
            "--config=" + cfg, "--ratio=0.800000"]
    f = cd.build_flags(argv)
    assert f.job_name == "ps" and f.ratio == "0.800000" and json.loads(f.config)["iter"] == "1000"
    assert cd.main(argv) == 0                      # the ps job is a no-op that returns
    f2 = cd.build_flags([])
    assert (f2.model_name, f2.save_path, f2.result, f2.iter) == ("model", "./train_model/", "./result.txt", 2000)
    from cloud_server_b200.compat import construct_inference as ci
    g = ci.build_flags(["--config=" + cfg, "--filename=infer/x.jpg"])
    assert (g.model_name, g.save_path, g.result) == ("model", "./train_model/", "result.json")


def test_dataset_matches_oracle_dataset():
    from cloud_server_b200.compat.construct_distribute import DataSet
    imgs = (np.arange(23 * 784).reshape(23, 28, 28, 1) % 251).astype(np.uint8)
    labs = np.eye(10)[np.arange(23) % 10]
    np.random.seed(5)
    a = DataSet(imgs, labs)
    b = tfref.DataSet(imgs, labs, rs=np.random.RandomState(5))
    for _ in range(9):
        xa, ya = a.next_batch(5)
        xb, yb = b.next_batch(5)
        assert np.array_equal(xa, xb) and np.array_equal(ya, yb)
    assert a.epochs_completed == 1 and a.num_examples == 23


def test_graph_builder_equals_json_interpreter():
    from cloud_server_b200.compat import graph as G
    x = G.placeholder()
    h = G.conv2d(x, [5, 5, 32], isBias=True)
    h = G.active(h, "relu")
    h = G.max_pool(h)
    h = G.connect_layer(h, 64)
    cfg = G.net_config_of(h)
    layers, params = tfref.parse_layers(cfg)
    assert [l["kind"] for l in layers] == ["conv", "active", "pool", "connect", "connect"]
    assert h.get_shape() == (None, 64)
    via_json = G.cnn(tfref.API_EXAMPLE, G.placeholder())
    l2, p2 = tfref.parse_layers(G.net_config_of(via_json))
    l3, p3 = tfref.parse_layers(tfref.API_EXAMPLE)
    assert [p[1] for p in p2] == [p[1] for p in p3]
    with pytest.raises(AssertionError):
        G.active(x, "tanh")
    assert G.loss_function("hinge", None) is None
    assert G.optimizer_function("GradientDescentOptimizer", "0.01") == ("sgd", 0.01)
    assert G.optimizer_function("AdamOptimizer", 0.5)[0] == "adagrad"      # the reference's else-branch (:310-311)


# ------------------------------------------------------------------------------------------ GPU
@pytest.mark.gpu
def test_preprocess_shim_files(ctx, tmp_path):
    import cv2
    from cloud_server_b200.compat.apps.preprocess import preprocess as pp
    from oracle import cvref
    pics = golden("pics.npz")
    big = np.random.default_rng(1).integers(0, 256, (96, 128, 3), np.uint8)
    p = str(tmp_path / "u.jpg")
    cv2.imwrite(p, big)
    decoded = cv2.imread(p)
    pp.resize(p)
    want = cv2.resize(decoded, (28, 28), interpolation=cv2.INTER_CUBIC)
    q = str(tmp_path / "w.jpg")
    cv2.imwrite(q, want)
    assert np.array_equal(cv2.imread(p), cv2.imread(q))            # same pixels into the same encoder
    for name, v1, v2 in [("flip_up_down", None, None), ("mean_filter", 3, None), ("erode", 3, None), ("equalize_hist", None, None),
                         ("adjust_brightness_contrast", 1.5, 30), ("clahe", None, None), ("gaussian_blur", 5, None)]:
        f = str(tmp_path / ("%s.jpg" % name))
        cv2.imwrite(f, pics["gray_cv"][3])
        src = cv2.imread(f, 0)
        getattr(pp, name)(f, True, v1, v2)
        assert os.path.exists(pp.copied_name(f))
        ref = str(tmp_path / "ref.jpg")
        want_arr = cvref.op(name, src, v1, v2)
        cv2.imwrite(ref, want_arr)
        from oracle import pp_restate as R
        ora = getattr(R, name)(src) if v1 is None else (getattr(R, name)(src, v1) if v2 is None else getattr(R, name)(src, v1, v2))
        got, want = cv2.imread(pp.copied_name(f), 0), cv2.imread(ref, 0)
        assert np.array_equal(got, want), (name, "file diff", int((got != want).sum()), "cv2 vs oracle", int((want_arr != ora).sum()),
                                           "threads", cv2.getNumThreads())
        assert np.array_equal(cv2.imread(f, 0), src)               # overlap=True leaves the original alone
    with pytest.raises(OverflowError):
        pp.adjust_brightness_contrast(f, False, 300, 0)


@pytest.mark.gpu
def test_train_and_inference_scripts_end_to_end(ctx, tmp_path, monkeypatch):
    import cv2
    from cloud_server_b200.compat import construct_distribute as cd, construct_inference as ci
    pics = golden("pics.npz")
    monkeypatch.chdir(tmp_path)
    os.mkdir("data")
    tags = {}
    for i, (name, lab) in enumerate(zip(pics["names"], pics["labels"])):
        cv2.imwrite(os.path.join("data", str(name)), pics["gray_cv"][i])
        tags[str(name)] = str(int(lab))
    json.dump(tags, open("tag.json", "w"))
    config = {"iter": "120", "learning_rate": "0.01", "loss_name": "entropy", "optimizer_name": "GradientDescentOptimizer",
              "net_type": "CNN", "net_config": {"middle_layer": [{"layer": "conv", "filter": [3, 3, 8], "isBias": "True"},
                                                                 {"layer": "active", "active_func": "relu"}, {"layer": "pool"},
                                                                 {"layer": "connect", "hidden": 64}, {"layer": "active", "active_func": "relu"}],
                                        "output_layer": {}}}
    argv = ["--mode=train", "--ps_hosts=a:1", "--worker_hosts=b:2", "--job_name=worker", "--task_index=0",
            "--config=" + json.dumps(config), "--ratio=0.800000", "--data_dir=.", "--save_path=./train_model/",
            "--honor_config", "--seed=3"]
    assert cd.main(argv) == 0
    # result.txt parses with the consumer's own logic (apps/runtime/views.py:55-70)
    lines = open("result.txt").read().split("\n")
    steps = [l for l in lines if l.startswith("step:")]
    assert len(steps) == 2 and re.fullmatch(r"step:\d+,accuracy:\d\.\d{6},duration:\d+\.\d{6}", steps[0])
    assert steps[0].startswith("step:0,") and steps[1].startswith("step:100,")
    final = [l for l in lines if l.startswith("final_accuracy:")]
    assert len(final) == 1 and 0.0 <= float(final[0].split(":")[1]) <= 1.0
    assert cd.get_model("./train_model/", "model") == 120
    assert os.path.exists("./train_model/model.ckpt-120.meta")
    # SGD at 0.01 for 120 steps on 79 pictures learns something: accuracy on the train split rises
    acc0 = float(steps[0].split(",")[1].split(":")[1])
    acc1 = float(steps[1].split(",")[1].split(":")[1])
    assert acc1 >= acc0
    # inference script: compat mode raises where the reference raises, --intended predicts a digit
    name = str(pics["names"][0])
    iargv = ["--config=" + json.dumps(config), "--filename=" + os.path.join("data", name)]
    from PIL import Image
    from oracle import pp_restate as R
    try:
        R.predict_chain(np.asarray(Image.open(os.path.join("data", name))), compat=True)
        ref_raises = False
    except IndexError:
        ref_raises = True
    if ref_raises:
        with pytest.raises(IndexError):
            ci.main(iargv)
    else:
        assert ci.main(iargv) == 0
    assert ci.main(iargv + ["--intended"]) == 0
    out = json.load(open("result.json"))
    assert out["result"] == "success" and out["message"] in list("0123456789")
    with pytest.raises(ValueError):
        ci.main(["--config=" + json.dumps(config), "--filename=x.jpg", "--save_path=./nowhere/"] if os.path.isdir("./nowhere/") else
                ["--config=" + json.dumps(config), "--filename=x.jpg", "--model_name=absent"])


@pytest.mark.gpu
def test_device_dataset_matches_reference_dataset(ctx):
    """DeviceDataSet.next_batch (HBM-resident pictures, gather kernel) == the reference DataSet semantics
    (oracle restatement of construct_distribute.py:526-569) batch for batch, under the same RNG state."""
    import torch
    from cloud_server_b200.dataset import DeviceDataSet
    rng = np.random.default_rng(0)
    imgs = rng.integers(0, 256, (203, 28, 28, 1), np.uint8)
    labs = np.eye(10, dtype=np.float32)[rng.integers(0, 10, 203)]
    np.random.seed(11)
    dev_ds = DeviceDataSet(imgs, labs, ctx)
    ref = tfref.DataSet(imgs, labs, rs=np.random.RandomState(11))
    for step in range(12):
        xb, yb = dev_ds.next_batch(50)
        xr, yr = ref.next_batch(50)
        assert np.array_equal(xb.cpu().numpy(), xr), step
        assert np.array_equal(yb.cpu().numpy(), yr), step
    assert dev_ds.epochs_completed == 2


@pytest.mark.gpu
def test_gather_rows_out_of_range_index_is_reported(ctx):
    """`self._images[perm]` (construct_distribute.py:541) raises IndexError for an index outside the set; the gather
    kernel zero-fills that row and nc_gather_rows_status reports (and clears) the sticky flag."""
    import ctypes as C
    import torch
    from cloud_server_b200 import _lib
    L = _lib.lib()
    src = torch.arange(5 * 8, dtype=torch.float32, device="cuda").reshape(5, 8)
    idx = torch.tensor([4, 0, 5, -1, 2], dtype=torch.int64, device="cuda")
    out = torch.full((5, 8), -7.0, device="cuda")
    _lib.check(ctx.handle, L.nc_gather_rows(ctx.handle, C.c_void_p(src.data_ptr()), 5, 8, C.c_void_p(idx.data_ptr()), 5,
                                            C.c_void_p(out.data_ptr()), ctx._stream()))
    with pytest.raises(IndexError):
        _lib.check(ctx.handle, L.nc_gather_rows_status(ctx.handle, ctx._stream()))
    assert torch.equal(out[[0, 1, 4]], src[[4, 0, 2]]) and float(out[[2, 3]].abs().max()) == 0.0
    _lib.check(ctx.handle, L.nc_gather_rows_status(ctx.handle, ctx._stream()))       # cleared by the read


@pytest.mark.gpu
def test_batched_view_adapter_equals_per_file_loop(ctx, tmp_path):
    """compat.preprocess_view.execute (one batch per step) leaves the same files and tag.json as the reference's
    per-file loop (resize(path); op(path, overlap, v1, v2) for every tag entry, views.py:118-132)."""
    import shutil
    import cv2
    from cloud_server_b200.compat import preprocess_view as pv
    from cloud_server_b200.compat.apps.preprocess import preprocess as pp
    pics = golden("pics.npz")
    rng = np.random.default_rng(4)
    for d in ("a", "b"):
        os.mkdir(tmp_path / d)
    tags = {}
    for i in range(6):
        name = "p%d.jpg" % i
        img = rng.integers(0, 256, (40 + 8 * (i % 2), 60, 3), np.uint8) if i < 4 else np.repeat(pics["gray_cv"][i][..., None], 3, 2)
        cv2.imwrite(str(tmp_path / "a" / name), img)
        tags[name] = str(i)
    for name in tags:
        shutil.copy(tmp_path / "a" / name, tmp_path / "b" / name)
    for d in ("a", "b"):
        json.dump(tags, open(tmp_path / ("tag_%s.json" % d), "w"))
    for opname, fname, overlap, v1, v2 in [("均值滤波", "mean_filter", False, 3, None), ("腐蚀", "erode", True, 3, None)]:
        pv.execute({"operationName": opname, "overlap": overlap, "value1": v1, "value2": v2}, str(tmp_path / "a") + "/",
                   str(tmp_path / "tag_a.json"), ctx)
        tb = json.load(open(tmp_path / "tag_b.json"))                     # the reference's loop, with the per-file shim
        tb_edit = tb.copy()
        for rel, tag in tb.items():
            p = str(tmp_path / "b" / rel)
            pp.resize(p)
            getattr(pp, fname)(p, overlap, v1, v2)
            if overlap:
                parts = os.path.splitext(rel)
                tb_edit[parts[0] + "_copy" + parts[1]] = tag
        json.dump(tb_edit, open(tmp_path / "tag_b.json", "w"))
    assert json.load(open(tmp_path / "tag_a.json")) == json.load(open(tmp_path / "tag_b.json"))
    files = sorted(os.listdir(tmp_path / "a"))
    assert files == sorted(os.listdir(tmp_path / "b")) and len(files) > 6
    for f in files:
        assert open(tmp_path / "a" / f, "rb").read() == open(tmp_path / "b" / f, "rb").read(), f


def test_read_train_results_matches_reference_golden():
    """compat.construct_view.read_train_results == TensorResultView.read_train_results (apps/runtime/views.py:44-72) on the
    vectors tests/golden/make_train_results.py produced by executing the reference's method."""
    from cloud_server_b200.compat import construct_view as cv
    cases = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "train_results.json")))
    assert len(cases) >= 16
    import tempfile
    for c in cases:
        with tempfile.NamedTemporaryFile("w", delete=False) as f:
            f.write(c["text"])
        try:
            assert cv.read_train_results(f.name, c["iter"]) == c["json"], c
        finally:
            os.unlink(f.name)


@pytest.mark.gpu
def test_in_process_views_train_infer_monitor(ctx, tmp_path):
    """compat.construct_view: ConstructView.create_file -> training thread in this process; TensorResultView's parser on
    the result.txt it leaves; InferenceView.create_file -> result.json from ONE resident net (no reload on the second
    request); the checkpoint on disk is a TensorFlow V2 bundle that restores into a fresh net bit for bit."""
    import cv2
    import torch
    from cloud_server_b200 import tf_bundle
    from cloud_server_b200.compat import construct_view as cv, construct_distribute as cd
    from cloud_server_b200.compat.graph import get_net, placeholder, compile_net
    pics = golden("pics.npz")
    mdir = tmp_path / "NJUCloud" / "7" / "model" / "digits"
    os.makedirs(mdir / "data")
    os.makedirs(mdir / "infer")
    tags = {}
    for i, (name, lab) in enumerate(zip(pics["names"], pics["labels"])):
        cv2.imwrite(str(mdir / "data" / str(name)), pics["gray_cv"][i])
        tags[str(name)] = str(int(lab))
    json.dump(tags, open(mdir / "tag.json", "w"))
    body = {"ratio": 0.8, "iter": "120", "learning_rate": "0.01", "loss_name": "entropy", "optimizer_name": "GradientDescentOptimizer",
            "net_type": "CNN", "net_config": {"middle_layer": [{"layer": "conv", "filter": [3, 3, 8], "isBias": "True"},
                                                               {"layer": "active", "active_func": "relu"}, {"layer": "pool"},
                                                               {"layer": "connect", "hidden": 64}, {"layer": "active", "active_func": "relu"}],
                                              "output_layer": {}}}
    json.dump(body, open(mdir / "model.json", "w"))                 # ConstructView.save_model_file (views.py:148-162)
    try:
        job = cv.create_file(body, str(mdir), "file", "digits", wait=True, extra_argv=["--seed=3"])
        assert not job["error"] and "--job_name=worker" in job["argv"] and "--ratio=0.800000" in job["argv"]
        mon = json.loads(cv.read_train_results(str(mdir / "result.txt"), "120"))
        assert [r["step"] for r in mon["every_result"]] == ["0", "100"] and 0.0 <= float(mon["final_accuracy"]) <= 1.0
        files = sorted(os.listdir(mdir / "train_model"))
        assert files == ["model.ckpt-120.data-00000-of-00001", "model.ckpt-120.index", "model.ckpt-120.meta"]
        # inference: a 28x28 bundled picture whose scan stays in range, twice; the second request reuses the resident net
        ok = [0]
        name = str(pics["names"][ok[0]])
        cv2.imwrite(str(mdir / "infer" / name), pics["gray_cv"][ok[0]])
        r1 = json.loads(cv.infer(str(mdir), name, intended=True))
        net1 = cv._LIVE[str(mdir)]["net"]
        before = ctx.launches
        r2 = json.loads(cv.infer(str(mdir), name, intended=True))
        assert r1 == r2 == json.load(open(mdir / "result.json")) and r1["result"] == "success" and r1["message"] in list("0123456789")
        assert cv._LIVE[str(mdir)]["net"] is net1                    # no rebuild, no checkpoint read
        assert ctx.launches - before <= 12                           # one preprocess kernel + one forward pass
        with pytest.raises(ValueError):
            cv.create_file(body, str(mdir), "url", "digits")
        # the bundle restores into a fresh net exactly (variables, Adagrad slots, step), by the reference's names
        t = tf_bundle.read_bundle(str(mdir / "train_model" / "model.ckpt-120"))
        assert {"Variable", "Variable_1", "Variable_2", "Variable_3", "Variable_4", "Variable_5", "Variable_6"} <= set(t)
        assert t["Variable"].shape == (3, 3, 1, 8) and t["Variable_6"].dtype == np.int32 and int(t["Variable_6"]) == 120
        fresh = compile_net(get_net("CNN", body["net_config"], placeholder(28, 28, 1), None), ctx, seed=99)
        assert cd.load_checkpoint(fresh, str(mdir / "train_model" / "model.ckpt-120")) == 120
        assert np.array_equal(fresh.get_params(), net1.get_params()) and fresh.global_step == 120
        fresh.close()
    finally:
        cv.close_all()
