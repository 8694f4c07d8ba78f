"""In-process stand-ins for the callers of the training, inference and monitoring path.

The reference reaches its training / inference scripts through SSH + `docker cp` + `docker exec` + `nohup python ... &`
(apps/construction/views.py:97-146, :209-250) and reads the metrics back with another `docker cp`
(apps/runtime/views.py:17-42).  One request therefore pays a process start, a graph rebuild and a checkpoint read --
for inference on ONE picture.  Here the same three steps are direct calls in the serving process:

  create_file(json_str, model_dir, datatype, modelname)   <- ConstructView.create_file   (views.py:97-146)
  infer(model_dir, infer_filename)                        <- InferenceView.create_file   (views.py:209-250)
  read_train_results(filepath, iter)                      <- TensorResultView.read_train_results (runtime/views.py:44-72)

Same inputs (the posted model JSON with its "ratio" key, the model directory that holds ./data, tag.json, model.json,
./infer/<upload>), same files left behind (result.txt, train_model/<model>.ckpt-<step>.*, result.json), same return
values (the result.json text; the JSON the monitoring view returns).  What changes is where the work runs: training is
a thread of this process driving the CUDA library (the `nohup ... &` of the reference), and ONE live net per model is
kept on the GPU, so a second inference request costs a preprocess kernel and a forward pass, not a process launch."""
import json
import os
import threading

_LIVE = {}            # model_dir -> {"net": Net, "step": int, "config": str}
_JOBS = {}            # model_dir -> {"thread": Thread, "stop": Event, "error": [exc]}
_LOCK = threading.Lock()


def _abs(model_dir):
    return os.path.abspath(model_dir)


def stop_training(model_dir, wait=True):
    """What `pkill -9 python` (views.py:129) does to a previous run of the same model -- politely: the loop checks the
    flag where the reference checks sv.should_stop() (construct_distribute.py:402)."""
    job = _JOBS.get(_abs(model_dir))
    if job and job["thread"].is_alive():
        job["stop"].set()
        if wait:
            job["thread"].join()


def create_file(json_str, model_dir, datatype="file", modelname=None, wait=False, extra_argv=()):
    """ConstructView.create_file: start training `model_dir` from the posted JSON (views.py:108-137).

    json_str: the request body (dict): {"ratio": .., "iter": .., "learning_rate": .., "loss_name": .., "optimizer_name": ..,
    "net_type": .., "net_config": {...}}.  Returns the job record (its thread runs compat.construct_distribute.main with
    the command line get_sample_train_cmd builds for the worker, cmd.py:63-69).  wait=True joins the thread and re-raises
    what it raised; the reference never learns about a failed run (nohup), so wait=False swallows nothing either: the
    exception is kept in job["error"]."""
    from cloud_server_b200.compat import construct_distribute as cd
    model_dir = _abs(model_dir)
    config = json.loads(json.dumps(json_str))
    ratio = config['ratio']
    config.pop('ratio')
    config = json.dumps(config)
    if datatype == "url":
        raise ValueError("datatype 'url' trains on TensorFlow's MNIST download (construct_distribute_url.py:384-385): no network here")
    stop_training(model_dir)
    argv = ["--mode=train", "--ps_hosts=", "--worker_hosts=", "--job_name=worker", "--task_index=0",
            "--config=" + config, "--ratio=%f" % float(ratio)] + list(extra_argv)
    stop = threading.Event()
    job = {"stop": stop, "error": [], "argv": argv}

    def run():
        try:
            keep = {}
            cd.main(argv, base_dir=model_dir, should_stop=stop.is_set, keep=keep)
            if "net" in keep:
                with _LOCK:
                    old = _LIVE.pop(model_dir, None)
                    _LIVE[model_dir] = {"net": keep["net"], "step": keep["step"], "config": config}
                if old:
                    old["net"].close()
        except BaseException as e:   # noqa: BLE001 -- kept for the caller, as a crashed nohup job leaves its traceback in nohup.out
            job["error"].append(e)

    job["thread"] = threading.Thread(target=run, name="train:" + (modelname or os.path.basename(model_dir)), daemon=True)
    _JOBS[model_dir] = job
    job["thread"].start()
    if wait:
        job["thread"].join()
        if job["error"]:
            raise job["error"][0]
    return job


def _live_net(model_dir, config, model_name="model", save_rel="./train_model/"):
    """The net of `model_dir`, resident on the GPU, at its newest checkpoint (reloaded only when a newer one appears)."""
    from cloud_server_b200.compat import construct_distribute as cd
    from cloud_server_b200.compat.graph import get_net, placeholder, compile_net
    from cloud_server_b200.context import default_context
    pre = os.path.join(model_dir, save_rel)
    step = cd.get_model(pre, model_name) if os.path.isdir(pre) else -1
    base = os.path.join(pre, model_name + ".ckpt-" + str(step))
    if not cd.is_already_save(base):
        raise ValueError("invalid mode: 'file'")                        # construct_inference.py:308 raises this before writing
    with _LOCK:
        live = _LIVE.get(model_dir)
        if live and live["config"] == config and live["step"] >= step:
            return live["net"]
    cfg = json.loads(config)
    net = compile_net(get_net(cfg["net_type"], cfg["net_config"], placeholder(28, 28, 1), None), default_context())
    cd.load_checkpoint(net, base)
    with _LOCK:
        old = _LIVE.pop(model_dir, None)
        _LIVE[model_dir] = {"net": net, "step": step, "config": config}
    if old:
        old["net"].close()
    return net


def infer(model_dir, infer_filename, intended=False):
    """InferenceView.create_file: classify model_dir/infer/<infer_filename>, write model_dir/result.json, return its text.

    The model JSON comes from model_dir/model.json (get_model_json, views.py:180-196) minus "ratio" (:225-226).  compat
    (default) keeps the reference's IndexError for pictures that hit `arr[bizar+i][bizar+j] = 220`
    (construct_inference.py:325-326); intended=True skips that statement."""
    import numpy as np
    import torch
    from PIL import Image
    from cloud_server_b200.context import default_context
    model_dir = _abs(model_dir)
    with open(os.path.join(model_dir, "model.json"), 'r') as load_f:
        config = json.load(load_f)
    config.pop('ratio')
    config = json.dumps(config)
    net = _live_net(model_dir, config)
    ctx = default_context()
    im = np.asarray(Image.open(os.path.join(model_dir, 'infer', infer_filename)))
    if im.ndim == 3 and im.shape[2] == 4:
        im = im[:, :, :3]
    src = torch.from_numpy(np.array(im)[None]).to(ctx.tdev)
    arr, status, _ = ctx.predict_chain(src, channel_order="RGB", compat=not intended, want_small20=True)
    if int(status[0].item()) != 0:
        raise IndexError("index %d is out of bounds for axis 0 with size 20" % 20)    # :326
    data = {"result": "success", "message": str(int(net.predict(arr)[0].item()))}
    with open(os.path.join(model_dir, "result.json"), 'w') as f:
        json.dump(data, f)
    return json.dumps(data)


def read_train_results(filepath, iter):
    """TensorResultView.read_train_results: result.txt -> the JSON the monitoring endpoint returns.

    Lines `step:<d>,accuracy:<f>,duration:<f>` become {"step","accuracy","duration"} string triples under "every_result";
    once more than iter/100 of them exist, "final_accuracy" is the text after `final_accuracy:` on the next line, or --
    when training has not written that line yet -- the mean of the accuracies so far (runtime/views.py:62-70)."""
    data = dict()
    rows, accs = [], []
    with open(filepath, 'r') as f:
        line = f.readline()                       # readline semantics matter: '' only at end of file, '\n' for a blank line
        while line.startswith("step"):
            parts = line.split(',')
            row = {'step': parts[0].split(':')[1], 'accuracy': parts[1].split(':')[1], 'duration': parts[2].split(':')[1].strip()}
            accs.append(float(row['accuracy']))
            rows.append(row)
            line = f.readline()
    data["every_result"] = rows
    if len(rows) > (int(iter) / 100):
        if len(line) < 1:
            data["final_accuracy"] = float(sum(accs)) / len(accs)
        else:
            data["final_accuracy"] = line.split(':')[1].strip()
    return json.dumps(data)


def close_all():
    """stop every training thread and release the resident nets (tests, server shutdown)"""
    for d in list(_JOBS):
        stop_training(d)
    with _LOCK:
        for live in _LIVE.values():
            live["net"].close()
        _LIVE.clear()
