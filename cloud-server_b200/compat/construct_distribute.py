#!/usr/bin/env python3
"""Drop-in for apps/construction/util/construct_distribute.py (and its _url twin): the same command
line (cmd.py:63-69: --mode --ps_hosts --worker_hosts --job_name --task_index --config --ratio, plus
--model_name --save_path --data_dir --filename --result --iter, :28-47), the same JSON --config, the
same outputs (result.txt lines of :409-420, checkpoints named <model_name>.ckpt-<step>.*), with the
TensorFlow graph replaced by the CUDA library.

Differences that follow from the design, not from the interface:
  * --job_name=ps returns at once: there is no parameter server, the weights live on the GPU.
  * under torchrun (WORLD_SIZE > 1) every rank trains on its slice of each batch and the flat
    gradient buffer is all-reduced over NCCL (net.DataParallel).
  * checkpoints are TensorFlow V2 bundles (<model>.ckpt-<step>.index / .data-00000-of-00001, written without TensorFlow by
    cloud_server_b200.tf_bundle with the variable names the reference's Saver would use) + an empty .meta marker so
    get_model() finds the newest step; the .npz files of earlier versions of this script are still read.
As in the reference the optimizer that runs is Adagrad(1e-4) and the batch is 50 whatever the
config says (:372-373, :403); pass --honor_config to use config["optimizer_name"/"learning_rate"].
"""
import argparse
import json
import os
import sys
import time

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(_HERE)))

from cloud_server_b200.compat.graph import (weight_variable, bias_variable, conv2d, max_pool, active, normalize,  # noqa: F401,E402
                                            connect_layer, cnn, get_net, loss_function, optimizer_function, placeholder,
                                            compile_net)


def build_flags(argv=None):
    ap = argparse.ArgumentParser(allow_abbrev=False)
    ap.add_argument("--mode", default="train")
    ap.add_argument("--ratio", default="0.8")
    ap.add_argument("--model_name", default="model")
    ap.add_argument("--save_path", default="./train_model/")
    ap.add_argument("--data_dir", default="")
    ap.add_argument("--filename", default="")
    ap.add_argument("--result", default="./result.txt")
    ap.add_argument("--ps_hosts", default="")
    ap.add_argument("--worker_hosts", default="")
    ap.add_argument("--job_name", default="")
    ap.add_argument("--task_index", type=int, default=0)
    ap.add_argument("--iter", type=int, default=2000)
    ap.add_argument("--config", default="")
    ap.add_argument("--honor_config", action="store_true", help="(extension) use the config's optimizer and learning rate")
    ap.add_argument("--batch", type=int, default=50, help="(extension) batch size; the reference hard-codes 50")
    ap.add_argument("--seed", type=int, default=None, help="(extension) seed numpy's shuffle RNG and the weight init")
    return ap.parse_args(argv)


def is_already_save(save_path):                                  # :53-54
    return os.path.exists(save_path + ".meta")


def get_model(path, name):                                       # :188-203
    maxstep = -1
    for filename in os.listdir(path):
        if filename.startswith(name + ".ckpt-") and filename.endswith(".meta"):
            step = int(filename[(len(name) + 6):][:-5])
            if step > maxstep:
                maxstep = step
    return maxstep


def dense_to_one_hot(labels_dense, num_classes):
    out = np.zeros((labels_dense.shape[0], num_classes))
    out[np.arange(labels_dense.shape[0]), labels_dense] = 1
    return out


class DataSet(object):
    """DataSet of :466-569 (numpy's global RNG shuffles, exactly as there)."""

    def __init__(self, images, labels, fake_data=False, one_hot=False, dtype=np.float32, reshape=True, seed=None):
        if seed is not None:
            np.random.seed(seed)
        if dtype not in (np.uint8, np.float32):
            raise TypeError('Invalid image dtype %r, expected uint8 or float32' % dtype)
        assert images.shape[0] == labels.shape[0], ('images.shape: %s labels.shape: %s' % (images.shape, labels.shape))
        self._num_examples = images.shape[0]
        if reshape:
            assert images.shape[3] == 1
            images = images.reshape(images.shape[0], images.shape[1] * images.shape[2])
        if dtype == np.float32:
            images = images.astype(np.float32)
            images = np.multiply(images, 1.0 / 255.0)
        self._images, self._labels = images, labels
        self._epochs_completed, self._index_in_epoch = 0, 0

    images = property(lambda self: self._images)
    labels = property(lambda self: self._labels)
    num_examples = property(lambda self: self._num_examples)
    epochs_completed = property(lambda self: self._epochs_completed)

    def next_batch(self, batch_size, fake_data=False, shuffle=True):
        start = self._index_in_epoch
        if self._epochs_completed == 0 and start == 0 and shuffle:
            perm0 = np.arange(self._num_examples)
            np.random.shuffle(perm0)
            self._images, self._labels = self.images[perm0], self.labels[perm0]
        if start + batch_size > self._num_examples:
            self._epochs_completed += 1
            rest_num_examples = self._num_examples - start
            images_rest_part = self._images[start:self._num_examples]
            labels_rest_part = self._labels[start:self._num_examples]
            if shuffle:
                perm = np.arange(self._num_examples)
                np.random.shuffle(perm)
                self._images, self._labels = self.images[perm], self.labels[perm]
            start = 0
            self._index_in_epoch = batch_size - rest_num_examples
            end = self._index_in_epoch
            return (np.concatenate((images_rest_part, self._images[start:end]), axis=0),
                    np.concatenate((labels_rest_part, self._labels[start:end]), axis=0))
        self._index_in_epoch += batch_size
        end = self._index_in_epoch
        return self._images[start:end], self._labels[start:end]


class Datasets(object):
    def __init__(self, train, validation, test):
        self.train, self.validation, self.test = train, validation, test


def read_user_data(path, ratio, base_dir=None):
    """:424-463: tag.json from `path`, pictures from ./data (hard-coded there, :441), *.jpg only,
    labels by file name, split at int(N*ratio) without pre-shuffle.  base_dir (extension): what the
    reference's working directory would be (the model directory the job is started in)."""
    from PIL import Image
    images, labels = [], []
    with open(os.path.join(base_dir or '', path, 'tag.json')) as file:
        lablejson = json.load(file)
    path = os.path.join(base_dir, 'data') if base_dir else 'data'
    for (root, dirs, files) in os.walk(path):
        for filename in files:
            if filename.endswith('.jpg'):
                img = np.array(Image.open(os.path.join(root, filename)))
                img = img.reshape([img.shape[0], img.shape[1], 1])
                labels.append(int(lablejson[filename]))
                images.append(img)
    images = np.array(images)
    labels = dense_to_one_hot(np.array(labels), 10)
    middle = int(labels.shape[0] * ratio)
    return Datasets(train=DataSet(images[:middle], labels[:middle]), validation=None,
                    test=DataSet(images[middle:], labels[middle:]))


def save_checkpoint(net, prefix, step, config, opt_id):
    """what the Supervisor's Saver leaves in logdir (:385-392): <prefix>-<step>.index / .data-00000-of-00001 / .meta"""
    from cloud_server_b200 import tf_bundle
    os.makedirs(os.path.dirname(prefix) or ".", exist_ok=True)
    base = "%s-%d" % (prefix, step)
    tf_bundle.save_net(base, net, opt_id)


def load_checkpoint(net, base):
    """saver.restore(sess, base) (construct_inference.py:336-339)"""
    if os.path.exists(base + ".index") and os.path.getsize(base + ".index") > 0:
        from cloud_server_b200 import tf_bundle
        return tf_bundle.load_net(base, net)
    z = np.load(base + ".npz")                     # written by earlier versions of this script
    net.set_params(z["params"])
    net.set_state(int(z["opt"]), z["state"], int(z["global_step"]))
    return int(z["global_step"])


def main(argv=None, base_dir=None, should_stop=None, keep=None):
    """base_dir / should_stop / keep (extensions, for in-process callers -- compat/construct_view.py): the directory the
    reference would `cd` into before starting the script; a callable polled where the reference polls sv.should_stop()
    (:402); a dict that receives the live net ("net") and its step ("step") when training ends."""
    FLAGS = build_flags(argv)
    model_name = FLAGS.model_name

    def _p(path):
        return path if (not base_dir or os.path.isabs(path)) else os.path.join(base_dir, path)
    pre = _p(FLAGS.save_path)
    save_path = pre + model_name + ".ckpt"
    result_path = _p(FLAGS.result)
    ratio = float(FLAGS.ratio)
    config = json.loads(FLAGS.config)
    iters = int(config["iter"])
    learning_rate = float(config["learning_rate"])
    loss_name = config["loss_name"]
    optimizer_name = config["optimizer_name"]
    net_type = config["net_type"]
    net_config = config["net_config"]

    if FLAGS.job_name == "ps":
        return 0                                   # nothing to serve: see the module docstring
    if FLAGS.job_name != "worker":
        return 0                                   # the reference does nothing for other job names either

    import torch
    from cloud_server_b200.context import Context
    from cloud_server_b200.net import DataParallel, OPT

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    if world > 1 and not torch.distributed.is_initialized():
        torch.distributed.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    if world > 1:
        # every rank must cut its rows out of the SAME shuffled batch: one shuffle seed for all (rank 0's draw unless --seed)
        box = [FLAGS.seed if FLAGS.seed is not None else int(np.random.SeedSequence().entropy % (2 ** 32))]
        torch.distributed.broadcast_object_list(box, src=0)
        np.random.seed(box[0])
        if FLAGS.batch % world:
            raise ValueError("batch %d is not a multiple of the %d ranks: the global-batch gradient would be mis-scaled" % (FLAGS.batch, world))
    elif FLAGS.seed is not None:
        np.random.seed(FLAGS.seed)
    ctx = Context(local_rank)
    mnist = read_user_data(FLAGS.data_dir, ratio, base_dir)
    x = placeholder(28, 28, 1)
    y_conv = get_net(net_type, net_config, x, None)
    loss = loss_function(loss_name, y_conv, None)
    if loss is None:
        raise ValueError("No gradients provided for any variable (loss_function returned None for %r)" % loss_name)
    net = compile_net(y_conv, ctx, seed=(FLAGS.seed or 0) + 1)
    dp = DataParallel(net)
    if FLAGS.honor_config:
        opt_name, lr = optimizer_function(optimizer_name, learning_rate) if optimizer_name not in OPT else (optimizer_name, learning_rate)
    else:
        opt_name, lr = "compat", 1e-4              # tf.train.AdagradOptimizer(1e-4), :372-373
    opt_id = 0 if opt_name == "compat" else OPT[opt_name]
    step = 0
    newest = get_model(pre, model_name) if os.path.isdir(pre) else -1
    if newest >= 0 and is_already_save("%s-%d" % (save_path, newest)):
        step = load_checkpoint(net, "%s-%d" % (save_path, newest))     # Supervisor restores from logdir (:385-392)
    duration = 0.0
    last_save = time.time()
    batch = FLAGS.batch
    dev = ctx.tdev

    graphed = None

    def accuracy_of(images, labels):
        xs = torch.from_numpy(np.ascontiguousarray(images, np.float32)).to(dev)
        pred = net.predict(xs).cpu().numpy()
        return float(np.mean(pred == np.asarray(labels).argmax(1))) if len(pred) else float("nan")

    while not (should_stop and should_stop()) and step < iters:
        bx, by = mnist.train.next_batch(batch)
        step = net.global_step
        if step % 100 == 0:
            train_accuracy = accuracy_of(bx, by)
            if rank == 0:
                with open(result_path, 'a') as file:
                    file.write('step:%d,accuracy:%f,duration:%f\n' % (step, train_accuracy, float(duration)))
        start_time = time.time()
        n_local = len(bx) // world
        lo = rank * n_local
        xs = torch.from_numpy(np.ascontiguousarray(bx[lo:lo + n_local], np.float32)).to(dev)
        ys = torch.from_numpy(np.ascontiguousarray(by[lo:lo + n_local], np.float32)).to(dev)
        if world == 1 and opt_id != 2:
            # fixed batch size, Adagrad / SGD: replay the step from a CUDA graph (one launch instead of ~30)
            if graphed is None or graphed.batch != n_local:
                from cloud_server_b200.net import GraphedTrainStep
                graphed = GraphedTrainStep(net, n_local, loss_name, opt_name, lr)
            graphed(xs, ys)
        else:
            dp.train_step(xs, ys, loss_name, opt_name, lr)
        torch.cuda.synchronize()
        duration = time.time() - start_time
        step = net.global_step
        if rank == 0 and time.time() - last_save >= 60:        # save_model_secs=60 (:391)
            save_checkpoint(net, save_path, step, config, opt_id)
            last_save = time.time()
    test_accuracy = accuracy_of(mnist.test.images, mnist.test.labels)
    if rank == 0:
        with open(result_path, 'a') as file:
            file.write('final_accuracy:%f\n' % test_accuracy + "\n")
        save_checkpoint(net, save_path, net.global_step, config, opt_id)
    if world > 1:
        torch.distributed.barrier()
    if keep is not None:
        keep["net"], keep["step"] = net, net.global_step
    return 0


if __name__ == "__main__":
    sys.exit(main())
