"""BASELINE config 1: LeNet weights trained by the ORACLE for the 99 bundled pictures.

The reference ships no checkpoint and TensorFlow is not installable, so the weights both sides of the C1
parity run load are produced by oracle/tfref.py itself: seeded init (default_rng(0)), SGD 0.01, STEPS
full-batch steps on the 99 pictures of test-data/test-pics (decoded as PIL 'L' by make_golden.py ->
pics.npz) after the predict-side chain of construct_inference.py:312-330 (intended semantics, i.e. without
the IndexError statement), labels from test-data/tag.json.  13 MB of float32 are not committed: the
training takes ~3 s of numpy, so tests and bench.py call c1_weights() (deterministic for a given numpy /
BLAS; both sides always load the very same array, which is what the parity run needs).

  python tests/golden/make_c1_weights.py        # prints the training trajectory and tfref's labels
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

STEPS = 30
LR = 0.01


def c1_inputs():
    """(x [99, 784] float32 in {0, 254/255}, labels [99] int) -- the oracle's predict-side chain on the bundled pictures"""
    from oracle import pp_restate as R
    z = np.load(os.path.join(HERE, "pics.npz"))
    x = np.stack([R.predict_canvas(R.pil_resize_bicubic_l(p), compat=False) for p in z["gray_pil"]])
    return x.reshape(len(x), 784).astype(np.float32), z["labels"].astype(np.int64), z["gray_pil"]


def c1_weights(steps=STEPS, verbose=False):
    """flat float32 LeNet parameters (variable-creation order) after `steps` oracle SGD steps"""
    from oracle import tfref
    x, labels, _ = c1_inputs()
    y = np.eye(10, dtype=np.float32)[labels]
    ref = tfref.Net(tfref.LENET)
    w = ref.flatten(tfref.init_params(ref.param_specs, np.random.default_rng(0)))
    opt = tfref.SGD(w.size, LR)
    for s in range(steps):
        loss, logits, _ = tfref.train_step(ref, w, x, y, opt)
        if verbose and (s % 5 == 0 or s == steps - 1):
            print("step %d loss %.4f train accuracy %.3f" % (s, loss, float((logits.argmax(1) == labels).mean())))
    return w


if __name__ == "__main__":
    from oracle import tfref
    w = c1_weights(verbose=True)
    x, labels, _ = c1_inputs()
    ref = tfref.Net(tfref.LENET)
    pred = ref.forward(x, ref.unflatten(w)).argmax(1)
    print("tfref labels:", "".join(str(int(p)) for p in pred))
    print("tag.json    :", "".join(str(int(p)) for p in labels))
    print("accuracy vs tag.json: %d / 99" % int((pred == labels).sum()))
