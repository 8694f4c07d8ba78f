"""Golden vectors of cv2.resize INTER_CUBIC in the wheel's DEFAULT arithmetic (Intel IPP enabled) -- what
apps/preprocess/preprocess.py:29 computes unless the host switches IPP off.  Run HERE:
  python tests/golden/make_golden_ipp.py  ->  tests/golden/train_chain_ipp.npz
Same seeded uploads as train_chain.npz (make_golden.py: TRAIN_CASES); stores the resized BGR picture, the gray picture
and how many outputs differ from the IPP-off arithmetic (mode 0)."""
import os
import sys

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, HERE)
from make_golden import TRAIN_CASES, synth_upload   # noqa: E402  (importing it switches IPP off: switch it back on below)

out = {"cases": np.array(TRAIN_CASES)}
diff = tot = 0
for seed, h, w, nat in TRAIN_CASES:
    img = synth_upload(seed, h, w, nat)
    cv2.ipp.setUseIPP(True)
    small = cv2.resize(img, (28, 28), interpolation=cv2.INTER_CUBIC)
    gray = cv2.cvtColor(small, cv2.COLOR_BGR2GRAY)
    cv2.ipp.setUseIPP(False)
    off = cv2.resize(img, (28, 28), interpolation=cv2.INTER_CUBIC)
    out["bgr_%d" % seed], out["u8_%d" % seed] = small, gray
    out["insum_%d" % seed] = np.array([img.astype(np.uint64).sum()])
    diff += int((small != off).sum())
    tot += small.size
out["mode0_vs_ipp"] = np.array([diff, tot])
np.savez_compressed(os.path.join(HERE, "train_chain_ipp.npz"), **out)
print("IPP-on vs IPP-off: %d of %d outputs differ (%.2f %%)" % (diff, tot, 100.0 * diff / tot), cv2.__version__, cv2.ipp.getIppVersion())
