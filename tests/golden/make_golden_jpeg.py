"""Golden vectors for the JPEG decoder.  Run HERE:  python tests/golden/make_golden_jpeg.py  ->  tests/golden/jpeg.npz
  blob / offsets          the bytes of the 99 bundled test pictures (/root/reference/test-data/test-pics/*.jpg, 423-631 B each)
  gray                    cv2.imread(path, cv2.IMREAD_UNCHANGED) of each (identical to PIL's 'L' decode: asserted)
  color_blob / color_offsets / color_bgr_k
                          four small colour files written by cv2.imwrite's defaults (q95, 4:2:0) and one 4:4:4, + cv2.imread"""
import glob
import io
import os

import cv2
import numpy as np
from PIL import Image

HERE = os.path.dirname(os.path.abspath(__file__))
files = sorted(glob.glob("/root/reference/test-data/test-pics/*.jpg"))
assert len(files) == 99
names = np.load(os.path.join(HERE, "pics.npz"))["names"]
files = [os.path.join("/root/reference/test-data/test-pics", str(n)) for n in names]       # same order as pics.npz
blob, offs, gray = b"", [0], []
for f in files:
    d = open(f, "rb").read()
    blob += d
    offs.append(len(blob))
    a = cv2.imread(f, cv2.IMREAD_UNCHANGED)
    assert np.array_equal(a, np.asarray(Image.open(f)))
    gray.append(a)
out = {"blob": np.frombuffer(blob, np.uint8), "offsets": np.array(offs), "gray": np.stack(gray)}
rng = np.random.default_rng(42)
cblob, coffs = b"", [0]
for k, (h, w, ss) in enumerate(((40, 56, None), (31, 45, None), (64, 64, cv2.IMWRITE_JPEG_SAMPLING_FACTOR_444), (50, 70, cv2.IMWRITE_JPEG_SAMPLING_FACTOR_422))):
    img = cv2.GaussianBlur(rng.integers(0, 256, (h, w, 3), np.uint8), (0, 0), 1.5)
    ok, buf = cv2.imencode(".jpg", img, [] if ss is None else [cv2.IMWRITE_JPEG_SAMPLING_FACTOR, ss])
    d = buf.tobytes()
    cblob += d
    coffs.append(len(cblob))
    out["color_bgr_%d" % k] = cv2.imdecode(buf, cv2.IMREAD_COLOR)
out["color_blob"], out["color_offsets"] = np.frombuffer(cblob, np.uint8), np.array(coffs)
np.savez_compressed(os.path.join(HERE, "jpeg.npz"), **out)
print("wrote jpeg.npz", len(blob), len(cblob))
