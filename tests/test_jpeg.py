"""JPEG decode (cv2.imread at apps/preprocess/preprocess.py:28,:40, Image.open at
construct_distribute.py:446).  CPU: the numpy oracle (oracle/jpeg_restate.py) is pinned against the live cv2 / PIL
decoders (both embed libjpeg-turbo) and against the committed golden decodes of the bundled pictures; the library's
host-side Huffman stage is checked against the oracle's.  GPU: nc_jpeg_decode == cv2.imread / PIL byte for byte."""
import ctypes as C
import io

import numpy as np
import pytest

from conftest import golden
from oracle import jpeg_restate as J

cv2 = pytest.importorskip("cv2")
from PIL import Image  # noqa: E402


def encode(img, q=95, ss=None, extra=()):
    params = [cv2.IMWRITE_JPEG_QUALITY, q]
    if ss is not None:
        params += [cv2.IMWRITE_JPEG_SAMPLING_FACTOR, ss]
    ok, buf = cv2.imencode(".jpg", img, params + list(extra))
    assert ok
    return buf.tobytes()


def cases(seed=0, sizes=((64, 64), (37, 53), (17, 9), (8, 8), (1, 1), (33, 16), (120, 161))):
    rng = np.random.default_rng(seed)
    out = []
    for (h, w) in sizes:
        noise = rng.integers(0, 256, (h, w, 3), np.uint8)
        smooth = cv2.GaussianBlur(noise, (0, 0), 2)
        for img, tag in ((noise, "noise"), (smooth, "smooth")):
            for q in (95, 60):
                for ss, sn in ((cv2.IMWRITE_JPEG_SAMPLING_FACTOR_420, "420"), (cv2.IMWRITE_JPEG_SAMPLING_FACTOR_444, "444"),
                               (cv2.IMWRITE_JPEG_SAMPLING_FACTOR_422, "422")):
                    out.append(("%dx%d %s q%d %s" % (h, w, tag, q, sn), encode(img, q, ss)))
        out.append(("%dx%d gray" % (h, w), encode(noise[..., 0].copy(), 95)))
        out.append(("%dx%d restart" % (h, w), encode(smooth, 90, cv2.IMWRITE_JPEG_SAMPLING_FACTOR_420, (cv2.IMWRITE_JPEG_RST_INTERVAL, 3))))
    return out


def test_oracle_equals_cv2_and_pil_on_synthetic_encodes():
    for name, data in cases():
        want_cv = cv2.imdecode(np.frombuffer(data, np.uint8), cv2.IMREAD_UNCHANGED)
        assert np.array_equal(J.decode(data, "BGR"), want_cv), name
        assert np.array_equal(J.decode(data, "RGB"), np.asarray(Image.open(io.BytesIO(data)))), name


def test_oracle_equals_golden_decodes_of_bundled_pictures():
    """tests/golden/jpeg.npz holds the 99 bundled files' bytes and their cv2.imread / PIL decodes (made by make_golden_jpeg.py)"""
    g = golden("jpeg.npz")
    blob, offs = g["blob"].tobytes(), g["offsets"]
    for i in range(len(offs) - 1):
        data = blob[offs[i]:offs[i + 1]]
        assert np.array_equal(J.decode(data), g["gray"][i]), i
    for k in range(len(g["color_offsets"]) - 1):
        data = g["color_blob"].tobytes()[g["color_offsets"][k]:g["color_offsets"][k + 1]]
        assert np.array_equal(J.decode(data, "BGR"), g["color_bgr_%d" % k]), k


def test_host_huffman_stage_equals_oracle(built_lib):
    L = built_lib
    for name, data in cases(seed=3):
        hdr = J.parse(data)
        want = np.concatenate([c.reshape(-1, 64) for c in J.entropy_decode(hdr)])
        buf = (C.c_ubyte * len(data)).from_buffer_copy(data)
        nb = C.c_int64()
        assert L.nc_jpeg_entropy_decode(buf, len(data), None, 0, C.byref(nb)) == 0, name
        assert nb.value == want.shape[0], name
        got = np.zeros((nb.value, 64), np.int16)
        assert L.nc_jpeg_entropy_decode(buf, len(data), got.ctypes.data_as(C.c_void_p), nb.value, C.byref(nb)) == 0
        assert np.array_equal(got, want), name
        w, h, nc, hs, vs = (C.c_int() for _ in range(5))
        assert L.nc_jpeg_info(buf, len(data), C.byref(w), C.byref(h), C.byref(nc), C.byref(hs), C.byref(vs)) == 0
        assert (w.value, h.value, nc.value) == (hdr["width"], hdr["height"], len(hdr["comps"]))
    # refusals: progressive files, truncated files, not a JPEG
    ok, prog = cv2.imencode(".jpg", np.zeros((16, 16, 3), np.uint8), [cv2.IMWRITE_JPEG_PROGRESSIVE, 1])
    pb = (C.c_ubyte * len(prog)).from_buffer_copy(prog.tobytes())
    assert L.nc_jpeg_info(pb, len(prog), None, None, None, None, None) == -4           # NC_EUNSUPPORTED
    junk = (C.c_ubyte * 8)(*b"notajpeg")
    assert L.nc_jpeg_info(junk, 8, None, None, None, None, None) == -1
    name, data = cases(seed=3)[0]
    cut = (C.c_ubyte * 200).from_buffer_copy(data[:200])
    nb = C.c_int64()
    assert L.nc_jpeg_entropy_decode(cut, 200, None, 0, C.byref(nb)) in (0, -1)         # zeros past the end or a clean error, never a crash


@pytest.mark.gpu
def test_device_decode_equals_cv2_and_pil(ctx):
    import torch
    g = golden("jpeg.npz")
    blob, offs = g["blob"].tobytes(), g["offsets"]
    files = [blob[offs[i]:offs[i + 1]] for i in range(len(offs) - 1)]
    got = ctx.decode_jpegs(files, channel_order="BGR", channels=3).cpu().numpy()          # cv2.imread(dir): gray replicated to BGR
    assert got.shape == (99, 28, 28, 3)
    for i in range(99):
        assert np.array_equal(got[i], np.repeat(g["gray"][i][..., None], 3, 2)), i
    got1 = ctx.decode_jpegs(files, channels=1).cpu().numpy()                              # Image.open(f) of an 'L' file
    assert np.array_equal(got1, g["gray"])
    for name, data in cases(seed=5, sizes=((64, 64), (37, 53), (17, 9), (1, 1), (120, 161), (768, 1024))):
        want = cv2.imdecode(np.frombuffer(data, np.uint8), cv2.IMREAD_COLOR)
        out = ctx.decode_jpegs([data, data], channel_order="BGR").cpu().numpy()
        assert np.array_equal(out[0], want) and np.array_equal(out[1], want), name
        rgb = ctx.decode_jpegs([data], channel_order="RGB").cpu().numpy()[0]
        assert np.array_equal(rgb, np.asarray(Image.open(io.BytesIO(data)).convert("RGB"))), name
    # q95 4:2:0 re-encodes of synthetic uploads (what cv2.imwrite leaves on disk, preprocess.py:30), one batch, feeding the chains
    rng = np.random.default_rng(9)
    ups = [cv2.GaussianBlur(rng.integers(0, 256, (768, 1024, 3), np.uint8), (0, 0), 3) for _ in range(4)]
    datas = [encode(u) for u in ups]
    dec = ctx.decode_jpegs(datas, channel_order="BGR")
    assert ctx.last_jpeg_h2d_bytes < 0.5 * dec.numel()                                     # the compressed domain crossed PCIe
    for i, d in enumerate(datas):
        assert np.array_equal(dec[i].cpu().numpy(), cv2.imdecode(np.frombuffer(d, np.uint8), cv2.IMREAD_COLOR)), i
    u8, f32 = ctx.resize_gray_norm(dec)
    from oracle import cvref
    for i, d in enumerate(datas):
        e8, ef = cvref.train_chain(cv2.imdecode(np.frombuffer(d, np.uint8), cv2.IMREAD_COLOR))
        assert np.array_equal(u8[i].cpu().numpy(), e8) and np.array_equal(f32[i].cpu().numpy(), ef)
    with pytest.raises(ValueError):
        ctx.decode_jpegs([datas[0], cases()[0][1]])                                       # mixed sizes in one batch
