"""numpy restatement of the JPEG decoder behind cv2.imread / PIL.Image.open (TEST INFRASTRUCTURE ONLY).

The reference decodes every upload with `cv2.imread(dir)` (apps/preprocess/preprocess.py:28, :40) and
`Image.open(...)` (apps/construction/util/construct_distribute.py:446, construct_inference.py:312).  Both wheels embed
libjpeg-turbo (opencv-python-headless 4.13.0: libjpeg-turbo 3.1.2; Pillow 12.2.0), whose source is NOT under
/root/reference; this file restates its baseline path from the published algorithm -- ITU T.81 Huffman decoding,
jidctint.c's "islow" integer inverse DCT (CONST_BITS 13, PASS1_BITS 2), jdsample.c's "fancy" triangle upsampling for 2x1 and
2x2 chroma, jdcolor.c's fixed-point YCbCr -> RGB -- and is pinned by tests/test_jpeg.py against the live cv2.imread /
PIL decodes of the 99 bundled pictures and of synthetic encodes (4:2:0, 4:2:2, 4:4:4, gray, odd sizes, restart
intervals, several qualities): bit exact.

Pure-Python bit reading: slow (seconds per megapixel) -- fine for a checker.
"""
from __future__ import annotations

import numpy as np

ZIGZAG = np.array([0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21, 28,
                   35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63])


class JpegError(ValueError):
    pass


def parse(data: bytes):
    """-> dict(width, height, comps=[(id, h, v, tq)], qt={id: int32[64] natural order}, dc/ac huffman tables, scan comps,
    restart interval, entropy-coded bytes)"""
    if data[:2] != b"\xff\xd8":
        raise JpegError("not a JPEG (no SOI)")
    i = 2
    out = {"qt": {}, "dc": {}, "ac": {}, "ri": 0}
    while i < len(data):
        if data[i] != 0xFF:
            raise JpegError("marker expected at %d" % i)
        while data[i] == 0xFF:
            i += 1
        m = data[i]
        i += 1
        if m in (0xD8, 0x01) or 0xD0 <= m <= 0xD7:
            continue
        if m == 0xD9:
            break
        ln = (data[i] << 8) | data[i + 1]
        seg = data[i + 2:i + ln]
        if m == 0xDB:
            j = 0
            while j < len(seg):
                pq, tq = seg[j] >> 4, seg[j] & 15
                j += 1
                if pq:
                    vals = [(seg[j + 2 * k] << 8) | seg[j + 2 * k + 1] for k in range(64)]
                    j += 128
                else:
                    vals = list(seg[j:j + 64])
                    j += 64
                q = np.zeros(64, np.int32)
                q[ZIGZAG] = vals
                out["qt"][tq] = q
        elif m in (0xC0, 0xC1):
            if seg[0] != 8:
                raise JpegError("only 8-bit precision")
            out["height"], out["width"] = (seg[1] << 8) | seg[2], (seg[3] << 8) | seg[4]
            n = seg[5]
            out["comps"] = [(seg[6 + 3 * k], seg[7 + 3 * k] >> 4, seg[7 + 3 * k] & 15, seg[8 + 3 * k]) for k in range(n)]
        elif m in (0xC2, 0xC3, 0xC5, 0xC6, 0xC7, 0xC9, 0xCA, 0xCB, 0xCD, 0xCE, 0xCF):
            raise JpegError("unsupported JPEG process (SOF%d): baseline / extended sequential Huffman only" % (m - 0xC0))
        elif m == 0xC4:
            j = 0
            while j < len(seg):
                tc, th = seg[j] >> 4, seg[j] & 15
                counts = list(seg[j + 1:j + 17])
                nsym = sum(counts)
                syms = list(seg[j + 17:j + 17 + nsym])
                j += 17 + nsym
                out["ac" if tc else "dc"][th] = (counts, syms)
        elif m == 0xDD:
            out["ri"] = (seg[0] << 8) | seg[1]
        elif m == 0xDA:
            ns = seg[0]
            out["scan"] = [(seg[1 + 2 * k], seg[2 + 2 * k] >> 4, seg[2 + 2 * k] & 15) for k in range(ns)]
            out["ecs"] = data[i + ln:]
            return out
        i += ln
    raise JpegError("no SOS marker")


def _huff_lookup(counts, syms):
    """(code length, code) -> symbol"""
    table, code, k = {}, 0, 0
    for ln in range(1, 17):
        for _ in range(counts[ln - 1]):
            table[(ln, code)] = syms[k]
            code += 1
            k += 1
        code <<= 1
    return table


class _Bits:
    def __init__(self, data):
        self.d, self.i, self.acc, self.n = data, 0, 0, 0

    def _fill(self):
        while self.n <= 24:
            if self.i >= len(self.d):
                b = 0
            else:
                b = self.d[self.i]
                if b == 0xFF:
                    nxt = self.d[self.i + 1] if self.i + 1 < len(self.d) else 0xD9
                    if nxt == 0:
                        self.i += 2
                    else:
                        b = 0           # a marker: feed zeros, do not advance
                else:
                    self.i += 1
            self.acc = ((self.acc << 8) | b) & 0xFFFFFFFFFF
            self.n += 8

    def get(self, k):
        if k == 0:
            return 0
        if self.n < k:
            self._fill()
        self.n -= k
        return (self.acc >> self.n) & ((1 << k) - 1)

    def decode(self, table):
        code = 0
        for ln in range(1, 17):
            code = (code << 1) | self.get(1)
            s = table.get((ln, code))
            if s is not None:
                return s
        raise JpegError("bad Huffman code")

    def restart(self):
        """skip to just behind the next RSTn marker; the bit buffer is discarded"""
        self.acc, self.n = 0, 0
        while self.i + 1 < len(self.d) and not (self.d[self.i] == 0xFF and 0xD0 <= self.d[self.i + 1] <= 0xD7):
            self.i += 1
        self.i += 2


def _extend(v, s):
    return v if v >= (1 << (s - 1)) else v - (1 << s) + 1


def entropy_decode(hdr):
    """-> list per component of int16 [blocks_v, blocks_h, 64] quantised coefficients in natural order (MCU-padded)"""
    comps = hdr["comps"]
    hmax, vmax = max(c[1] for c in comps), max(c[2] for c in comps)
    W, H = hdr["width"], hdr["height"]
    mcux, mcuy = -(-W // (8 * hmax)), -(-H // (8 * vmax))
    order = {c[0]: k for k, c in enumerate(comps)}
    if len(hdr["scan"]) != len(comps):
        raise JpegError("non-interleaved scans are not supported")
    coefs = [np.zeros((mcuy * c[2], mcux * c[1], 64), np.int16) for c in comps]
    dct = {k: _huff_lookup(*v) for k, v in hdr["dc"].items()}
    act = {k: _huff_lookup(*v) for k, v in hdr["ac"].items()}
    bits = _Bits(hdr["ecs"])
    pred = [0] * len(comps)
    ri, cnt = hdr["ri"], 0
    for my in range(mcuy):
        for mx in range(mcux):
            if ri and cnt and cnt % ri == 0:
                bits.restart()
                pred = [0] * len(comps)
            cnt += 1
            for cid, td, ta in hdr["scan"]:
                ci = order[cid]
                _, h, v, _ = comps[ci]
                for by in range(v):
                    for bx in range(h):
                        blk = coefs[ci][my * v + by, mx * h + bx]
                        s = bits.decode(dct[td])
                        pred[ci] += _extend(bits.get(s), s) if s else 0
                        blk[0] = pred[ci]
                        k = 1
                        while k < 64:
                            rs = bits.decode(act[ta])
                            r, s = rs >> 4, rs & 15
                            if s == 0:
                                if r == 15:
                                    k += 16
                                    continue
                                break
                            k += r
                            blk[ZIGZAG[k]] = _extend(bits.get(s), s)
                            k += 1
    return coefs


# ---------------------------------------------------------------------------------------------- jidctint.c (islow)
_F = dict(f0_298=2446, f0_390=3196, f0_541=4433, f0_765=6270, f0_899=7373, f1_175=9633, f1_501=12299, f1_847=15137,
          f1_961=16069, f2_053=16819, f2_562=20995, f3_072=25172)


def _idct_1d(d, shift_in, descale):
    """one pass of jpeg_idct_islow over the leading axis of d[8, ...] (int64); shift_in: the even part's left shift"""
    z2, z3 = d[2], d[6]
    z1 = (z2 + z3) * _F["f0_541"]
    tmp2 = z1 + z3 * (-_F["f1_847"])
    tmp3 = z1 + z2 * _F["f0_765"]
    z2, z3 = d[0], d[4]
    tmp0 = (z2 + z3) << shift_in
    tmp1 = (z2 - z3) << shift_in
    tmp10, tmp13, tmp11, tmp12 = tmp0 + tmp3, tmp0 - tmp3, tmp1 + tmp2, tmp1 - tmp2
    t0, t1, t2, t3 = d[7], d[5], d[3], d[1]
    z1, z2, z3, z4 = t0 + t3, t1 + t2, t0 + t2, t1 + t3
    z5 = (z3 + z4) * _F["f1_175"]
    t0, t1, t2, t3 = t0 * _F["f0_298"], t1 * _F["f2_053"], t2 * _F["f3_072"], t3 * _F["f1_501"]
    z1, z2, z3, z4 = z1 * (-_F["f0_899"]), z2 * (-_F["f2_562"]), z3 * (-_F["f1_961"]), z4 * (-_F["f0_390"])
    z3, z4 = z3 + z5, z4 + z5
    t0, t1, t2, t3 = t0 + z1 + z3, t1 + z2 + z4, t2 + z2 + z3, t3 + z1 + z4
    rnd = 1 << (descale - 1)
    return np.stack([(tmp10 + t3 + rnd) >> descale, (tmp11 + t2 + rnd) >> descale, (tmp12 + t1 + rnd) >> descale,
                     (tmp13 + t0 + rnd) >> descale, (tmp13 - t0 + rnd) >> descale, (tmp12 - t1 + rnd) >> descale,
                     (tmp11 - t2 + rnd) >> descale, (tmp10 - t3 + rnd) >> descale])


def _range_limit(v):
    """IDCT_range_limit(cinfo)[v & RANGE_MASK]: clamp(v + 128) on [-512, 511], the table's wrap beyond"""
    idx = v & 1023
    return np.where(idx < 128, idx + 128, np.where(idx < 512, 255, np.where(idx < 896, 0, idx - 896))).astype(np.uint8)


def idct_islow(coefs, q):
    """coefs int16 [by, bx, 64] (quantised, natural order), q int32[64] -> uint8 plane [by*8, bx*8]"""
    by, bx, _ = coefs.shape
    d = (coefs.astype(np.int64) * q.astype(np.int64)).reshape(by, bx, 8, 8)      # [.., row, col]
    ws = _idct_1d(np.moveaxis(d, 2, 0), 13, 13 - 2)                               # pass 1: columns -> ws[row, by, bx, col]
    ws = np.moveaxis(ws, 0, 2)                                                    # [by, bx, row, col]
    out = _idct_1d(np.moveaxis(ws, 3, 0), 13, 13 + 2 + 3)                         # pass 2: rows -> out[col, by, bx, row]
    out = np.moveaxis(out, 0, 3)                                                  # [by, bx, row, col]
    return _range_limit(out).transpose(0, 2, 1, 3).reshape(by * 8, bx * 8)


# ---------------------------------------------------------------------------------------------- jdsample.c (fancy)
def _h2v1_fancy_rows(p):
    """p int32 [rows, w] -> [rows, 2w]"""
    rows, w = p.shape
    out = np.empty((rows, 2 * w), np.int32)
    if w == 1:
        out[:, 0] = p[:, 0]
        out[:, 1] = p[:, 0]
        return out
    t = 3 * p
    out[:, 0] = p[:, 0]
    out[:, 1] = (t[:, 0] + p[:, 1] + 2) >> 2
    out[:, 2:-2:2] = (t[:, 1:-1] + p[:, :-2] + 1) >> 2
    out[:, 3:-1:2] = (t[:, 1:-1] + p[:, 2:] + 2) >> 2
    out[:, -2] = (t[:, -1] + p[:, -2] + 1) >> 2
    out[:, -1] = p[:, -1]
    return out


def upsample_h2v1(plane, dw):
    return _h2v1_fancy_rows(plane[:, :dw].astype(np.int32)).astype(np.uint8)


def upsample_h2v2(plane, dw, dh):
    """plane uint8 (MCU-padded); dw, dh = downsampled_width / height (real columns / rows) -> [2*dh, 2*dw]"""
    p = plane[:dh, :dw].astype(np.int32)
    up = np.vstack([p[:1], p[:-1]])           # row above (first row: itself, jdmainct's duplicated context row)
    dn = np.vstack([p[1:], p[-1:]])           # row below (last real row: itself)
    out = np.empty((2 * dh, 2 * dw), np.int32)
    for v, other in ((0, up), (1, dn)):
        cs = 3 * p + other                    # thiscolsum
        o = np.empty((dh, 2 * dw), np.int32)
        if dw == 1:
            o[:, 0] = (cs[:, 0] * 4 + 8) >> 4
            o[:, 1] = (cs[:, 0] * 4 + 7) >> 4
        else:
            o[:, 0] = (cs[:, 0] * 4 + 8) >> 4
            o[:, 1] = (cs[:, 0] * 3 + cs[:, 1] + 7) >> 4
            o[:, 2:-2:2] = (cs[:, 1:-1] * 3 + cs[:, :-2] + 8) >> 4
            o[:, 3:-1:2] = (cs[:, 1:-1] * 3 + cs[:, 2:] + 7) >> 4
            o[:, -2] = (cs[:, -1] * 3 + cs[:, -2] + 8) >> 4
            o[:, -1] = (cs[:, -1] * 4 + 7) >> 4
        out[v::2] = o
    return out.astype(np.uint8)


# ---------------------------------------------------------------------------------------------- jdcolor.c
def ycc_to_rgb(y, cb, cr):
    y, cb, cr = y.astype(np.int32), cb.astype(np.int32) - 128, cr.astype(np.int32) - 128
    r = y + ((91881 * cr + 32768) >> 16)
    g = y + ((-22554 * cb + 32768 - 46802 * cr) >> 16)
    b = y + ((116130 * cb + 32768) >> 16)
    return np.clip(np.stack([r, g, b], -1), 0, 255).astype(np.uint8)


def decode(data: bytes, order: str = "RGB"):
    """-> uint8 [H, W, 3] (order RGB as PIL, BGR as cv2.imread) or [H, W] for a one-component file"""
    hdr = parse(data)
    coefs = entropy_decode(hdr)
    return reconstruct(hdr, coefs, order)


def reconstruct(hdr, coefs, order="RGB"):
    comps = hdr["comps"]
    W, H = hdr["width"], hdr["height"]
    hmax, vmax = max(c[1] for c in comps), max(c[2] for c in comps)
    planes = [idct_islow(cf, hdr["qt"][c[3]]) for cf, c in zip(coefs, comps)]
    if len(comps) == 1:
        return planes[0][:H, :W]
    if len(comps) != 3:
        raise JpegError("%d components" % len(comps))
    full = []
    for p, (_, h, v, _) in zip(planes, comps):
        if h == hmax and v == vmax:
            full.append(p[:H, :W])
        elif h * 2 == hmax and v * 2 == vmax:
            dw, dh = -(-W * h // hmax), -(-H * v // vmax)
            full.append(upsample_h2v2(p, dw, dh)[:H, :W])
        elif h * 2 == hmax and v == vmax:
            dw = -(-W * h // hmax)
            full.append(upsample_h2v1(p, dw)[:H, :W])
        else:
            raise JpegError("unsupported sampling factors %dx%d of %dx%d" % (h, v, hmax, vmax))
    rgb = ycc_to_rgb(*full)
    return rgb if order == "RGB" else rgb[..., ::-1]
