"""Device context + the batched preprocess surface (additive to the reference's per-file API).

PyTorch is used for device memory and streams only; every number is produced by the CUDA
kernels behind the C ABI (``include/njucloud_b200.h``)."""
import ctypes as C

import numpy as np
import torch

from . import _lib

_default_ctx = {}


class Context:
    """One nc_ctx: bound to one device, not thread-safe (one per thread / process)."""

    def __init__(self, device=0):
        if not torch.cuda.is_available():
            raise RuntimeError("no CUDA device: njucloud-b200 has no CPU fallback")
        self.device = int(device)
        h = C.c_void_p()
        _lib.check(None, _lib.lib().nc_init(self.device, C.byref(h)))
        self.handle = h
        self.tdev = torch.device("cuda", self.device)

    def close(self):
        if getattr(self, "handle", None):
            _lib.lib().nc_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.tdev).cuda_stream)

    @property
    def launches(self):
        return int(_lib.lib().nc_launch_count(self.handle))

    # ---------------------------------------------------------------- train-side chain (P2-P4)
    def resize_gray_norm(self, src, out_u8=None, out_f32=None, want_u8=True, want_f32=True, mode=0):
        """src: uint8 tensor [N,H,W,3] (BGR, as cv2.imread) or [N,H,W] / [N,H,W,1].

        CUDA tensor -> nc_pp_resize_gray_norm; CPU tensor (pinned or not) ->
        nc_pp_resize_gray_norm_host.  Returns (u8[N,28,28] | None, f32[N,784] | None) on device.
        Stands in for preprocess.resize + imread(,0) + DataSet normalise
        (apps/preprocess/preprocess.py:28-30,40; construct_distribute.py:503-504).
        mode 0: OpenCV's own resize arithmetic (cv2 with IPP disabled; bit exact), mode 1: the arithmetic of the wheel's
        default (Intel IPP) as a float model (>= 99.9 % of the outputs equal)."""
        if src.dtype != torch.uint8:
            raise ValueError("src must be uint8")
        if src.dim() == 3:
            src = src.unsqueeze(-1)
        if src.dim() != 4:
            raise ValueError("src must be [N,H,W,C]")
        n, h, w, c = src.shape
        if not src[0].is_contiguous() if n else False:
            raise ValueError("each picture must be contiguous HWC")
        stride = src.stride(0) if n > 1 else h * w * c
        if want_u8 and out_u8 is None:
            out_u8 = torch.empty((n, 28, 28), dtype=torch.uint8, device=self.tdev)
        if want_f32 and out_f32 is None:
            out_f32 = torch.empty((n, 784), dtype=torch.float32, device=self.tdev)
        p_u8 = C.c_void_p(out_u8.data_ptr()) if out_u8 is not None else None
        p_f32 = C.c_void_p(out_f32.data_ptr()) if out_f32 is not None else None
        L = _lib.lib()
        if src.is_cuda:
            rc = L.nc_pp_resize_gray_norm(self.handle, C.c_void_p(src.data_ptr()), n, h, w, c, stride, int(mode),
                                          p_u8, p_f32, self._stream())
        else:
            staged = C.c_int(0)
            rc = L.nc_pp_resize_gray_norm_host(self.handle, C.c_void_p(src.data_ptr()), n, h, w, c, stride, int(mode),
                                               p_u8, p_f32, C.byref(staged), self._stream())
        _lib.check(self.handle, rc)
        return out_u8, out_f32

    def resize_cubic(self, src, mode=0):
        """uint8 CUDA [N,H,W,C] -> uint8 [N,28,28,C]: cv2.resize(.., (28,28), INTER_CUBIC) alone
        (what preprocess.resize writes back, apps/preprocess/preprocess.py:29-30)."""
        if src.dtype != torch.uint8 or not src.is_cuda or src.dim() != 4:
            raise ValueError("src must be a uint8 CUDA tensor [N,H,W,C]")
        n, h, w, c = src.shape
        src = src.contiguous()
        out = torch.empty((n, 28, 28, c), dtype=torch.uint8, device=self.tdev)
        rc = _lib.lib().nc_pp_resize_cubic(self.handle, C.c_void_p(src.data_ptr()), n, h, w, c, h * w * c, int(mode),
                                           C.c_void_p(out.data_ptr()), self._stream())
        _lib.check(self.handle, rc)
        return out

    # ---------------------------------------------------------------- JPEG decode (P1)
    @staticmethod
    def jpeg_info(data):
        """(width, height, components, hsamp, vsamp) of a JPEG file's bytes -- header only, host side"""
        buf = (C.c_ubyte * len(data)).from_buffer_copy(data)
        v = [C.c_int() for _ in range(5)]
        _lib.check(None, _lib.lib().nc_jpeg_info(buf, len(data), *[C.byref(x) for x in v]))
        return tuple(x.value for x in v)

    def decode_jpegs(self, files, channel_order="BGR", channels=3, host_threads=0, out=None):
        """files: list of bytes objects (JPEG files of ONE width x height).  Returns a uint8 CUDA tensor
        [N, H, W, channels] (channels 1: [N, H, W]) equal to cv2.imread (BGR) / PIL.Image.open (RGB) of each file
        (apps/preprocess/preprocess.py:28, construct_distribute.py:446).  self.last_jpeg_h2d_bytes = bytes that crossed PCIe."""
        n = len(files)
        if n == 0:
            return torch.empty((0, 0, 0, channels), dtype=torch.uint8, device=self.tdev)
        w, h, _, _, _ = self.jpeg_info(files[0])
        files = [f if isinstance(f, bytes) else bytes(f) for f in files]
        ptrs = (C.c_void_p * n)(*[C.cast(C.c_char_p(f), C.c_void_p).value for f in files])       # no copies: bytes are immutable and kept alive by `files`
        sizes = (C.c_size_t * n)(*[len(f) for f in files])
        if out is None:
            out = torch.empty((n, h, w, channels), dtype=torch.uint8, device=self.tdev)
        moved = C.c_int64(0)
        rc = _lib.lib().nc_jpeg_decode(self.handle, ptrs, sizes, n, w, h, channels, {"RGB": 0, "BGR": 1}[channel_order],
                                       C.c_void_p(out.data_ptr()), h * w * channels, int(host_threads), C.byref(moved), self._stream())
        _lib.check(self.handle, rc)
        self.last_jpeg_h2d_bytes = int(moved.value)
        return out[..., 0] if channels == 1 else out

    # ---------------------------------------------------------------- predict-side chain (P5)
    def predict_chain(self, src, channel_order="RGB", compat=False, want_small20=False, out=None, status=None):
        """src: uint8 CUDA tensor [N,H,W,3] or [N,H,W] ('L').  Returns (f32[N,784], status[N] int32
        (1 where the reference raises IndexError; only meaningful with compat), small20 | None).
        Stands in for construct_inference.py:312-330."""
        if src.dtype != torch.uint8 or not src.is_cuda:
            raise ValueError("src must be a uint8 CUDA tensor")
        if src.dim() == 3:
            src = src.unsqueeze(-1)
        n, h, w, c = src.shape
        stride = src.stride(0) if n > 1 else h * w * c
        if out is None:
            out = torch.empty((n, 784), dtype=torch.float32, device=self.tdev)
        if status is None:
            status = torch.zeros((n,), dtype=torch.int32, device=self.tdev)
        small = torch.empty((n, 20, 20), dtype=torch.uint8, device=self.tdev) if want_small20 else None
        order = {"RGB": 0, "BGR": 1}[channel_order]
        rc = _lib.lib().nc_pp_predict_chain(self.handle, C.c_void_p(src.data_ptr()), n, h, w, c, stride, order,
                                            1 if compat else 0, C.c_void_p(out.data_ptr()),
                                            C.c_void_p(small.data_ptr()) if small is not None else None,
                                            C.c_void_p(status.data_ptr()), self._stream())
        _lib.check(self.handle, rc)
        return out, status, small

    def _status_buf(self, n):
        b = getattr(self, "_status", None)
        if b is None or b.numel() < n:
            b = self._status = torch.zeros((max(n, 256),), dtype=torch.int32, device=self.tdev)
        return b[:n]

    # ---------------------------------------------------------------- both chains over one batch
    def both_chains(self, src, channel_order="BGR", compat=False, want_small20=False, out=None, status=None):
        """Train-side chain (P2-P4) and predict-side chain (P5) of a CUDA batch [N,H,W,C] in one call; for
        aligned 3-channel uploads of >= ~160 rows that is ONE kernel reading every upload once
        (nc_pp_both_chains).  Returns (u8[N,28,28], f32_train[N,784], f32_predict[N,784], status[N],
        small20 | None, fused: bool); bytes identical to resize_gray_norm + predict_chain."""
        if src.dtype != torch.uint8 or not src.is_cuda or src.dim() != 4:
            raise ValueError("src must be a uint8 CUDA tensor [N,H,W,C]")
        src = src if src.is_contiguous() else src.contiguous()
        n, h, w, c = src.shape
        if out is None:
            out = (torch.empty((n, 28, 28), dtype=torch.uint8, device=self.tdev),
                   torch.empty((n, 784), dtype=torch.float32, device=self.tdev),
                   torch.empty((n, 784), dtype=torch.float32, device=self.tdev))
        u8, f32, pred = out
        if status is None:
            status = torch.zeros((n,), dtype=torch.int32, device=self.tdev)
        small = torch.empty((n, 20, 20), dtype=torch.uint8, device=self.tdev) if want_small20 else None
        fused = C.c_int(0)
        rc = _lib.lib().nc_pp_both_chains(self.handle, C.c_void_p(src.data_ptr()), n, h, w, c, h * w * c,
                                          {"RGB": 0, "BGR": 1}[channel_order], 1 if compat else 0,
                                          C.c_void_p(u8.data_ptr()), C.c_void_p(f32.data_ptr()), C.c_void_p(pred.data_ptr()),
                                          C.c_void_p(small.data_ptr()) if small is not None else None,
                                          C.c_void_p(status.data_ptr()), C.byref(fused), self._stream())
        _lib.check(self.handle, rc)
        return u8, f32, pred, status, small, bool(fused.value)

    def preprocess_uploads(self, src, channel_order="BGR", chunk=256, out=None):
        """The whole preprocess half for a batch of decoded uploads [N,H,W,3] uint8: the train-side
        chain (P2-P4) and the predict-side chain (P5).  Returns device tensors
        (u8[N,28,28], f32_train[N,784], f32_predict[N,784]).

        src on the GPU: one launch (two for shapes outside the fused kernel's plan).  src in (pinned) host memory: the batch is streamed through two
        device staging buffers of `chunk` pictures -- H2D copies on a side stream overlap the kernels of
        the previous chunk, so the end-to-end rate is the PCIe rate."""
        if src.dtype != torch.uint8 or src.dim() != 4:
            raise ValueError("src must be uint8 [N,H,W,C]")
        n, h, w, c = src.shape
        if out is None:
            out = (torch.empty((n, 28, 28), dtype=torch.uint8, device=self.tdev),
                   torch.empty((n, 784), dtype=torch.float32, device=self.tdev),
                   torch.empty((n, 784), dtype=torch.float32, device=self.tdev))
        u8, f32, pred = out
        if src.is_cuda:
            self.both_chains(src, channel_order=channel_order, out=out, status=self._status_buf(n))
            return out
        key = (chunk, h, w, c)
        st = getattr(self, "_stage", None)
        if st is None or st["key"] != key:
            st = {"key": key, "buf": [torch.empty((chunk, h, w, c), dtype=torch.uint8, device=self.tdev) for _ in range(2)],
                  "ready": [torch.cuda.Event() for _ in range(2)], "free": [torch.cuda.Event() for _ in range(2)],
                  "copy": torch.cuda.Stream(device=self.tdev)}
            self._stage = st
        main = torch.cuda.current_stream(self.tdev)
        for e in st["free"]:
            e.record(main)
        for i, lo in enumerate(range(0, n, chunk)):
            hi = min(n, lo + chunk)
            b = i & 1
            st["copy"].wait_event(st["free"][b])
            with torch.cuda.stream(st["copy"]):
                st["buf"][b][:hi - lo].copy_(src[lo:hi], non_blocking=True)
                st["ready"][b].record(st["copy"])
            main.wait_event(st["ready"][b])
            d = st["buf"][b][:hi - lo]
            self.both_chains(d, channel_order=channel_order, out=(u8[lo:hi], f32[lo:hi], pred[lo:hi]),
                             status=self._status_buf(hi - lo))
            st["free"][b].record(main)
        return out

    # ---------------------------------------------------------------- the 14 ops (P6-P19)
    def op(self, op_id, imgs, v1=None, v2=None, draws=None, out=None):
        """imgs: uint8 CUDA tensor [N,28,28]; one reference op applied to every picture."""
        if imgs.dtype != torch.uint8 or not imgs.is_cuda or tuple(imgs.shape[1:]) != (28, 28):
            raise ValueError("imgs must be a uint8 CUDA tensor [N,28,28]")
        imgs = imgs.contiguous()
        n = imgs.shape[0]
        if out is None:
            out = torch.empty_like(imgs)
        is_int = (1 if isinstance(v1, (int, np.integer)) and not isinstance(v1, bool) else 0) | \
                 (2 if isinstance(v2, (int, np.integer)) and not isinstance(v2, bool) else 0)
        dptr, per = None, 0
        if draws is not None:
            d = np.ascontiguousarray(np.asarray(draws, np.float64).reshape(n, -1))
            per = d.shape[1]
            dptr = d.ctypes.data_as(C.c_void_p)
        rc = _lib.lib().nc_pp_op(self.handle, int(op_id), C.c_void_p(imgs.data_ptr()), C.c_void_p(out.data_ptr()), n,
                                 float(0.0 if v1 is None else v1), float(0.0 if v2 is None else v2), is_int,
                                 dptr, per, self._stream())
        _lib.check(self.handle, rc)
        return out

    def matmul(self, a, b, bias=None, trans_a=False, trans_b=False, precision=0):
        """tf.matmul(a, b) + bias on the TF32 tcgen05 path (construct_distribute.py:182, :264)."""
        a, b = a.contiguous(), b.contiguous()
        m, k = (a.shape[1], a.shape[0]) if trans_a else (a.shape[0], a.shape[1])
        n = b.shape[0] if trans_b else b.shape[1]
        out = torch.empty((m, n), dtype=torch.float32, device=self.tdev)
        rc = _lib.lib().nc_matmul(self.handle, C.c_void_p(a.data_ptr()), C.c_void_p(b.data_ptr()),
                                  C.c_void_p(bias.data_ptr()) if bias is not None else None, m, k, n,
                                  1 if trans_a else 0, 1 if trans_b else 0, int(precision), C.c_void_p(out.data_ptr()),
                                  self._stream())
        _lib.check(self.handle, rc)
        return out

    def normalise(self, imgs_u8):
        imgs_u8 = imgs_u8.contiguous()
        n = imgs_u8.shape[0]
        out = torch.empty((n, imgs_u8.numel() // max(n, 1)), dtype=torch.float32, device=self.tdev)
        rc = _lib.lib().nc_pp_normalise(self.handle, C.c_void_p(imgs_u8.data_ptr()), imgs_u8.numel(),
                                        C.c_void_p(out.data_ptr()), self._stream())
        _lib.check(self.handle, rc)
        return out


def default_context(device=None):
    if device is None:
        device = torch.cuda.current_device() if torch.cuda.is_available() else 0
    if device not in _default_ctx:
        _default_ctx[device] = Context(device)
    return _default_ctx[device]


def preprocess_batch(src, device=None):
    """Batched train-side chain: decoded uploads [N,H,W,3] uint8 (BGR) -> (u8[N,28,28], f32[N,784])."""
    return default_context(device).resize_gray_norm(src)
