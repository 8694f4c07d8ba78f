"""CPU tests of the drop-in boundary: the C-ABI library builds, loads, exports every symbol the
header declares, fails loudly without a GPU, and its host-side coefficient tables equal the
oracle's (no compute call is made here)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT
from oracle import pp_restate as R


def header_symbols():
    txt = open(os.path.join(ROOT, "include", "njucloud_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(nc_[a-z0-9_]+)\s*\(", txt)))


def test_header_symbols_all_exported(built_lib):
    from cloud_server_b200 import _lib
    syms = header_symbols()
    assert len(syms) >= 30
    assert syms == _lib.exported_symbols(), "binding table and header disagree"
    for s in syms:
        assert hasattr(built_lib, s), s


def test_every_entry_point_cites_reference():
    txt = open(os.path.join(ROOT, "include", "njucloud_b200.h")).read()
    for needle in ["preprocess.py:29", "construct_inference.py:312-330", "construct_distribute.py:208-281",
                   "construct_distribute.py:372-373", "views.py:118-126"]:
        assert needle in txt, needle


def test_no_gpu_fails_loudly(built_lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    h = C.c_void_p()
    rc = built_lib.nc_init(0, C.byref(h))
    assert rc < 0 and not h.value
    assert b"no CPU fallback" in built_lib.nc_last_error(None)
    from cloud_server_b200.context import Context
    with pytest.raises(RuntimeError):
        Context(0)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "cloud-server_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cpp", ".h", ".cuh")):
                src = open(os.path.join(dp, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f
                assert "import cv2" not in src or f in ("preprocess.py", "preprocess_view.py", "imageio.py"), f   # host JPEG codec only


@pytest.mark.parametrize("src", [28, 29, 37, 56, 100, 255, 480, 640, 768, 1024, 1080, 1920, 4001])
def test_cubic_table_equals_oracle(built_lib, src):
    ofs = np.zeros(112, np.int32)
    coef = np.zeros(112, np.int16)
    assert built_lib.nc_pp_cubic_table(src, ofs.ctypes.data_as(C.c_void_p), coef.ctypes.data_as(C.c_void_p)) == 0
    o, c = R.cv_cubic_tables(src, 28)
    assert np.array_equal(ofs.reshape(28, 4), o)
    assert np.array_equal(coef.reshape(28, 4), c)


@pytest.mark.parametrize("size", [20, 21, 23, 28, 64, 300, 768, 1024, 977, 4096])
def test_pil_table_equals_oracle(built_lib, size):
    ks = built_lib.nc_pp_pil_table(size, 20, None, None, None, 0)
    xmin = np.zeros(20, np.int32)
    cnt = np.zeros(20, np.int32)
    kk = np.zeros(20 * ks, np.int32)
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    assert built_lib.nc_pp_pil_table(size, 20, p(xmin), p(cnt), p(kk), kk.size) == ks
    x0, c0, k0 = R.pil_bicubic_tables(size, 20)
    assert k0.shape[1] == ks
    assert np.array_equal(xmin, x0) and np.array_equal(cnt, c0)
    assert np.array_equal(kk.reshape(20, ks), k0)


@pytest.mark.parametrize("k", [1, 3, 5, 7, 9, 11, 13, 15, 21, 27])
def test_gaussian_q8_equals_oracle(built_lib, k):
    q = np.zeros(k, np.int32)
    assert built_lib.nc_pp_gaussian_q8(k, q.ctypes.data_as(C.c_void_p)) == 0
    assert np.array_equal(q, R.gaussian_kernel_q8(k))
    assert q.sum() == 256
