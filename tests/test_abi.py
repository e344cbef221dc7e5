"""The C-ABI library loads and exports every symbol include/lsb200.h declares (no compute calls: CPU-safe)."""
import ctypes
import os
import re

from lidarshooter_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols(header):
    text = open(os.path.join(ROOT, "include", header)).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(lsb(?:200|h)_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    names = declared_symbols("lsb200.h")
    assert len(names) >= 30
    assert sorted(capi.SYMBOLS) == names
    for path in (capi.LIB_PATH, capi.LAB_LIB_PATH):       # the product and the lab build export the same ABI
        L = ctypes.CDLL(path)
        for n in names:
            assert hasattr(L, n), (path, n)


def test_version_and_pure_host_entry_points():
    L = capi.lib()
    assert b"sm_100a" in L.lsb200_version()
    import numpy as np
    A = capi.affine_from_components((1, 2, 3), (0, 0, 0))
    assert np.array_equal(A, np.array([1, 0, 0, 1, 0, 1, 0, 2, 0, 0, 1, 3], np.float32))


def test_create_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        return
    h = ctypes.c_void_p()
    assert capi.lib().lsb200_create(0, ctypes.byref(h)) < 0
    try:
        capi.Context(0)
    except capi.Lsb200Error:
        pass
    else:
        raise AssertionError("Context() must raise without a CUDA device")
