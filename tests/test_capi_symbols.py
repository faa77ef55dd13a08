"""The C-ABI library loads and exports every symbol include/hmb200.h declares; without a GPU it fails loudly."""
import ctypes
import os
import re

import numpy as np
import pytest

import helpers
from htool_python_b200 import _lib

HEADER = os.path.join(helpers.ROOT, "include", "hmb200.h")


def declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(hmb200_[a-z_0-9]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(str(_lib.LIB_PATH))
    names = declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), n
    assert set(names) == set(_lib.EXPORTS), set(names) ^ set(_lib.EXPORTS)
    assert b"sm_100a" in _lib.lib().hmb200_version()


def test_struct_layouts_match_header():
    from htool_python_b200.flat import hmb200_flat

    assert ctypes.sizeof(hmb200_flat) == 96
    assert ctypes.sizeof(_lib.hmb200_options) == 64
    assert ctypes.sizeof(_lib.hmb200_info) == 152


def test_no_cpu_fallback_without_gpu():
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from htool_python_b200 import Engine

    rng = np.random.default_rng(0)
    flat = helpers.synthetic_flat(rng, 40, 40)
    with pytest.raises(RuntimeError, match="no CUDA device|no CPU fallback"):
        Engine(flat)
    with pytest.raises(RuntimeError, match="no CUDA device"):
        Engine(flat, device=0)


def test_host_flags_publish_and_wait():
    """hmb200_host_flag_publish / hmb200_host_flags_wait (host code only): strided words, a late publisher, the timeout"""
    import threading
    import time

    lib = _lib.lib()
    words = np.zeros(4 * 8, np.int64)
    fl = words[::8]  # one 64-byte line per rank, as the operator lays them out
    for q in (0, 1, 3):
        assert lib.hmb200_host_flag_publish(fl.ctypes.data + fl.strides[0] * q, 5) == 0
    assert list(fl) == [5, 5, 0, 5]
    laggard = ctypes.c_int(-1)
    t0 = time.perf_counter()
    assert lib.hmb200_host_flags_wait(fl.ctypes.data, 8, 4, 5, 0.05, ctypes.byref(laggard)) == 2
    assert laggard.value == 2 and 0.04 < time.perf_counter() - t0 < 2.0
    assert "rank 2" in _lib.last_error()
    threading.Timer(0.05, lambda: lib.hmb200_host_flag_publish(fl.ctypes.data + fl.strides[0] * 2, 6)).start()
    assert lib.hmb200_host_flags_wait(fl.ctypes.data, 8, 4, 5, 10.0, ctypes.byref(laggard)) == 0  # >= value counts
    assert lib.hmb200_host_flags_wait(fl.ctypes.data, 8, 0, 5, 0.0, None) == 0
