"""Locate and preload the BLAS/LAPACK the reference's header-only core links against.

The htool headers declare the plain Fortran BLAS/LAPACK symbols (lib/htool/include/htool/wrappers/wrapper_blas.hpp:9-63).
This image has no system BLAS; the only library exporting the un-prefixed LP64 symbols is the OpenBLAS bundled with
opencv_python_headless.  Its libgfortran sits next to it without an rpath, so it is loaded explicitly first.
"""
from __future__ import annotations

import ctypes
import importlib.util
import os
from pathlib import Path

_done = False


def openblas_dir() -> Path:
    spec = importlib.util.find_spec("cv2")
    if spec is None or spec.origin is None:
        raise RuntimeError("opencv_python_headless (bundled OpenBLAS) not found")
    return Path(spec.origin).resolve().parent.parent / "opencv_python_headless.libs"


def preload() -> None:
    global _done
    if _done:
        return
    # the bundled OpenBLAS is a pthreads build: nested inside the reference's OpenMP loops it must stay single-threaded
    os.environ.setdefault("OPENBLAS_NUM_THREADS", "1")
    d = openblas_dir()
    for pat in ("libquadmath-*.so*", "libgfortran-*.so*", "libopenblasp-*.so"):
        for lib in sorted(d.glob(pat)):
            ctypes.CDLL(str(lib), mode=ctypes.RTLD_GLOBAL)
    _done = True
