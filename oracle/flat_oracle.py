"""oracle/flat_oracle.py -- TEST INFRASTRUCTURE ONLY: ctypes wrapper of oracle/_ref/libflat_oracle.so (flat_oracle.c).

Parity status: pinned (see the header of flat_oracle.c).  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this module.
"""
from __future__ import annotations

import ctypes
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
_LIB = None


def lib():
    global _LIB
    if _LIB is None:
        p = _HERE / "_ref" / "libflat_oracle.so"
        if not p.exists():
            raise RuntimeError(f"{p} missing: run `make -C oracle _ref/libflat_oracle.so`")
        _LIB = ctypes.CDLL(str(p))
    return _LIB


def _scal(v, flat):
    c = complex(v)
    return (ctypes.c_double * 2)(c.real, c.imag)


def add_vector_product(flat, trans, alpha, x, beta, out):
    """out = beta*out + alpha*op(A)*x on the flat leaf table (cluster numbering), in place."""
    x = np.ascontiguousarray(x, dtype=flat.np_dtype)
    assert out.dtype == flat.np_dtype and out.flags.c_contiguous
    cs = flat.c_struct()
    fn = lib().flat_add_vector_product_z if flat.dtype == "z" else lib().flat_add_vector_product_d
    fn(ctypes.byref(cs), ctypes.c_char(trans.encode()), _scal(alpha, flat), ctypes.c_void_p(x.ctypes.data), _scal(beta, flat), ctypes.c_void_p(out.ctypes.data))
    return out


def add_matrix_product_row_major(flat, trans, alpha, x, beta, out, mu):
    x = np.ascontiguousarray(x, dtype=flat.np_dtype)
    assert out.dtype == flat.np_dtype and out.flags.c_contiguous
    cs = flat.c_struct()
    fn = lib().flat_add_matrix_product_row_major_z if flat.dtype == "z" else lib().flat_add_matrix_product_row_major_d
    fn(ctypes.byref(cs), ctypes.c_char(trans.encode()), _scal(alpha, flat), ctypes.c_void_p(x.ctypes.data), _scal(beta, flat), ctypes.c_void_p(out.ctypes.data), ctypes.c_int(mu))
    return out


def permute_gather(perm, off, n, x):
    perm = np.ascontiguousarray(perm, dtype=np.int32)
    x = np.ascontiguousarray(x)
    out = np.empty(n, dtype=x.dtype)
    fn = lib().flat_permute_gather_z if x.dtype == np.complex128 else lib().flat_permute_gather_d
    fn(ctypes.c_void_p(perm.ctypes.data), ctypes.c_int(off), ctypes.c_int(n), ctypes.c_void_p(x.ctypes.data), ctypes.c_void_p(out.ctypes.data))
    return out


def permute_scatter(perm, off, n, x):
    perm = np.ascontiguousarray(perm, dtype=np.int32)
    x = np.ascontiguousarray(x)
    out = np.empty(n, dtype=x.dtype)
    fn = lib().flat_permute_scatter_z if x.dtype == np.complex128 else lib().flat_permute_scatter_d
    fn(ctypes.c_void_p(perm.ctypes.data), ctypes.c_int(off), ctypes.c_int(n), ctypes.c_void_p(x.ctypes.data), ctypes.c_void_p(out.ctypes.data))
    return out


def matvec_user(flat, x):
    """y = A x in user numbering: gather -> product -> scatter (add_hmatrix_vector_product.hpp:172-197)."""
    xb = permute_gather(flat.source_permutation, flat.source_offset, flat.n_cols, x.astype(flat.np_dtype))
    yb = np.zeros(flat.n_rows, dtype=flat.np_dtype)
    add_vector_product(flat, "N", 1.0, xb, 0.0, yb)
    return permute_scatter(flat.target_permutation, flat.target_offset, flat.n_rows, yb)
