"""Shared test helpers: golden fixtures, synthetic flat H-matrices, the CPU packet emulator, the reference modules."""
from __future__ import annotations

import ctypes
import glob
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
REF_DIR = os.path.join(ROOT, "oracle", "_ref")

from htool_python_b200 import FlatHMatrix, _lib  # noqa: E402

TOL = 1e-12  # BASELINE.json north_star: relative 2-norm, reordered summation allowed


def golden_names():
    return sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "[dz]_*.npz")))  # H-matrix fixtures


def load_golden(name):
    path = os.path.join(GOLDEN, name + ".npz")
    return FlatHMatrix.load_npz(path), dict(np.load(path, allow_pickle=False))


def rel_err(a, b):
    nb = np.linalg.norm(b)
    return float(np.linalg.norm(np.asarray(a) - np.asarray(b)) / (nb if nb > 0 else 1.0))


def rand_vec(rng, n, dtype, cols=None):
    shape = (n,) if cols is None else (n, cols)
    v = rng.random(shape) - 0.5
    if np.dtype(dtype) == np.complex128:
        v = v + 1j * (rng.random(shape) - 0.5)
    return v.astype(dtype)


# ------------------------------------------------------------------------------------------------------------------
# synthetic flat H-matrices: a random recursive tiling of an M x N block into dense / low-rank leaves
# ------------------------------------------------------------------------------------------------------------------
def synthetic_flat(rng, n_rows, n_cols, dtype="d", min_block=1, max_leaf=48, p_lowrank=0.5, max_rank=9, allow_rank0=True, allow_empty=False, p_stop=0.5):
    leaves = []

    def rec(r0, m, c0, n, depth):
        if m == 0 or n == 0:
            if allow_empty:
                leaves.append((r0, m, c0, n))
            return
        if (m <= max_leaf and n <= max_leaf and rng.random() < p_stop) or (m <= min_block or n <= min_block) or depth > 12:
            leaves.append((r0, m, c0, n))
            return
        ms = int(rng.integers(1, m)) if m > 1 else m
        ns = int(rng.integers(1, n)) if n > 1 else n
        for (rr, mm) in ((r0, ms), (r0 + ms, m - ms)):
            for (cc, nn) in ((c0, ns), (c0 + ns, n - ns)):
                if mm > 0 and nn > 0:
                    rec(rr, mm, cc, nn, depth + 1)

    rec(0, n_rows, 0, n_cols, 0)
    nl = len(leaves)
    np_dt = np.float64 if dtype == "d" else np.complex128
    row_off = np.array([l[0] for l in leaves], np.int32)
    nrows = np.array([l[1] for l in leaves], np.int32)
    col_off = np.array([l[2] for l in leaves], np.int32)
    ncols = np.array([l[3] for l in leaves], np.int32)
    rank = np.full(nl, -1, np.int32)
    for i in range(nl):
        if rng.random() < p_lowrank:
            lo = 0 if allow_rank0 else 1
            rank[i] = int(rng.integers(lo, max_rank + 1))
    sizes_a = np.where(rank < 0, nrows.astype(np.int64) * ncols, nrows.astype(np.int64) * np.maximum(rank, 0))
    sizes_b = np.where(rank < 0, 0, np.maximum(rank, 0).astype(np.int64) * ncols)
    a_off = np.concatenate([[0], np.cumsum(sizes_a)])
    b_off = a_off[-1] + np.concatenate([[0], np.cumsum(sizes_b)])
    coef = rand_vec(rng, int(b_off[-1]) + 1, np_dt)
    stored = float(np.sum(np.where(rank < 0, nrows.astype(np.float64) * ncols, np.maximum(rank, 0) * (nrows + ncols).astype(np.float64))))
    f = FlatHMatrix(dtype=dtype, n_rows=n_rows, n_cols=n_cols, row_off=row_off, col_off=col_off, nrows=nrows, ncols=ncols, rank=rank,
                    flags=np.zeros(nl, np.uint8), a_ptr=np.zeros(nl, np.uint64), b_ptr=np.zeros(nl, np.uint64), symmetry_for_leaves="N",
                    target_permutation=rng.permutation(n_rows).astype(np.int32), source_permutation=rng.permutation(n_cols).astype(np.int32),
                    stored_coefficients=stored)
    return f.with_coefficients(coef, a_off[:-1], b_off[:-1])


def flat_to_dense(flat):
    """dense matrix of a flat table with symmetry 'N' (numpy, tiny cases only)."""
    A = np.zeros((flat.n_rows, flat.n_cols), dtype=flat.np_dtype)
    for l in range(flat.n_leaves):
        a, b = flat.leaf_arrays(l)
        blk = a if b is None else a @ b
        r0, c0 = int(flat.row_off[l]), int(flat.col_off[l])
        A[r0:r0 + blk.shape[0], c0:c0 + blk.shape[1]] += blk
    return A


# ------------------------------------------------------------------------------------------------------------------
# CPU packet emulator (tests/native/plan_emulator.cpp): planner + packer of the product, interpreted on the host
# ------------------------------------------------------------------------------------------------------------------
_emu = None


def emulator():
    global _emu
    if _emu is None:
        so = os.path.join(ROOT, "tests", "native", "libplan_emulator.so")
        srcs = [os.path.join(ROOT, "tests", "native", "plan_emulator.cpp"), os.path.join(ROOT, "htool_python_b200", "csrc", "capi", "planner.cpp")]
        if not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
            gxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
            subprocess.check_call([gxx, "-O2", "-std=c++17", "-fPIC", "-shared", "-fopenmp", "-I", os.path.join(ROOT, "include"),
                                   "-I", os.path.join(ROOT, "htool_python_b200", "csrc", "capi"), *srcs, "-o", so])
        _emu = ctypes.CDLL(so)
        _emu.plan_emulate_error.restype = ctypes.c_char_p
    return _emu


def emulate(flat, op, alpha, x, beta, y, **opts):
    """y = beta*y + alpha*op(A)*x through the packed streams, on the CPU; returns plan statistics."""
    emu = emulator()
    o = _lib.hmb200_options()
    for k, v in opts.items():
        setattr(o, k, v)
    cs = flat.c_struct()
    stats = (ctypes.c_int64 * 16)()
    a = (ctypes.c_double * 2)(complex(alpha).real, complex(alpha).imag)
    b = (ctypes.c_double * 2)(complex(beta).real, complex(beta).imag)
    x = np.ascontiguousarray(x, dtype=flat.np_dtype)
    rc = emu.plan_emulate(ctypes.byref(cs), flat.elem_bytes, ctypes.c_char(op.encode()), ctypes.byref(o), a,
                          ctypes.c_void_p(x.ctypes.data), b, ctypes.c_void_p(y.ctypes.data), stats)
    if rc:
        raise RuntimeError(emu.plan_emulate_error().decode())
    keys = ("stream1_bytes", "stream2_bytes", "pieces1", "pieces2", "stages1", "stages2", "reduce", "bins", "xz_elems", "stored", "zcopies1", "zcopies2")
    return dict(zip(keys, list(stats)))


# ------------------------------------------------------------------------------------------------------------------
# the reference itself (only in the build container: oracle/_ref/Htool*.so needs nothing else, but is optional)
# ------------------------------------------------------------------------------------------------------------------
def reference_modules():
    """(Htool, htool_ref_ops) or None when oracle/_ref has not been built."""
    if not glob.glob(os.path.join(REF_DIR, "Htool*.so")) or not glob.glob(os.path.join(REF_DIR, "htool_ref_ops*.so")):
        return None
    from htool_python_b200 import _blas

    _blas.preload()
    if REF_DIR not in sys.path:
        sys.path.insert(0, REF_DIR)
    import Htool
    import htool_ref_ops

    return Htool, htool_ref_ops


def bridge_available():
    try:
        from htool_python_b200 import bridge

        bridge.module()
        return True
    except Exception:
        return False


def emulate_mm(flat, op, alpha, X, beta, Y, **opts):
    """Y = beta*Y + alpha*op(A)*X (row-major n x mu) through the multi-RHS packet streams, on the CPU."""
    emu = emulator()
    o = _lib.hmb200_options()
    for k, v in opts.items():
        setattr(o, k, v)
    cs = flat.c_struct()
    stats = (ctypes.c_int64 * 16)()
    a = (ctypes.c_double * 2)(complex(alpha).real, complex(alpha).imag)
    b = (ctypes.c_double * 2)(complex(beta).real, complex(beta).imag)
    X = np.ascontiguousarray(X, dtype=flat.np_dtype)
    assert Y.flags.c_contiguous and Y.dtype == flat.np_dtype
    rc = emu.plan_emulate_mm(ctypes.byref(cs), flat.elem_bytes, ctypes.c_char(op.encode()), ctypes.byref(o), a,
                             ctypes.c_void_p(X.ctypes.data), b, ctypes.c_void_p(Y.ctypes.data), int(X.shape[1]), stats)
    if rc:
        raise RuntimeError(emu.plan_emulate_error().decode())
    keys = ("mm1_bytes", "mm2_bytes", "pieces_mm1", "pieces_mm2", "stages_mm1", "stages_mm2", "reduce_mm", "units_mm1", "mm_rows", "stored", "zcopies_mm1", "zcopies_mm2")
    return dict(zip(keys, list(stats)))
