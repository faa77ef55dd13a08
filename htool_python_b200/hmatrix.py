"""`B200HMatrix`: API-compatible stand-in for the product surface of the reference's `Htool.HMatrix`.

The single-process `HMatrix` has no plugin hook -- `__mul__`/`__matmul__` are hard-wired to the CPU loop
(src/htool/hmatrix/hmatrix.hpp:101-138) -- so the drop-in is a wrapper: `B200HMatrix(hmatrix)` flattens the assembled
reference object once, uploads it, and serves `shape`, `*` and `@` from the GPU with the same argument checks and
error behaviour (RuntimeError on wrong dimension / size, src/htool/hmatrix/hmatrix.hpp:103-108,121-126).

The flat copy is a snapshot: re-create it after `recompression()` / `lu_factorization()` mutate the H-matrix.
"""
from __future__ import annotations

import numpy as np

from .engine import Engine, _is_torch
from .flat import FlatHMatrix


class B200HMatrix:
    def __init__(self, hmatrix, device=None, **options):
        flat = hmatrix if isinstance(hmatrix, FlatHMatrix) else FlatHMatrix.from_htool(hmatrix)
        self.flat_info = dict(n_leaves=flat.n_leaves, stored_coefficients=flat.stored_coefficients, dtype=flat.dtype,
                              algorithmic_bytes=flat.algorithmic_bytes(1), symmetry_for_leaves=flat.symmetry_for_leaves)
        self.engine = Engine(flat, device=device, **options)
        self._np_dtype = flat.np_dtype

    @property
    def shape(self):
        return self.engine.shape

    @property
    def dtype(self):
        return self._np_dtype

    def algorithmic_bytes(self, mu: int = 1) -> float:
        es = 16 if self.engine.is_complex else 8
        return es * (self.flat_info["stored_coefficients"] + mu * (self.shape[0] + self.shape[1]))

    # y = H * x  (user numbering)
    def __mul__(self, x):
        if _is_torch(x):
            if x.dim() != 1:
                raise RuntimeError("Wrong dimension for HMatrix-vector product")
            if x.shape[0] != self.shape[1]:
                raise RuntimeError("Wrong size for HMatrix-vector product")
            return self.engine.matvec_user(x.contiguous())
        x = np.asarray(x)
        if x.ndim != 1:
            raise RuntimeError("Wrong dimension for HMatrix-vector product")
        if x.shape[0] != self.shape[1]:
            raise RuntimeError("Wrong size for HMatrix-vector product")
        x = np.ascontiguousarray(x, dtype=self._np_dtype)  # forcecast, like py::array_t<T, f_style>
        return self.engine.matvec_user(x)

    def matvec(self, x, out=None):
        """`H * x` writing into the caller's `out` (user numbering).  With page-locked numpy buffers (e.g.
        `torch.empty(n, dtype=torch.float64).pin_memory().numpy()`) the library DMAs straight from / into them and the
        call makes no host-side copy at all; pageable buffers go through the pinned staging area."""
        if _is_torch(x):
            return self.engine.matvec_user(x.contiguous(), out)
        x = np.asarray(x)
        if x.ndim != 1:
            raise RuntimeError("Wrong dimension for HMatrix-vector product")
        if x.shape[0] != self.shape[1]:
            raise RuntimeError("Wrong size for HMatrix-vector product")
        if x.dtype != self._np_dtype or not x.flags.c_contiguous:
            x = np.ascontiguousarray(x, dtype=self._np_dtype)
        return self.engine.matvec_user(x, out)

    # Y = H @ X  (user numbering, Fortran-ordered result like the reference)
    def __matmul__(self, X):
        X = np.asarray(X)
        if X.ndim != 2:
            raise RuntimeError("Wrong dimension for HMatrix-matrix product")
        if X.shape[0] != self.shape[1]:
            raise RuntimeError("Wrong size for HMatrix-matrix product")
        X = np.asfortranarray(X, dtype=self._np_dtype)
        return self.engine.matmat_user(X)

    def matmat(self, X, out=None):
        """`H @ X` writing into the caller's Fortran-ordered `out` (user numbering).  With page-locked blocks the library DMAs
        straight from / into them (no host-side copy, no fresh allocation per call); see `matvec`."""
        X = np.asarray(X)
        if X.ndim != 2:
            raise RuntimeError("Wrong dimension for HMatrix-matrix product")
        if X.shape[0] != self.shape[1]:
            raise RuntimeError("Wrong size for HMatrix-matrix product")
        if X.dtype != self._np_dtype or not X.flags.f_contiguous:
            X = np.asfortranarray(X, dtype=self._np_dtype)
        if out is not None and (not isinstance(out, np.ndarray) or out.dtype != self._np_dtype or out.shape != (self.shape[0], X.shape[1]) or not out.flags.f_contiguous):
            raise ValueError("out must be a Fortran-ordered numpy array of the operator's dtype and shape (n_target, mu)")
        return self.engine.matmat_user(X, out)

    def close(self):
        self.engine.close()
