"""Python handle on one packed, device-resident operator (libhmb200.so `hmb200_handle`).

Host arrays (numpy) go through the library's pinned staging (HMB200_HOST, synchronous like the reference's CPU calls);
torch CUDA tensors are passed as device pointers on torch's current stream (HMB200_DEVICE, asynchronous).
"""
from __future__ import annotations

import ctypes
from typing import Optional

import numpy as np

from . import _lib
from .flat import FlatHMatrix


def _is_torch(x) -> bool:
    return type(x).__module__.startswith("torch")


def _scalar(v, is_complex: bool):
    c = complex(v)
    if not is_complex and c.imag != 0:
        raise ValueError("complex scalar passed to a real operator")
    return (ctypes.c_double * 2)(c.real, c.imag)


class Engine:
    def __init__(self, flat: FlatHMatrix, device: Optional[int] = None, stream: Optional[int] = None, **options):
        self._h = None
        lib = _lib.lib()
        self.flat_meta = dict(dtype=flat.dtype, n_rows=flat.n_rows, n_cols=flat.n_cols, stored_coefficients=flat.stored_coefficients,
                              symmetry_for_leaves=flat.symmetry_for_leaves)
        self.dtype = flat.dtype
        self.is_complex = flat.dtype == "z"
        self.np_dtype = flat.np_dtype
        self.shape = (flat.n_rows, flat.n_cols)
        if device is None:
            import torch

            if not torch.cuda.is_available():
                raise RuntimeError("hmb200: no CUDA device available (the engine has no CPU fallback)")
            device = torch.cuda.current_device()
        self.device = int(device)
        opt = _lib.hmb200_options()
        for k, v in options.items():
            if not hasattr(opt, k):
                raise TypeError(f"unknown engine option {k!r}")
            setattr(opt, k, int(v))
        cs = flat.c_struct()
        create = lib.hmb200_create_z if self.is_complex else lib.hmb200_create_d
        h = create(ctypes.byref(cs), self.device, ctypes.c_void_p(stream or 0), ctypes.byref(opt))
        if not h:
            raise RuntimeError("hmb200_create failed: " + _lib.last_error())
        self._h = ctypes.c_void_p(h)
        self.target_offset, self.source_offset = int(flat.target_offset or 0), int(flat.source_offset or 0)
        _lib.check(lib.hmb200_set_cluster_offsets(self._h, self.target_offset, self.source_offset))
        self._has_perms = False
        # user-numbering products need the root block of both trees (a row slab's slots map to arbitrary user indices)
        if (flat.target_permutation is not None and len(flat.target_permutation) == flat.n_rows and flat.target_offset == 0
                and flat.source_permutation is not None and len(flat.source_permutation) == flat.n_cols and flat.source_offset == 0):
            tp, sp = flat.local_target_permutation(), flat.local_source_permutation()
            _lib.check(lib.hmb200_set_permutations(self._h, tp.ctypes.data, sp.ctypes.data))
            self._has_perms = True

    # ---------------------------------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None):
            _lib.lib().hmb200_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def info(self) -> dict:
        i = _lib.hmb200_info()
        _lib.check(_lib.lib().hmb200_get_info(self._h, ctypes.byref(i)))
        return {n: getattr(i, n) for n, _ in i._fields_}

    def enable_timing(self, on: bool = True):
        _lib.check(_lib.lib().hmb200_enable_timing(self._h, int(on)))

    def last_kernel_ms(self):
        ms = (ctypes.c_float * 4)()
        _lib.check(_lib.lib().hmb200_last_kernel_ms(self._h, ctypes.byref(ms)))
        return dict(phase1=ms[0], reduce=ms[1], phase2=ms[2], total=ms[3])

    def kernel_ms_avg(self):
        """per-kernel CUDA-event times averaged over the products issued since enable_timing(True)"""
        ms, n = (ctypes.c_float * 4)(), ctypes.c_int(0)
        _lib.check(_lib.lib().hmb200_kernel_ms_avg(self._h, ctypes.byref(ms), ctypes.byref(n)))
        return dict(phase1=ms[0], reduce=ms[1], phase2=ms[2], total=ms[3], products=n.value)

    @property
    def launch_count(self) -> int:
        return int(_lib.lib().hmb200_launch_count(self._h))

    def set_stream(self, stream: int):
        _lib.check(_lib.lib().hmb200_set_stream(self._h, ctypes.c_void_p(stream)))

    # ---------------------------------------------------------------------------------------------------
    def _sizes(self, trans: str):
        n_out, n_in = self.shape if trans == "N" else self.shape[::-1]
        return n_in, n_out

    def _ptr(self, a, n_elems: int, writable: bool):
        """returns (pointer, mem) after validating dtype / contiguity / size; no copies are made here."""
        if _is_torch(a):
            import torch

            want = torch.complex128 if self.is_complex else torch.float64
            if a.dtype != want or not a.is_contiguous() or a.numel() != n_elems or not a.is_cuda:
                raise ValueError(f"expected a contiguous CUDA tensor of {n_elems} {want} elements")
            if a.device.index != self.device:
                raise ValueError("tensor lives on another device than the engine")
            self.set_stream(torch.cuda.current_stream(a.device).cuda_stream)
            return a.data_ptr(), _lib.DEVICE
        if not isinstance(a, np.ndarray) or a.dtype != self.np_dtype or a.size != n_elems or not (a.flags.c_contiguous or a.flags.f_contiguous):
            raise ValueError(f"expected a contiguous numpy array of {n_elems} {np.dtype(self.np_dtype).name} elements")
        if writable and not a.flags.writeable:
            raise ValueError("output array is read-only")
        return a.ctypes.data, _lib.HOST

    def add_vector_product(self, trans: str, alpha, x, beta, out):
        """out = beta*out + alpha*op(A)*x, cluster numbering, IN PLACE (operator convention of the reference,
        core/distributed_operator/interfaces/virtual_global_to_local_operator.hpp:16)."""
        n_in, n_out = self._sizes(trans)
        px, mx = self._ptr(x, n_in, False)
        po, mo = self._ptr(out, n_out, True)
        if mx != mo:
            raise ValueError("in and out must both be host arrays or both CUDA tensors")
        fn = _lib.lib().hmb200_add_vector_product_z if self.is_complex else _lib.lib().hmb200_add_vector_product_d
        _lib.check(fn(self._h, trans.encode(), _scalar(alpha, self.is_complex), px, _scalar(beta, self.is_complex), po, mx))
        return out

    def add_matrix_product_row_major(self, trans: str, alpha, x, beta, out, mu: Optional[int] = None):
        """row-major blocks (rows x mu), in place (virtual_global_to_local_operator.hpp:25)."""
        n_in, n_out = self._sizes(trans)
        if mu is None:
            mu = x.shape[1] if x.ndim == 2 else x.size // max(n_in, 1)
        if not _is_torch(x) and x.ndim == 2 and not x.flags.c_contiguous:
            raise ValueError("row-major (C-contiguous) blocks expected")
        px, mx = self._ptr(x, n_in * mu, False)
        po, mo = self._ptr(out, n_out * mu, True)
        if mx != mo:
            raise ValueError("in and out must both be host arrays or both CUDA tensors")
        fn = _lib.lib().hmb200_add_matrix_product_row_major_z if self.is_complex else _lib.lib().hmb200_add_matrix_product_row_major_d
        _lib.check(fn(self._h, trans.encode(), _scalar(alpha, self.is_complex), px, _scalar(beta, self.is_complex), po, int(mu), mx))
        return out

    def add_sub_matrix_product_to_local(self, x, out, mu: int, offset: int, size: int):
        """out += A[:, offset:offset+size] * x, IN PLACE; x = (size x mu) row-major window of the source in source-tree
        numbering (virtual_global_to_local_operator.hpp:33; restricted_operator.hpp:173-194)."""
        if size < 0 or mu <= 0:
            raise ValueError("size must be >= 0 and mu > 0")
        px, mx = self._ptr(x, size * mu, False) if size else (0, None)
        po, mo = self._ptr(out, self.shape[0] * mu, True)
        if size and mx != mo:
            raise ValueError("in and out must both be host arrays or both CUDA tensors")
        fn = _lib.lib().hmb200_add_sub_matrix_product_to_local_z if self.is_complex else _lib.lib().hmb200_add_sub_matrix_product_to_local_d
        _lib.check(fn(self._h, px, po, int(mu), int(offset), int(size), mo))
        return out

    def matvec_user(self, x, out=None):
        """y = A x in USER numbering (HMatrix.__mul__, src/htool/hmatrix/hmatrix.hpp:101-117)."""
        if not self._has_perms:
            raise RuntimeError("engine was created without permutations")
        if out is None:
            out = self._empty_like(x, self.shape[0])
        px, mx = self._ptr(x, self.shape[1], False)
        po, mo = self._ptr(out, self.shape[0], True)
        if mx != mo:
            raise ValueError("in and out must both be host arrays or both CUDA tensors")
        fn = _lib.lib().hmb200_matvec_user_z if self.is_complex else _lib.lib().hmb200_matvec_user_d
        _lib.check(fn(self._h, px, po, mx))
        return out

    def matmat_user(self, X, out=None):
        """Y = A X, user numbering, column-major n x mu blocks (HMatrix.__matmul__, hmatrix.hpp:119-138)."""
        if not self._has_perms:
            raise RuntimeError("engine was created without permutations")
        mu = X.shape[1]
        if _is_torch(X):
            raise NotImplementedError("matmat_user takes Fortran-ordered numpy arrays; use add_matrix_product_row_major for CUDA tensors")
        if not X.flags.f_contiguous:
            raise ValueError("Fortran-ordered (column-major) input expected")
        if out is None:
            out = np.zeros((self.shape[0], mu), dtype=self.np_dtype, order="F")
        px, mx = self._ptr(X, self.shape[1] * mu, False)
        po, mo = self._ptr(out, self.shape[0] * mu, True)
        fn = _lib.lib().hmb200_matmat_user_z if self.is_complex else _lib.lib().hmb200_matmat_user_d
        _lib.check(fn(self._h, px, po, int(mu), mx))
        return out

    def _empty_like(self, x, n):
        if _is_torch(x):
            import torch

            return torch.zeros(n, dtype=x.dtype, device=x.device)
        return np.zeros(n, dtype=self.np_dtype)


def permute_gather(perm: np.ndarray, off: int, n: int, x: np.ndarray, mu: int = 1) -> np.ndarray:
    """out[i] = x[perm[off+i]-off] on the device (user_to_cluster, clustering/cluster_node.hpp:162-175); host arrays."""
    perm = np.ascontiguousarray(perm, dtype=np.int32)
    x = np.ascontiguousarray(x) if x.ndim == 1 else np.asfortranarray(x)
    out = np.empty_like(x, order="F" if x.ndim == 2 else "C")
    _lib.check(_lib.lib().hmb200_permute_gather(perm.ctypes.data, off, n, x.ctypes.data, out.ctypes.data, x.dtype.itemsize, mu, _lib.HOST, None))
    return out


def permute_scatter(perm: np.ndarray, off: int, n: int, x: np.ndarray, mu: int = 1) -> np.ndarray:
    """out[perm[off+i]-off] = x[i] on the device (cluster_to_user, clustering/cluster_node.hpp:149-161); host arrays."""
    perm = np.ascontiguousarray(perm, dtype=np.int32)
    x = np.ascontiguousarray(x) if x.ndim == 1 else np.asfortranarray(x)
    out = np.empty_like(x, order="F" if x.ndim == 2 else "C")
    _lib.check(_lib.lib().hmb200_permute_scatter(perm.ctypes.data, off, n, x.ctypes.data, out.ctypes.data, x.dtype.itemsize, mu, _lib.HOST, None))
    return out
