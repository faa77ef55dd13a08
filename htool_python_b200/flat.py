"""Flat leaf table of an assembled H-matrix (host side of include/hmb200.h `hmb200_flat`).

A `FlatHMatrix` is the hand-over format between the reference's `htool::HMatrix` and the engine: per leaf its
target/source range, rank, symmetry flags and the ADDRESS of its U/V/dense buffer (zero-copy view into the reference's
memory), or -- for fixtures and files -- offsets into one owned coefficient array.
Reference: leaves as enumerated by get_leaves_from (lib/htool/include/htool/hmatrix/hmatrix.hpp:247-274).
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass, field
from typing import Any, Optional

import numpy as np

MIRROR = 1


class hmb200_flat(ctypes.Structure):
    _fields_ = [
        ("n_rows", ctypes.c_int32),
        ("n_cols", ctypes.c_int32),
        ("n_leaves", ctypes.c_int64),
        ("row_off", ctypes.c_void_p),
        ("col_off", ctypes.c_void_p),
        ("nrows", ctypes.c_void_p),
        ("ncols", ctypes.c_void_p),
        ("rank", ctypes.c_void_p),
        ("flags", ctypes.c_void_p),
        ("a", ctypes.c_void_p),
        ("b", ctypes.c_void_p),
        ("symmetry_for_leaves", ctypes.c_int32),
        ("n_row_cuts", ctypes.c_int32),
        ("row_cuts", ctypes.c_void_p),
    ]


def _c(a, dtype):
    return np.ascontiguousarray(a, dtype=dtype)


@dataclass
class FlatHMatrix:
    dtype: str  # "d" or "z"
    n_rows: int
    n_cols: int
    row_off: np.ndarray
    col_off: np.ndarray
    nrows: np.ndarray
    ncols: np.ndarray
    rank: np.ndarray
    flags: np.ndarray
    a_ptr: np.ndarray  # uint64 host addresses
    b_ptr: np.ndarray
    symmetry_for_leaves: str = "N"
    target_offset: int = 0
    source_offset: int = 0
    target_permutation: Optional[np.ndarray] = None  # full tree permutation (cluster slot -> user index)
    source_permutation: Optional[np.ndarray] = None
    stored_coefficients: float = 0.0
    row_cuts: Optional[np.ndarray] = None
    keepalive: Any = field(default=None, repr=False)  # object owning the memory a_ptr/b_ptr point into

    # ------------------------------------------------------------------------------------------------------
    @property
    def np_dtype(self):
        return np.float64 if self.dtype == "d" else np.complex128

    @property
    def elem_bytes(self) -> int:
        return 8 if self.dtype == "d" else 16

    @property
    def n_leaves(self) -> int:
        return int(self.rank.shape[0])

    @property
    def shape(self):
        return (self.n_rows, self.n_cols)

    def algorithmic_bytes(self, mu: int = 1) -> float:
        """sizeof(T)*(W + mu*(n_src+n_tgt))  -- SURVEY.md section 8(d)."""
        return self.elem_bytes * (self.stored_coefficients + mu * (self.n_rows + self.n_cols))

    def local_target_permutation(self) -> np.ndarray:
        """perm[off+i]-off for the root block's target range (cluster_node.hpp:149-175)."""
        o = self.target_offset
        return _c(self.target_permutation[o:o + self.n_rows] - o, np.int32)

    def local_source_permutation(self) -> np.ndarray:
        o = self.source_offset
        return _c(self.source_permutation[o:o + self.n_cols] - o, np.int32)

    # ------------------------------------------------------------------------------------------------------
    @classmethod
    def from_view(cls, view: dict, keepalive=None) -> "FlatHMatrix":
        """`view` = dict returned by _htool_bridge.flat_view."""
        return cls(
            dtype=view["dtype"], n_rows=int(view["n_rows"]), n_cols=int(view["n_cols"]),
            row_off=_c(view["row_off"], np.int32), col_off=_c(view["col_off"], np.int32),
            nrows=_c(view["nrows"], np.int32), ncols=_c(view["ncols"], np.int32), rank=_c(view["rank"], np.int32),
            flags=_c(view["flags"], np.uint8), a_ptr=_c(view["a_ptr"], np.uint64), b_ptr=_c(view["b_ptr"], np.uint64),
            symmetry_for_leaves=str(view["symmetry_for_leaves"]),
            target_offset=int(view["target_offset"]), source_offset=int(view["source_offset"]),
            target_permutation=_c(view["target_permutation"], np.int32),
            source_permutation=_c(view["source_permutation"], np.int32),
            stored_coefficients=float(view["stored_coefficients"]), row_cuts=_c(view["row_cuts"], np.int32),
            keepalive=keepalive,
        )

    @classmethod
    def from_htool(cls, hmatrix) -> "FlatHMatrix":
        """Flatten the reference's Htool.HMatrix / Htool.ComplexHMatrix or a bridge NativeHMatrix (zero copy)."""
        from . import bridge

        br = bridge.module()
        return cls.from_view(br.flat_view(hmatrix), keepalive=hmatrix)

    # ------------------------------------------------------------------------------------------------------
    def leaf_arrays(self, l: int):
        """(A, B) numpy views of leaf l: dense -> (D m x n, None); low-rank -> (U m x r, V r x n). Column-major."""
        m, n, r = int(self.nrows[l]), int(self.ncols[l]), int(self.rank[l])
        dt = self.np_dtype

        def view(addr, rows, cols):
            if rows * cols == 0:
                return np.zeros((rows, cols), dtype=dt, order="F")
            buf = (ctypes.c_char * (rows * cols * self.elem_bytes)).from_address(int(addr))
            return np.frombuffer(buf, dtype=dt).reshape((rows, cols), order="F")

        if r < 0:
            return view(self.a_ptr[l], m, n), None
        return view(self.a_ptr[l], m, r), view(self.b_ptr[l], r, n)

    def to_owned(self) -> "FlatHMatrix":
        """Deep copy of the coefficients into one owned array (for fixtures / serialisation)."""
        sizes_a = np.where(self.rank < 0, self.nrows.astype(np.int64) * self.ncols, self.nrows.astype(np.int64) * np.maximum(self.rank, 0))
        sizes_b = np.where(self.rank < 0, 0, np.maximum(self.rank, 0).astype(np.int64) * self.ncols)
        a_off = np.concatenate([[0], np.cumsum(sizes_a)]).astype(np.int64)
        b_off = a_off[-1] + np.concatenate([[0], np.cumsum(sizes_b)]).astype(np.int64)
        coef = np.zeros(int(b_off[-1]) + 1, dtype=self.np_dtype)
        for l in range(self.n_leaves):
            A, B = self.leaf_arrays(l)
            coef[a_off[l]:a_off[l + 1]] = A.reshape(-1, order="F")
            if B is not None:
                coef[b_off[l]:b_off[l + 1]] = B.reshape(-1, order="F")
        return self.with_coefficients(coef, a_off[:-1], b_off[:-1])

    def with_coefficients(self, coef: np.ndarray, a_off: np.ndarray, b_off: np.ndarray) -> "FlatHMatrix":
        coef = np.ascontiguousarray(coef, dtype=self.np_dtype)
        base = coef.ctypes.data
        a_ptr = (base + _c(a_off, np.int64) * self.elem_bytes).astype(np.uint64)
        b_ptr = np.where(self.rank < 0, 0, base + _c(b_off, np.int64) * self.elem_bytes).astype(np.uint64)
        out = FlatHMatrix(**{**self.__dict__, "a_ptr": a_ptr, "b_ptr": b_ptr, "keepalive": (coef, _c(a_off, np.int64), _c(b_off, np.int64))})
        return out

    def save_npz(self, path) -> None:
        owned = self if (isinstance(self.keepalive, tuple) and len(self.keepalive) == 3) else self.to_owned()
        coef, a_off, b_off = owned.keepalive
        np.savez_compressed(
            path, dtype=np.array(self.dtype), n_rows=self.n_rows, n_cols=self.n_cols, row_off=self.row_off, col_off=self.col_off,
            nrows=self.nrows, ncols=self.ncols, rank=self.rank, flags=self.flags, coef=coef, a_off=a_off, b_off=b_off,
            symmetry_for_leaves=np.array(self.symmetry_for_leaves), target_offset=self.target_offset, source_offset=self.source_offset,
            target_permutation=self.target_permutation if self.target_permutation is not None else np.zeros(0, np.int32),
            source_permutation=self.source_permutation if self.source_permutation is not None else np.zeros(0, np.int32),
            stored_coefficients=self.stored_coefficients,
        )

    @classmethod
    def load_npz(cls, path) -> "FlatHMatrix":
        z = np.load(path, allow_pickle=False)
        nl = z["rank"].shape[0]
        f = cls(
            dtype=str(z["dtype"]), n_rows=int(z["n_rows"]), n_cols=int(z["n_cols"]), row_off=_c(z["row_off"], np.int32),
            col_off=_c(z["col_off"], np.int32), nrows=_c(z["nrows"], np.int32), ncols=_c(z["ncols"], np.int32),
            rank=_c(z["rank"], np.int32), flags=_c(z["flags"], np.uint8), a_ptr=np.zeros(nl, np.uint64), b_ptr=np.zeros(nl, np.uint64),
            symmetry_for_leaves=str(z["symmetry_for_leaves"]), target_offset=int(z["target_offset"]), source_offset=int(z["source_offset"]),
            target_permutation=_c(z["target_permutation"], np.int32), source_permutation=_c(z["source_permutation"], np.int32),
            stored_coefficients=float(z["stored_coefficients"]),
        )
        return f.with_coefficients(z["coef"], z["a_off"], z["b_off"])

    # ------------------------------------------------------------------------------------------------------
    def c_struct(self) -> hmb200_flat:
        """ctypes view (valid while `self` is alive)."""
        s = hmb200_flat()
        s.n_rows, s.n_cols, s.n_leaves = self.n_rows, self.n_cols, self.n_leaves
        for name in ("row_off", "col_off", "nrows", "ncols", "rank", "flags"):
            setattr(s, name, getattr(self, name).ctypes.data)
        s.a = self.a_ptr.ctypes.data
        s.b = self.b_ptr.ctypes.data
        s.symmetry_for_leaves = ord(self.symmetry_for_leaves)
        if self.row_cuts is not None and len(self.row_cuts):
            s.n_row_cuts = len(self.row_cuts)
            s.row_cuts = self.row_cuts.ctypes.data
        return s
