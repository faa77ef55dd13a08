"""Generates the golden fixtures of tests/golden/ by IMPORTING THE REFERENCE (oracle/_ref/Htool*.so, built from
/root/reference by oracle/Makefile) in the build container.  The reference cannot travel to the GPU box, so the
vectors it produces are committed here together with this script.

Each fixture holds: the flat leaf table of an H-matrix assembled by the reference's own Python API
(Htool.ClusterTreeBuilder / Htool.HMatrixTreeBuilder.build with a Python VirtualGenerator, as tests/test_hmatrix.py of
the reference does), an input x / X and the outputs of the reference's own products:
    y_ref  = H * x    (Htool.HMatrix.__mul__,    src/htool/hmatrix/hmatrix.hpp:101-117, sequential policy)
    Y_ref  = H @ X    (Htool.HMatrix.__matmul__, src/htool/hmatrix/hmatrix.hpp:119-138)
    dense  = H.to_dense_in_user_numbering() @ x  (the reference test's own 1e-10 check, tests/test_hmatrix.py:84)

Run:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))
from htool_python_b200 import _blas  # noqa: E402

_blas.preload()
import Htool  # noqa: E402  (the reference module)

from htool_python_b200 import FlatHMatrix  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def geometry(kind, n, rng):
    if kind == "cylinder":  # use_hmatrix.py-style: points on a cylinder
        th = rng.random(n) * 2 * np.pi
        return np.asfortranarray(np.vstack([np.cos(th), np.sin(th), 4 * rng.random(n)]))
    g = rng.standard_normal((3, n))
    return np.asfortranarray(g / np.linalg.norm(g, axis=0))


def kernel(name, X, Y):
    d = np.linalg.norm(X[:, :, None] - Y[:, None, :], axis=0)
    if name == "invdist":  # example/define_generators.py:15-17 (regularised)
        return 1.0 / (0.1 + d)
    if name == "complex":  # (1+i)/(1e-5+4 pi r): testing/generator_test.hpp:189-196
        return (1 + 1j) / (1e-5 + 4 * np.pi * d)
    if name == "hermitian":  # testing/generator_test.hpp:198-205
        s = np.sign(X[0][:, None] - Y[0][None, :])
        return (1 + 1j * s) / (1e-5 + 4 * np.pi * d)
    raise ValueError(name)


def make(name, kern, geo, n_t, n_s, sym, uplo, eps, eta, leaf, seed, mu=3):
    rng = np.random.default_rng(seed)
    tp = geometry(geo, n_t, rng)
    square = sym != "N" or n_s is None
    sp = tp if square else geometry(geo, n_s, rng)
    is_complex = kern in ("complex", "hermitian")

    Base = Htool.ComplexVirtualGenerator if is_complex else Htool.VirtualGenerator

    class Gen(Base):
        def build_submatrix(self, J, K, mat):
            mat[:] = kernel(kern, tp[:, J], sp[:, K])

    cb = Htool.ClusterTreeBuilder()
    cb.set_maximal_leaf_size(leaf)
    tc = cb.create_cluster_tree(tp, 2)
    sc = tc if square else cb.create_cluster_tree(sp, 2)
    Builder = Htool.ComplexHMatrixTreeBuilder if is_complex else Htool.HMatrixTreeBuilder
    gen = Gen()
    H = Builder(eps, eta, sym, uplo).build(gen, tc, sc)
    flat = FlatHMatrix.from_htool(H).to_owned()
    n_src = sp.shape[1]
    dt = np.complex128 if is_complex else np.float64
    x = (rng.random(n_src) + (1j * rng.random(n_src) if is_complex else 0)).astype(dt)
    X = np.asfortranarray((rng.random((n_src, mu)) + (1j * rng.random((n_src, mu)) if is_complex else 0)).astype(dt))
    y_ref = H * x
    Y_ref = H @ X if square else None  # the reference's __matmul__ sizes its result with X.shape[0]: square only (hmatrix.hpp:127)
    dense = H.to_dense_in_user_numbering()
    y_dense = dense @ x
    assert np.linalg.norm(y_ref - y_dense) / np.linalg.norm(y_dense) < 1e-10  # tests/test_hmatrix.py:84
    info = H.get_local_information()
    path = os.path.join(OUT, name + ".npz")
    flat.save_npz(path)
    z = dict(np.load(path))
    z.update(x=x, X=X, y_ref=y_ref, y_dense=y_dense, compression_ratio=float(info["Compression_ratio"]),
             n_dense=int(info["Number_of_dense_blocks"]), n_low_rank=int(info["Number_of_low_rank_blocks"]))
    if Y_ref is not None:
        z["Y_ref"] = Y_ref
    np.savez_compressed(path, **z)
    print(f"{name}: {flat.n_rows}x{flat.n_cols} leaves={flat.n_leaves} dense={z['n_dense']} lr={z['n_low_rank']} "
          f"CR={z['compression_ratio']:.2f} |H x - dense x|={np.linalg.norm(y_ref - y_dense) / np.linalg.norm(y_dense):.1e} "
          f"size={os.path.getsize(path) / 1e6:.2f} MB")


if __name__ == "__main__":
    #     name                 kernel       geometry    n_t  n_s   sym  uplo eps   eta  leaf seed
    make("d_N_cylinder",      "invdist",   "cylinder", 700, None, "N", "N", 1e-3, 10., 16, 0)
    make("d_N_rect",          "invdist",   "sphere",   600, 350,  "N", "N", 1e-6, 10., 10, 1)
    make("d_SL_cylinder",     "invdist",   "cylinder", 700, None, "S", "L", 1e-3, 10., 16, 2)
    make("d_SU_sphere",       "invdist",   "sphere",   500, None, "S", "U", 1e-6, 100., 10, 3)
    make("z_N_sphere",        "complex",   "sphere",   500, None, "N", "N", 1e-3, 10., 16, 4)
    make("z_SL_sphere",       "complex",   "sphere",   450, None, "S", "L", 1e-4, 10., 12, 5)
    make("z_HU_sphere",       "hermitian", "sphere",   450, None, "H", "U", 1e-4, 10., 12, 6)
    make("z_HL_cylinder",     "hermitian", "cylinder", 400, None, "H", "L", 1e-3, 10., 20, 7)
