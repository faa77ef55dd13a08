"""Pins the oracle against the REAL reference running here (oracle/_ref, compiled from /root/reference):
same htool::HMatrix object, reference CPU products vs the flat C restatement.  Skipped where oracle/_ref is absent."""
import numpy as np
import pytest

import helpers
from helpers import TOL, rel_err
from oracle import flat_oracle as fo

mods = helpers.reference_modules() if helpers.bridge_available() else None
pytestmark = pytest.mark.skipif(mods is None, reason="oracle/_ref (reference build) or the bridge is not available")

CASES = [
    ("invdist_reg", "N", "N", 1500, 1e-3, 16),
    ("laplace_reg", "S", "L", 1500, 1e-4, 32),
    ("laplace_reg", "S", "U", 1200, 1e-6, 10),
    ("complex_reg", "N", "N", 1000, 1e-3, 16),
    ("complex_reg", "S", "U", 1000, 1e-4, 16),
    ("complex_herm", "H", "L", 1000, 1e-4, 16),
    ("helmholtz", "N", "N", 1200, 1e-3, 32),
]


def _build(kernel, sym, uplo, n, eps, leaf, seed=0, partition=1, slab=-1):
    from htool_python_b200 import FlatHMatrix, bridge

    rng = np.random.default_rng(seed)
    g = rng.standard_normal((3, n))
    pts = np.asfortranarray(g / np.linalg.norm(g, axis=0))
    tree = bridge.create_cluster_tree(pts, 2, partition, leaf)
    H = bridge.build_hmatrix(kernel, tree, epsilon=eps, eta=10.0, symmetry=sym, UPLO=uplo, wavenumber=5.0,
                             target_partition_number=slab, partition_number_for_symmetry=slab)
    return H, FlatHMatrix.from_htool(H), rng


@pytest.mark.parametrize("kernel,sym,uplo,n,eps,leaf", CASES)
def test_oracle_equals_reference_internal_product(kernel, sym, uplo, n, eps, leaf):
    _, ro = mods
    H, flat, rng = _build(kernel, sym, uplo, n, eps, leaf)
    alpha, beta = (1.5, -0.5) if flat.dtype == "d" else (1.5 - 0.3j, 0.2 + 0.7j)
    ops = ["N"] + (["T"] if sym != "H" else []) + (["C"] if flat.dtype == "z" and sym != "S" else [])
    for op in ops:
        x = helpers.rand_vec(rng, flat.n_rows if op != "N" else flat.n_cols, flat.np_dtype)
        y0 = helpers.rand_vec(rng, flat.n_cols if op != "N" else flat.n_rows, flat.np_dtype)
        for parallel in (False, True):
            y_ref = y0.copy()
            getattr(ro, "internal_vector_product_addr_" + flat.dtype)(H.address, op, alpha, x, beta, y_ref, parallel)
            y = y0.copy()
            fo.add_vector_product(flat, op, alpha, x, beta, y)
            assert rel_err(y, y_ref) < TOL, (op, parallel)


@pytest.mark.parametrize("kernel,sym,uplo,n,eps,leaf", CASES[:5])
def test_oracle_equals_reference_user_products(kernel, sym, uplo, n, eps, leaf):
    _, ro = mods
    H, flat, rng = _build(kernel, sym, uplo, n, eps, leaf, seed=1)
    x = helpers.rand_vec(rng, flat.n_cols, flat.np_dtype)
    y_ref = getattr(ro, "vector_product_user_addr_" + flat.dtype)(H.address, x, True)
    assert rel_err(fo.matvec_user(flat, x), y_ref) < TOL
    mu = 4
    X = np.asfortranarray(helpers.rand_vec(rng, flat.n_cols, flat.np_dtype, mu))
    Y_ref = getattr(ro, "matrix_product_user_addr_" + flat.dtype)(H.address, X, True)
    Y = np.stack([fo.matvec_user(flat, np.ascontiguousarray(X[:, j])) for j in range(mu)], axis=1)
    assert rel_err(Y, Y_ref) < TOL
    # row-major internal entry point
    Xb = np.ascontiguousarray(helpers.rand_vec(rng, flat.n_cols, flat.np_dtype, mu))
    Yb_ref = np.zeros((flat.n_rows, mu), dtype=flat.np_dtype)
    getattr(ro, "internal_matrix_product_row_major_addr_" + flat.dtype)(H.address, "N", 1.0, Xb, 0.0, Yb_ref, mu, True)
    Yb = np.zeros((flat.n_rows, mu), dtype=flat.np_dtype)
    fo.add_matrix_product_row_major(flat, "N", 1.0, Xb, 0.0, Yb, mu)
    assert rel_err(Yb, Yb_ref) < TOL


def test_reference_python_object_and_native_flatten_identically():
    """flat_view of the reference's own Python HMatrix (cross-module cast) reproduces Htool.HMatrix.__mul__."""
    Htool, _ = mods
    from htool_python_b200 import FlatHMatrix

    rng = np.random.default_rng(3)
    n = 400
    pts = np.asfortranarray(rng.random((3, n)))

    class Gen(Htool.VirtualGenerator):
        def build_submatrix(self, J, K, mat):
            mat[:] = 1.0 / (0.1 + np.linalg.norm(pts[:, J][:, :, None] - pts[:, K][:, None, :], axis=0))

    cb = Htool.ClusterTreeBuilder()
    cb.set_maximal_leaf_size(10)
    c = cb.create_cluster_tree(pts, 2)
    H = Htool.HMatrixTreeBuilder(1e-3, 10.0, "S", "L").build(Gen(), c, c)
    flat = FlatHMatrix.from_htool(H)
    x = rng.random(n)
    assert rel_err(fo.matvec_user(flat, x), H * x) < TOL
    info = H.get_local_information()
    assert abs(n * n / flat.stored_coefficients - float(info["Compression_ratio"])) < 1e-4 * float(info["Compression_ratio"])


def test_row_slab_operators_reassemble_the_full_product():
    """block-row split of DistributedOperator (distributed_operator/utility.hpp:56): concat of slab products = full."""
    _, ro = mods
    P = 4
    H, flat, rng = _build("laplace_reg", "N", "N", 2000, 1e-4, 16, seed=5, partition=P)
    x = helpers.rand_vec(rng, flat.n_cols, flat.np_dtype)
    y_full = np.zeros(flat.n_rows)
    fo.add_vector_product(flat, "N", 1.0, x, 0.0, y_full)
    parts = []
    for p in range(P):
        Hp, fp, _ = _build("laplace_reg", "N", "N", 2000, 1e-4, 16, seed=5, partition=P, slab=p)
        yp = np.zeros(fp.n_rows)
        fo.add_vector_product(fp, "N", 1.0, x, 0.0, yp)
        yp_ref = np.zeros(fp.n_rows)
        ro.internal_vector_product_addr_d(Hp.address, "N", 1.0, x, 0.0, yp_ref, True)
        assert rel_err(yp, yp_ref) < TOL
        parts.append((fp.target_offset, yp))
    parts.sort(key=lambda t: t[0])
    assert rel_err(np.concatenate([p[1] for p in parts]), y_full) < 1e-3  # slabs are compressed independently (eps)
