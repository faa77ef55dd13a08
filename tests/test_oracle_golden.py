"""The oracle (oracle/flat_oracle.c) against the golden vectors produced by the reference itself (tests/golden/)."""
import numpy as np
import pytest

import helpers
from helpers import TOL, load_golden, rel_err
from oracle import flat_oracle as fo


@pytest.mark.parametrize("name", helpers.golden_names())
def test_oracle_matvec_matches_reference_golden(name):
    flat, z = load_golden(name)
    y = fo.matvec_user(flat, z["x"])
    assert rel_err(y, z["y_ref"]) < TOL            # reference's own H * x
    assert rel_err(y, z["y_dense"]) < 1e-10        # reference test's densification check (tests/test_hmatrix.py:84)


@pytest.mark.parametrize("name", [n for n in helpers.golden_names() if "rect" not in n])
def test_oracle_matmat_matches_reference_golden(name):
    flat, z = load_golden(name)
    X = z["X"]
    mu = X.shape[1]
    # user numbering -> cluster numbering per column, row-major product, back (add_hmatrix_matrix_product.hpp:170-205)
    Xb = np.stack([fo.permute_gather(flat.source_permutation, flat.source_offset, flat.n_cols, np.ascontiguousarray(X[:, j])) for j in range(mu)], axis=1)
    Yb = np.zeros((flat.n_rows, mu), dtype=flat.np_dtype)
    fo.add_matrix_product_row_major(flat, "N", 1.0, np.ascontiguousarray(Xb), 0.0, Yb, mu)
    Y = np.stack([fo.permute_scatter(flat.target_permutation, flat.target_offset, flat.n_rows, np.ascontiguousarray(Yb[:, j])) for j in range(mu)], axis=1)
    assert rel_err(Y, z["Y_ref"]) < TOL


@pytest.mark.parametrize("name", helpers.golden_names())
def test_stored_coefficients_match_reference_compression_ratio(name):
    flat, z = load_golden(name)
    # Compression_ratio = rows*cols / stored coefficients (hmatrix_output.hpp:153-208) -- defines the algorithmic bytes
    assert abs(flat.n_rows * flat.n_cols / flat.stored_coefficients - float(z["compression_ratio"])) < 1e-4 * float(z["compression_ratio"])
    assert int((flat.rank < 0).sum()) == int(z["n_dense"]) and int((flat.rank >= 0).sum()) == int(z["n_low_rank"])


def test_permutations_bit_exact_roundtrip():
    rng = np.random.default_rng(0)
    n, off = 1000, 37
    perm = np.concatenate([np.arange(off), off + rng.permutation(n)]).astype(np.int32)
    for dt in (np.float64, np.complex128):
        x = helpers.rand_vec(rng, n, dt)
        g = fo.permute_gather(perm, off, n, x)
        assert np.array_equal(g, x[perm[off:] - off])
        s = fo.permute_scatter(perm, off, n, g)
        assert np.array_equal(s.view(np.uint8), x.view(np.uint8))  # bit-exact round trip
