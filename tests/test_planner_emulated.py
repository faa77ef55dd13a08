"""Host logic of the product (planner + packer, csrc/capi/planner.cpp) validated on the CPU: the packed packet streams
are interpreted by tests/native/plan_emulator.cpp and compared with the oracle.  Covers every layout decision the
CUDA kernels rely on (alignment, operand copies, partial slots, one-writer-per-row) without a GPU."""
import numpy as np
import pytest

import helpers
from helpers import TOL, emulate, load_golden, rel_err, synthetic_flat
from oracle import flat_oracle as fo

OPTION_SETS = [
    {},
    dict(stage_bytes=4096, stage_z_bytes=1024, piece1_bytes=512, piece2_bytes=700, bin_rows=20, unit_stages=3),
    dict(stage_bytes=2048, stage_z_bytes=1024, piece1_bytes=256, piece2_bytes=256, bin_rows=7, unit_stages=1),
    dict(stage_bytes=65536, stage_z_bytes=16384, piece1_bytes=32768, piece2_bytes=32768, bin_rows=64, unit_stages=1000),
]


def _check(flat, op, opts, rng, alpha=None, beta=None):
    n_in = flat.n_cols if op == "N" else flat.n_rows
    n_out = flat.n_rows if op == "N" else flat.n_cols
    if alpha is None:
        alpha, beta = (1.25, -0.75) if flat.dtype == "d" else (1.25 - 0.5j, 0.3 + 0.9j)
    x = helpers.rand_vec(rng, n_in, flat.np_dtype)
    y0 = helpers.rand_vec(rng, n_out, flat.np_dtype)
    y_ref = y0.copy()
    fo.add_vector_product(flat, op, alpha, x, beta, y_ref)
    y = y0.copy()
    stats = emulate(flat, op, alpha, x, beta, y, **opts)
    assert rel_err(y, y_ref) < TOL
    return stats


@pytest.mark.parametrize("name", helpers.golden_names())
@pytest.mark.parametrize("oi", range(len(OPTION_SETS)))
def test_packed_streams_reproduce_golden(name, oi):
    flat, z = load_golden(name)
    rng = np.random.default_rng(oi)
    sym = flat.symmetry_for_leaves
    ops = ["N"] + (["T"] if sym != "H" else []) + (["C"] if flat.dtype == "z" and sym != "S" else [])
    for op in ops:
        _check(flat, op, OPTION_SETS[oi], rng)
    # and the reference's own vector, through the permutations
    xb = fo.permute_gather(flat.source_permutation, flat.source_offset, flat.n_cols, z["x"])
    yb = np.zeros(flat.n_rows, dtype=flat.np_dtype)
    emulate(flat, "N", 1.0, xb, 0.0, yb, **OPTION_SETS[oi])
    y = fo.permute_scatter(flat.target_permutation, flat.target_offset, flat.n_rows, yb)
    assert rel_err(y, z["y_ref"]) < TOL


@pytest.mark.parametrize("seed", range(12))
def test_synthetic_tilings(seed):
    rng = np.random.default_rng(100 + seed)
    dtype = "dz"[seed % 2]
    m, n = int(rng.integers(1, 400)), int(rng.integers(1, 400))
    flat = synthetic_flat(rng, m, n, dtype=dtype, max_leaf=int(rng.integers(4, 150)), max_rank=int(rng.integers(1, 70)))
    A = helpers.flat_to_dense(flat)
    x = helpers.rand_vec(rng, n, flat.np_dtype)
    y = np.zeros(m, dtype=flat.np_dtype)
    fo.add_vector_product(flat, "N", 1.0, x, 0.0, y)
    assert rel_err(y, A @ x) < 1e-11  # the oracle itself against plain dense algebra
    for opts in OPTION_SETS:
        _check(flat, "N", opts, rng)
        _check(flat, "T", opts, rng)
        if dtype == "z":
            _check(flat, "C", opts, rng)


def test_edge_cases():
    rng = np.random.default_rng(7)
    # beta = 0 must overwrite NaNs (BLAS semantics); alpha = 0
    flat = synthetic_flat(rng, 130, 90, max_leaf=40)
    x = helpers.rand_vec(rng, 90, np.float64)
    y = np.full(130, np.nan)
    emulate(flat, "N", 1.0, x, 0.0, y)
    y_ref = np.zeros(130)
    fo.add_vector_product(flat, "N", 1.0, x, 0.0, y_ref)
    assert rel_err(y, y_ref) < TOL
    y0 = helpers.rand_vec(rng, 130, np.float64)
    y = y0.copy()
    emulate(flat, "N", 0.0, x, 2.0, y)
    assert np.allclose(y, 2.0 * y0, rtol=0, atol=1e-15)
    # a single dense leaf larger than a bin, a single low-rank leaf with rank > 32, 1x1
    for m, n, r in [(200, 3, -1), (3, 200, -1), (150, 170, 45), (1, 1, -1), (1, 1, 1), (70, 70, 0)]:
        f = synthetic_flat(rng, m, n, max_leaf=1000, p_lowrank=0.0, p_stop=1.0)
        assert f.n_leaves == 1
        f.rank[:] = r
        if r >= 0:
            coef = helpers.rand_vec(rng, r * (m + n) + 1, np.float64)
            f = f.with_coefficients(coef, np.array([0]), np.array([m * r]))
        for opts in OPTION_SETS:
            _check(f, "N", opts, rng)
            _check(f, "T", opts, rng)
    # rows that no leaf covers still receive beta*y (slab operators may be sparse in rows)
    f = synthetic_flat(rng, 100, 50, max_leaf=20)
    keep = f.row_off >= 40
    f2 = helpers.FlatHMatrix(**{**f.__dict__, "row_off": f.row_off[keep], "col_off": f.col_off[keep], "nrows": f.nrows[keep], "ncols": f.ncols[keep],
                               "rank": f.rank[keep], "flags": f.flags[keep], "a_ptr": f.a_ptr[keep], "b_ptr": f.b_ptr[keep]})
    _check(f2, "N", {}, rng)
    # empty operator
    f0 = helpers.FlatHMatrix(**{**f.__dict__, "row_off": f.row_off[:0], "col_off": f.col_off[:0], "nrows": f.nrows[:0], "ncols": f.ncols[:0],
                               "rank": f.rank[:0], "flags": f.flags[:0], "a_ptr": f.a_ptr[:0], "b_ptr": f.b_ptr[:0]})
    _check(f0, "N", {}, rng)


def test_invalid_inputs_are_rejected():
    rng = np.random.default_rng(1)
    f = synthetic_flat(rng, 50, 50)
    f.row_off[0] = 49
    f.nrows[0] = 5  # leaf sticks out of the root block
    with pytest.raises(RuntimeError, match="outside"):
        emulate(f, "N", 1.0, np.zeros(50), 0.0, np.zeros(50))
    flat, _ = load_golden("z_HU_sphere")
    with pytest.raises(RuntimeError, match="not supported"):  # reference logs the same error (add_hmatrix_vector_product.hpp:58-61)
        emulate(flat, "T", 1.0, np.zeros(flat.n_rows, complex), 0.0, np.zeros(flat.n_cols, complex))


def test_plan_statistics_and_overhead():
    """packed bytes stay close to the algorithmic bytes (no mirrored leaf: < 6 % overhead at default options)."""
    flat, _ = load_golden("d_N_cylinder")
    y = np.zeros(flat.n_rows)
    st = emulate(flat, "N", 1.0, np.ones(flat.n_cols), 0.0, y)
    packed = st["stream1_bytes"] + st["stream2_bytes"]
    assert st["stored"] == int(flat.stored_coefficients)
    assert packed >= 8 * flat.stored_coefficients
    assert packed < 1.12 * 8 * flat.stored_coefficients  # tiny leaves (leaf size 16) -> descriptor overhead is at its worst here


MM_OPTION_SETS = [
    {},
    dict(mm_task_cols=40, mm1_kg=2, mm2_kg=4, bin_rows=20),   # split phase-1 units (partial T rows + reduce), short packets
    dict(mm_task_cols=8, mm1_kg=2, mm2_kg=2, bin_rows=9),     # every leaf wider than 8 columns is split; two k-groups per packet
]


def _check_mm(flat, op, opts, rng, mu):
    n_in = flat.n_cols if op == "N" else flat.n_rows
    n_out = flat.n_rows if op == "N" else flat.n_cols
    alpha, beta = (1.25, -0.75) if flat.dtype == "d" else (1.25 - 0.5j, 0.3 + 0.9j)
    X = np.ascontiguousarray(helpers.rand_vec(rng, n_in, flat.np_dtype, mu))
    Y0 = np.ascontiguousarray(helpers.rand_vec(rng, n_out, flat.np_dtype, mu))
    Y_ref = Y0.copy()
    fo.add_matrix_product_row_major(flat, op, alpha, X, beta, Y_ref, mu)
    Y = Y0.copy()
    stats = helpers.emulate_mm(flat, op, alpha, X, beta, Y, **opts)
    assert rel_err(Y, Y_ref) < TOL
    return stats


@pytest.mark.parametrize("name", helpers.golden_names())
@pytest.mark.parametrize("oi", range(len(MM_OPTION_SETS)))
def test_multi_rhs_streams_reproduce_oracle(name, oi):
    """the multi-RHS plan (stacked V panels per source cluster with split units and partial rows, K-concatenated bin panels,
    fragment-major packing, operand rows, K padding on zero rows) on the golden H-matrices"""
    flat, _ = load_golden(name)
    rng = np.random.default_rng(40 + oi)
    sym = flat.symmetry_for_leaves
    ops = ["N"] + (["T"] if sym != "H" else []) + (["C"] if flat.dtype == "z" and sym != "S" else [])
    for op in ops:
        st = _check_mm(flat, op, MM_OPTION_SETS[oi], rng, mu=3)
    assert st["units_mm1"] >= 1 and st["mm1_bytes"] > 0


@pytest.mark.parametrize("seed", range(6))
def test_multi_rhs_synthetic_tilings(seed):
    rng = np.random.default_rng(300 + seed)
    dtype = "dz"[seed % 2]
    m, n = int(rng.integers(1, 300)), int(rng.integers(1, 300))
    flat = synthetic_flat(rng, m, n, dtype=dtype, max_leaf=int(rng.integers(4, 150)), max_rank=int(rng.integers(1, 150)))
    for opts in MM_OPTION_SETS:
        _check_mm(flat, "N", opts, rng, mu=int(rng.integers(1, 6)))


@pytest.mark.parametrize("name,mu", [("d_N_cylinder", 70), ("d_N_rect", 37), ("z_N_sphere", 33), ("z_HL_cylinder", 9)])
def test_multi_rhs_streams_column_tails_and_passes(name, mu):
    """several passes of 64 (32 complex) columns, a column tail in the last one, NaN pad columns in the emulated workspace"""
    flat, _ = load_golden(name)
    rng = np.random.default_rng(mu)
    for op in ["N"] + (["T"] if flat.symmetry_for_leaves != "H" else []):
        _check_mm(flat, op, MM_OPTION_SETS[1], rng, mu)
