"""The Python plugin form of the boundary (-m gpu): `htool_python_b200.plugin` subclasses of the REFERENCE's own plugin classes,
driven through the reference's real trampolines (oracle/_ref/htool_plugin_host = src/htool/local_operator/*.hpp compiled without MPI):
the C++ side calls the virtual operator interface with raw pointers, the trampoline wraps them as zero-copy numpy views, and the
GPU engine must write the caller's buffer in place.  Results against the oracle at 1e-12."""
import glob
import os

import numpy as np
import pytest

import helpers
from helpers import TOL, load_golden, rel_err
from oracle import flat_oracle as fo
from oracle import helpers_ref

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not glob.glob(os.path.join(helpers_ref.REF_DIR, "htool_plugin_host*.so")), reason="oracle/_ref/htool_plugin_host not built")]


def _alpha_beta(cplx):
    return [(1.0, 0.0), (1.25, -0.5)] + ([(0.5 - 0.25j, 0.75 + 0.5j)] if cplx else [])


@pytest.mark.skipif(not helpers.bridge_available(), reason="bridge not built")
@pytest.mark.parametrize("kernel", ["laplace_reg", "complex_reg"])
def test_global_to_local_plugin_under_reference_trampoline(kernel):
    """two block-row slabs assembled by the reference; each one as a B200GlobalToLocalOperator: global input, local output"""
    from htool_python_b200 import FlatHMatrix, bridge
    from htool_python_b200.plugin import global_to_local_operator_class

    hp = helpers_ref.plugin_host()
    cplx = kernel == "complex_reg"
    sfx = "_g2l_z" if cplx else "_g2l_d"
    rng = np.random.default_rng(3)
    n, mu = 3000, 5
    g = rng.standard_normal((3, n))
    tree = bridge.create_cluster_tree(np.asfortranarray(g / np.linalg.norm(g, axis=0)), 2, 2, 32)
    Op = global_to_local_operator_class(hp, complex=cplx)
    ident = np.arange(n, dtype=np.int32)  # LocalRenumbering keeps the pointer: alive for the whole test
    for p, (off, size) in enumerate(tree.partition):
        H = bridge.build_hmatrix(kernel, tree, epsilon=1e-4, eta=10.0, target_partition_number=p, partition_number_for_symmetry=p)
        flat = FlatHMatrix.from_htool(H)
        dt = flat.np_dtype
        op = Op(H, hp.LocalRenumbering(int(off), int(size), ident), hp.LocalRenumbering(0, n, ident))
        assert op.local_target_renumbering.offset == off and op.local_source_renumbering.size == n
        for alpha, beta in _alpha_beta(cplx):
            x, y0 = helpers.rand_vec(rng, n, dt), helpers.rand_vec(rng, size, dt)
            y = y0.copy()
            getattr(hp, "drive_add_vector_product" + sfx)(op, "N", alpha, x, beta, y)  # C++ -> trampoline -> Python -> GPU, in place
            ref = y0.copy()
            fo.add_vector_product(flat, "N", alpha, x, beta, ref)
            assert rel_err(y, ref) < TOL, (p, alpha, beta)
            X, Y0 = helpers.rand_vec(rng, n, dt, mu), helpers.rand_vec(rng, size, dt, mu)
            Y = Y0.copy()
            getattr(hp, "drive_add_matrix_product_row_major" + sfx)(op, "N", alpha, X, beta, Y, mu)
            Ref = Y0.copy()
            fo.add_matrix_product_row_major(flat, "N", alpha, X, beta, Ref, mu)
            assert rel_err(Y, Ref) < TOL, (p, alpha, beta)
        # out += A[:, offset:offset+size] in through the reference's own clipping (restricted_operator.hpp:173-194)
        w0, ws = n // 3, n // 4
        Xs, Ys = helpers.rand_vec(rng, ws, dt, mu), np.zeros((size, mu), dt)
        getattr(hp, "drive_add_sub_matrix_product_to_local" + sfx)(op, Xs, Ys, mu, w0, ws)
        Xz = np.zeros((n, mu), dt)
        Xz[w0:w0 + ws] = Xs
        Ref = np.zeros((size, mu), dt)
        fo.add_matrix_product_row_major(flat, "N", 1.0, Xz, 0.0, Ref, mu)
        assert rel_err(Ys, Ref) < TOL, p
        # the in-place contract is enforced: a wrong output is refused instead of being silently lost
        with pytest.raises(RuntimeError, match="in place"):
            op.add_vector_product("N", 1.0, np.zeros(n, dt), 0.0, np.zeros(size + 1, dt))
        op.close()


@pytest.mark.parametrize("name", ["d_N_cylinder", "z_N_sphere"])
def test_local_to_local_plugin_under_reference_trampoline(name):
    """a square operator as B200LocalToLocalOperator: 'N' and the transposed products through local_add_*"""
    from htool_python_b200.plugin import local_to_local_operator_class

    hp = helpers_ref.plugin_host()
    flat, _ = load_golden(name)
    cplx = flat.dtype == "z"
    sfx = "_l2l_z" if cplx else "_l2l_d"
    n, mu, dt = flat.n_rows, 4, flat.np_dtype
    ident = np.arange(n, dtype=np.int32)
    op = local_to_local_operator_class(hp, complex=cplx)(flat, hp.LocalRenumbering(0, n, ident), hp.LocalRenumbering(0, n, ident), build_transpose=1)
    rng = np.random.default_rng(5)
    for tr in ["N", "T"] + (["C"] if cplx else []):
        for alpha, beta in _alpha_beta(cplx):
            x, y0 = helpers.rand_vec(rng, n, dt), helpers.rand_vec(rng, n, dt)
            y = y0.copy()
            getattr(hp, "drive_add_vector_product" + sfx)(op, tr, alpha, x, beta, y)
            ref = y0.copy()
            fo.add_vector_product(flat, tr, alpha, x, beta, ref)
            assert rel_err(y, ref) < TOL, (tr, alpha, beta)
    X, Y0 = helpers.rand_vec(rng, n, dt, mu), helpers.rand_vec(rng, n, dt, mu)
    Y = Y0.copy()
    getattr(hp, "drive_add_matrix_product_row_major" + sfx)(op, "N", 1.25, X, -0.5, Y, mu)
    Ref = Y0.copy()
    fo.add_matrix_product_row_major(flat, "N", 1.25, X, -0.5, Ref, mu)
    assert rel_err(Y, Ref) < TOL
    op.close()
