"""The reference-side binding (include/hmb200_htool.hpp: RestrictedGlobalToLocalB200) compiled against the reference's headers and
run UNDER THE REFERENCE'S OWN DRIVERS on one rank (tests/native/reference_binding.cpp, built into oracle/_ref by oracle/Makefile):
the GPU operator is registered in a genuine htool::DistributedOperator through add_global_to_local_operator and compared with the
stock RestrictedGlobalToLocalHMatrix on the same htool::HMatrix -- global-to-global vector / row-major matrix products,
add_sub_matrix_product_to_local through the virtual interface, and DDM::solve -> HPDDM GMRES -> HPDDMOperator::GMV."""
import glob
import os
import sys

import numpy as np
import pytest

import helpers

REF_DIR = os.path.join(helpers.ROOT, "oracle", "_ref")
TOL = 1e-12
HPDDM_ARGS = "-hpddm_krylov_method gmres -hpddm_schwarz_method none -hpddm_variant right -hpddm_tol 1e-8 -hpddm_gmres_restart 30 -hpddm_max_it 120"


def binding():
    if not glob.glob(os.path.join(REF_DIR, "htool_b200_binding*.so")):
        pytest.skip("oracle/_ref/htool_b200_binding is not built (make -C oracle ref, needs /root/reference)")
    from htool_python_b200 import _blas, _lib

    _blas.preload()
    _lib.lib()  # libhmb200.so first: the binding links it
    if REF_DIR not in sys.path:
        sys.path.insert(0, REF_DIR)
    import htool_b200_binding

    return htool_b200_binding


def test_binding_module_is_built_against_the_reference():
    """not a GPU test: the adapter compiled against the reference's headers and links the C ABI"""
    mod = binding()
    assert hasattr(mod, "check_d") and hasattr(mod, "check_z")


def _points(n, seed=0):
    rng = np.random.default_rng(seed)
    theta = rng.random(n) * 2 * np.pi
    return np.asfortranarray(np.vstack([np.cos(theta), np.sin(theta), 3 * rng.random(n)]))


@pytest.mark.gpu
@pytest.mark.skipif(not helpers.bridge_available(), reason="bridge not built")
@pytest.mark.parametrize("kernel,sym,uplo", [("laplace_reg", "N", "N"), ("laplace_reg", "S", "L"), ("complex_reg", "N", "N"), ("complex_reg", "S", "L"), ("complex_herm", "H", "L")])
def test_gpu_operator_under_the_reference_drivers(kernel, sym, uplo):
    from htool_python_b200 import bridge

    mod = binding()
    n, mu = 3000, 5
    tree = bridge.create_cluster_tree(_points(n), 2, 1, 50)
    H = bridge.build_hmatrix(kernel, tree, epsilon=1e-6, eta=10.0, symmetry=sym, UPLO=uplo)
    fn = mod.check_z if H.dtype == "z" else mod.check_d
    res = fn(H.address, 0, mu, HPDDM_ARGS, True, True)
    print(kernel, sym, uplo, res)
    for key, val in res.items():
        if key.startswith(("vector_", "matrix_", "sub_")):
            assert val < TOL, (key, val)
    for sfx in ("_mu1", "_mu"):
        assert res["ddm_iterations_gpu" + sfx] == res["ddm_iterations_reference" + sfx], res
        assert res["ddm_iterations_gpu" + sfx] < 120
        # both iterates solve to tol 1e-8; GMRES amplifies summation-order differences by the condition number
        assert res["ddm_solution_rel_diff" + sfx] < 1e-6, res
        assert res["ddm_error_vs_exact_gpu" + sfx] < 1e-4, res
