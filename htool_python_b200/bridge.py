"""Lazy access to the pybind11 flattening shim (`_htool_bridge`, built from csrc/bridge/htool_bridge.cpp)."""
from __future__ import annotations

import importlib

from . import _blas

_mod = None


def module():
    global _mod
    if _mod is None:
        _blas.preload()  # the reference's header-only core links BLAS/LAPACK
        try:
            _mod = importlib.import_module("htool_python_b200._htool_bridge")
        except ImportError as e:  # fail loudly: there is no pure-Python flattening path
            raise ImportError(
                "htool_python_b200._htool_bridge is not built; run `python -m htool_python_b200.build bridge` "
                "(needs the reference headers under /root/reference)") from e
    return _mod


KERNEL_NAMES = ("laplace", "laplace_reg", "invdist_reg", "helmholtz", "complex_reg", "complex_herm")
COMPLEX_KERNELS = ("helmholtz", "complex_reg", "complex_herm")


def create_cluster_tree(points, number_of_children=2, size_of_partition=1, maximal_leaf_size=64):
    """Reference ClusterTreeBuilder::create_cluster_tree (clustering/tree_builder/tree_builder.hpp:52) on (dim, N) points."""
    return module().create_cluster_tree(points, number_of_children, size_of_partition, maximal_leaf_size)


def build_hmatrix(kernel, target_tree, source_tree=None, epsilon=1e-3, eta=10.0, symmetry="N", UPLO="N",
                  target_partition_number=-1, partition_number_for_symmetry=-1, wavenumber=0.0, parallel=True,
                  minimal_target_depth=0, minimal_source_depth=0):
    """Assemble with the reference's HMatrixTreeBuilder::build and a native C++ generator for `kernel`."""
    br = module()
    source_tree = target_tree if source_tree is None else source_tree
    kid = br.KERNELS[kernel]
    fn = br.build_complex_hmatrix if kernel in COMPLEX_KERNELS else br.build_hmatrix
    return fn(kid, float(wavenumber), target_tree, source_tree, float(epsilon), float(eta), symmetry, UPLO,
              int(target_partition_number), int(partition_number_for_symmetry), bool(parallel),
              int(minimal_target_depth), int(minimal_source_depth))
