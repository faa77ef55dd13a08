"""The Python plugin form of the drop-in boundary: subclasses of the reference's own plugin classes whose products
run on the GPU engine.

The reference lets Python code provide a rank's operator by subclassing `Htool.RestrictedGlobalToLocalOperator(target_renumbering,
source_renumbering, target_use_permutation, source_use_permutation)` and overriding `add_vector_product(trans, alpha, in, beta,
out)` / `add_matrix_product_row_major(trans, alpha, in, beta, out)` (src/htool/local_operator/local_operator.hpp:41-70,73-87), or
`Htool.VirtualLocalToLocalOperator(target_renumbering, source_renumbering)` with `local_add_vector_product` /
`local_add_matrix_product_row_major` (src/htool/local_operator/virtual_local_to_local_operator.hpp:58-84,89-99); worked example
example/advanced/define_custom_local_operator.py:6-109.  The trampolines hand over zero-copy numpy views of the C++ buffers
(local_operator.hpp:14-28): `in` has the size of the local source renumbering, `out` of the local target renumbering, blocks are
row-major (rows, mu), and the result must be written INTO `out` (rebinding the name is silently lost).

The reference module is a run-time choice of the user (and only registers these classes when built with MPI), so the classes are
made from the module that is passed in:

    import Htool
    from htool_python_b200.plugin import global_to_local_operator_class
    B200Op = global_to_local_operator_class(Htool)                   # complex=True for ComplexRestrictedGlobalToLocalOperator
    op = B200Op(builder.hmatrix, Htool.LocalRenumbering(target_cluster_on_partition), Htool.LocalRenumbering(source_root))
    dist_op = Htool.CustomApproximationBuilder(target_cluster, source_cluster, comm, op).distributed_operator   # keep `op` alive

There is no CPU fallback: creating an operator without a CUDA device raises.
"""
from __future__ import annotations

import numpy as np

from .engine import Engine
from .flat import FlatHMatrix


def _make_engine(hmatrix, engine_options):
    flat = hmatrix if isinstance(hmatrix, FlatHMatrix) else FlatHMatrix.from_htool(hmatrix)
    return flat, Engine(flat, **engine_options)


def _check_in_place(engine, out, rows, mu=None):
    """the trampoline's `out` is a view of the caller's C++ buffer: it must be used as it is (no copy, no cast)"""
    want = (rows,) if mu is None else (rows, mu)
    if not isinstance(out, np.ndarray) or out.dtype != engine.np_dtype or not out.flags.c_contiguous or not out.flags.writeable or out.shape != want:
        raise RuntimeError(f"output must be a writeable C-contiguous {np.dtype(engine.np_dtype).name} array of shape {want} (it is written in place)")


def global_to_local_operator_class(htool_module, complex: bool = False):
    """`B200GlobalToLocalOperator`, a subclass of `htool_module.[Complex]RestrictedGlobalToLocalOperator` backed by the GPU engine.

    B200GlobalToLocalOperator(hmatrix, target_renumbering, source_renumbering, **engine_options): `hmatrix` is the rank's slab as
    assembled by the reference (`Htool.HMatrix`, e.g. `DefaultApproximationBuilder.hmatrix`) or a `FlatHMatrix`; it is flattened,
    packed and uploaded once.  `build_transpose=1` among the options makes trans = 'T' / 'C' available."""
    base = getattr(htool_module, ("Complex" if complex else "") + "RestrictedGlobalToLocalOperator")

    class B200GlobalToLocalOperator(base):
        def __init__(self, hmatrix, target_renumbering, source_renumbering, **engine_options):
            # no extra permutation on either side: the engine works in the cluster numbering of the slab (restricted_operator.hpp:22-23)
            base.__init__(self, target_renumbering, source_renumbering, False, False)
            self.flat, self.engine = _make_engine(hmatrix, engine_options)
            if self.engine.is_complex != bool(complex):
                raise RuntimeError("the H-matrix and the operator class differ in their coefficient type")

        def add_vector_product(self, trans, alpha, input, beta, output):  # local_operator.hpp:46-57
            _check_in_place(self.engine, output, self.engine._sizes(trans)[1])
            self.engine.add_vector_product(trans, alpha, np.ascontiguousarray(input, self.engine.np_dtype), beta, output)

        def add_matrix_product_row_major(self, trans, alpha, input, beta, output):  # local_operator.hpp:58-69
            mu = int(output.shape[1]) if getattr(output, "ndim", 0) == 2 else 0
            _check_in_place(self.engine, output, self.engine._sizes(trans)[1], mu)
            self.engine.add_matrix_product_row_major(trans, alpha, np.ascontiguousarray(input, self.engine.np_dtype), beta, output, mu=mu)

        def close(self):
            self.engine.close()

    return B200GlobalToLocalOperator


def local_to_local_operator_class(htool_module, complex: bool = False):
    """`B200LocalToLocalOperator`, a subclass of `htool_module.[Complex]VirtualLocalToLocalOperator` (the block-diagonal operator of a
    rank: local input, local output) backed by the GPU engine; same constructor as `B200GlobalToLocalOperator`."""
    base = getattr(htool_module, ("Complex" if complex else "") + "VirtualLocalToLocalOperator")

    class B200LocalToLocalOperator(base):
        def __init__(self, hmatrix, target_renumbering, source_renumbering, **engine_options):
            base.__init__(self, target_renumbering, source_renumbering)
            self.flat, self.engine = _make_engine(hmatrix, engine_options)
            if self.engine.is_complex != bool(complex):
                raise RuntimeError("the H-matrix and the operator class differ in their coefficient type")

        def local_add_vector_product(self, trans, alpha, input, beta, output):  # virtual_local_to_local_operator.hpp:58-70
            _check_in_place(self.engine, output, self.engine._sizes(trans)[1])
            self.engine.add_vector_product(trans, alpha, np.ascontiguousarray(input, self.engine.np_dtype), beta, output)

        def local_add_matrix_product_row_major(self, trans, alpha, input, beta, output):  # :71-84
            mu = int(output.shape[1]) if getattr(output, "ndim", 0) == 2 else 0
            _check_in_place(self.engine, output, self.engine._sizes(trans)[1], mu)
            self.engine.add_matrix_product_row_major(trans, alpha, np.ascontiguousarray(input, self.engine.np_dtype), beta, output, mu=mu)

        def close(self):
            self.engine.close()

    return B200LocalToLocalOperator
