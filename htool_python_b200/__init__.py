"""htool_python_b200 -- B200-native engine for the H-matrix product path of htool-ddm/htool_python.

Scope: `HMatrix.__mul__/__matmul__`, the operators behind `DistributedOperator`'s matvec/matmat and the
unpreconditioned GMRES of the DDM solver on top of them (`B200Solver`, Krylov basis resident on the device).
The H-matrix is assembled by the reference; `FlatHMatrix.from_htool` flattens it, `Engine`/`B200HMatrix` serve the
products from hand-written sm_100a kernels (libhmb200.so, C ABI in include/hmb200.h).
"""
from .flat import FlatHMatrix  # noqa: F401

__all__ = ["FlatHMatrix", "Engine", "B200HMatrix", "B200DistributedOperator", "B200Solver", "bridge"]


def __getattr__(name):
    if name == "Engine":
        from .engine import Engine
        return Engine
    if name == "B200HMatrix":
        from .hmatrix import B200HMatrix
        return B200HMatrix
    if name == "B200DistributedOperator":
        from .distributed import B200DistributedOperator
        return B200DistributedOperator
    if name == "B200Solver":
        from .solver import B200Solver
        return B200Solver
    if name == "bridge":
        import importlib
        return importlib.import_module(".bridge", __name__)
    raise AttributeError(name)
