"""oracle/helpers_ref.py -- TEST INFRASTRUCTURE ONLY: loader of oracle/_ref/htool_ref_ops (the reference's own CPU products
compiled from /root/reference by oracle/Makefile).  Used by tests/ and by bench.py's cpu_baseline / --impl reference legs."""
import glob
import os
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(_HERE, "_ref")


def ref_ops():
    if not glob.glob(os.path.join(REF_DIR, "htool_ref_ops*.so")):
        raise RuntimeError("oracle/_ref/htool_ref_ops*.so is missing (make -C oracle ref, needs /root/reference)")
    root = os.path.dirname(_HERE)
    if root not in sys.path:
        sys.path.insert(0, root)
    from htool_python_b200 import _blas

    _blas.preload()
    if REF_DIR not in sys.path:
        sys.path.insert(0, REF_DIR)
    import htool_ref_ops

    return htool_ref_ops
