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


def hpddm_solve(matvec, b, x0=None, args="", mu=1):
    """Runs the REFERENCE's Krylov solver (HPDDM::IterativeMethod::solve, the call of DDM::solve,
    lib/htool/include/htool/solvers/ddm.hpp:193) on one rank with `matvec(v) -> A v` as the operator (no preconditioner).
    b: (n,) or (n, mu) F-ordered like HPDDM; returns (x, iterations, operator_calls)."""
    import ctypes

    import numpy as np

    path = os.path.join(REF_DIR, "libhpddm_ref.so")
    if not os.path.exists(path):
        raise RuntimeError("oracle/_ref/libhpddm_ref.so is missing (make -C oracle ref, needs /root/reference)")
    root = os.path.dirname(_HERE)
    if root not in sys.path:
        sys.path.insert(0, root)
    from htool_python_b200 import _blas

    _blas.preload()
    lib = ctypes.CDLL(path)
    b = np.asarray(b)
    cplx = np.iscomplexobj(b)
    dt = np.complex128 if cplx else np.float64
    n = b.shape[0]
    bf = np.asfortranarray(b, dtype=dt)
    x = np.zeros_like(bf) if x0 is None else np.asfortranarray(x0, dtype=dt).copy(order="F")
    calls = [0]
    CB = ctypes.CFUNCTYPE(ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int)

    def cb(_ctx, pin, pout, m):
        w = 2 if cplx else 1
        vin = np.ctypeslib.as_array(ctypes.cast(pin, ctypes.POINTER(ctypes.c_double)), (n * m * w,)).view(dt)
        vout = np.ctypeslib.as_array(ctypes.cast(pout, ctypes.POINTER(ctypes.c_double)), (n * m * w,)).view(dt)
        for j in range(m):
            vout[j * n:(j + 1) * n] = matvec(vin[j * n:(j + 1) * n].copy())
        calls[0] += 1
        return 0

    # '-hpddm_schwarz_method' is registered by htool's DDM wrapper, not by HPDDM itself (it would only print a warning on stdout)
    toks = args.split()
    args = " ".join(t for i, t in enumerate(toks) if t != "-hpddm_schwarz_method" and not (i > 0 and toks[i - 1] == "-hpddm_schwarz_method"))
    fn = lib.hpddm_ref_solve_z if cplx else lib.hpddm_ref_solve_d
    fn.restype = ctypes.c_int
    it = fn(ctypes.c_int(n), CB(cb), None, ctypes.c_void_p(bf.ctypes.data), ctypes.c_void_p(x.ctypes.data), ctypes.c_int(mu), args.encode())
    return x, int(it), calls[0]


def plugin_host():
    """oracle/_ref/htool_plugin_host: the reference's own Python plugin classes (LocalRenumbering, RestrictedGlobalToLocalOperator,
    VirtualLocalToLocalOperator; src/htool/local_operator/*.hpp) in a module without MPI + drivers through the C++ interface"""
    if not glob.glob(os.path.join(REF_DIR, "htool_plugin_host*.so")):
        raise RuntimeError("oracle/_ref/htool_plugin_host*.so is missing (make -C oracle ref, needs /root/reference)")
    root = os.path.dirname(_HERE)
    if root not in sys.path:
        sys.path.insert(0, root)
    from htool_python_b200 import _blas

    _blas.preload()
    if REF_DIR not in sys.path:
        sys.path.insert(0, REF_DIR)
    import htool_plugin_host

    return htool_plugin_host
