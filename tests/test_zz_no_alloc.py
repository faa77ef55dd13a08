"""hmb200_krylov_dots_static (-m gpu): the dot products for callers that cannot provide a device workspace.

Some hosts of this library have no way to allocate device memory when they need h = V^H w: a callback that a foreign solver
invokes with bare pointers, code that must not stall the stream it was given (cudaMalloc / cudaFree synchronise the device), a
process whose HBM is planned to the last byte, a plain C or ctypes binding without an allocator.  For them the library keeps
the partial sums in a 4 MiB array that is part of its own device code.  This file checks, through the C ABI, that the variant
gives what hmb200_krylov_dots gives with a caller-allocated workspace -- twice in a row with different shapes, because all
calls share that one array and the second call finds the leftovers of the first -- and that a request that does not fit is
refused instead of overrunning it.

The file name sorts last on purpose: the driver runs pytest -x, so a failure here must not hide the parity tests.  Nothing
here depends on other tests having run (or not): the first call in the process does the per-device address look-up itself.
"""
import numpy as np
import pytest

import helpers
from helpers import rel_err

pytestmark = pytest.mark.gpu

TOL = 1e-13  # the bound test_gpu_solver.py::test_krylov_kernels_against_numpy puts on the same kernels


def _both_ways(lib, _lib, torch, rng, n, k, ld_pad, cplx):
    """h from the existing entry point (workspace allocated here, by torch) and from the static-workspace one."""
    dev = torch.device("cuda", 0)
    dt = np.complex128 if cplx else np.float64
    es = 16 if cplx else 8
    ld = n + ld_pad
    V = helpers.rand_vec(rng, k * ld, dt).reshape(k, ld)
    w = helpers.rand_vec(rng, n, dt)
    Vd, wd = torch.from_numpy(V).to(dev), torch.from_numpy(w).to(dev)
    s = torch.cuda.current_stream().cuda_stream
    scratch = torch.empty(int(lib.hmb200_krylov_scratch_elems(n, k)), dtype=Vd.dtype, device=dev)
    h_old = torch.empty(k, dtype=Vd.dtype, device=dev)
    _lib.check(lib.hmb200_krylov_dots(Vd.data_ptr(), ld, k, wd.data_ptr(), n, h_old.data_ptr(), scratch.data_ptr(), es, s))
    h_new = torch.full((k,), float("nan"), dtype=Vd.dtype, device=dev)  # every entry must be overwritten
    _lib.check(lib.hmb200_krylov_dots_static(Vd.data_ptr(), ld, k, wd.data_ptr(), n, h_new.data_ptr(), es, s))
    torch.cuda.synchronize()
    return h_old.cpu().numpy(), h_new.cpu().numpy(), V[:, :n].conj() @ w


@pytest.mark.parametrize("cplx", [False, True])
def test_static_workspace_dots_match_allocated_workspace_dots(cplx):
    import torch

    from htool_python_b200 import _lib

    lib = _lib.lib()
    rng = np.random.default_rng(7 + cplx)
    with torch.cuda.device(0):  # the address look-up of the first call is for the CURRENT device
        # two calls in a row on the shared array: a large one, then a smaller one with another k, so the layout of the
        # partial sums (chunk * k + j) changes and stale entries of the first call lie where the second one writes and beyond
        for n, k, ld_pad in [(100003, 41, 1), (5000, 7, 3)]:
            h_old, h_new, h_ref = _both_ways(lib, _lib, torch, rng, n, k, ld_pad, cplx)
            assert np.all(np.isfinite(h_new.view(np.float64)))
            assert rel_err(h_new, h_old) < TOL, (n, k)
            assert rel_err(h_new, h_ref) < TOL, (n, k)
        # empty vectors give zeros, like the existing entry point
        h_old, h_new, _ = _both_ways(lib, _lib, torch, rng, 0, 3, 0, cplx)
        assert not h_new.any() and not h_old.any()


def test_static_workspace_refuses_what_does_not_fit():
    import torch

    from htool_python_b200 import _lib

    lib = _lib.lib()
    limit = 4 * 1024 * 1024  # HMB200_STATIC_SCRATCH_BYTES
    with torch.cuda.device(0):
        dev = torch.device("cuda", 0)
        n = 2048 * 1024  # 1024 chunks
        k_fit = limit // (1024 * 8)  # 512 real vectors fill the array exactly ...
        w = torch.ones(n, dtype=torch.float64, device=dev)
        h = torch.zeros(k_fit + 1, dtype=torch.float64, device=dev)
        s = torch.cuda.current_stream().cuda_stream
        # ... one more does not: refused before anything is launched (V is never dereferenced, ldv = 0 reuses w as every vector)
        rc = lib.hmb200_krylov_dots_static(w.data_ptr(), 0, k_fit + 1, w.data_ptr(), n, h.data_ptr(), 8, s)
        assert rc != 0
        msg = _lib.last_error()
        assert "built-in workspace" in msg and str(limit) in msg, msg
        torch.cuda.synchronize()
        assert not h.cpu().numpy().any()
        # the exact fit is accepted: <w, w> = n for each of the k_fit (identical) vectors
        _lib.check(lib.hmb200_krylov_dots_static(w.data_ptr(), 0, k_fit, w.data_ptr(), n, h.data_ptr(), 8, s))
        torch.cuda.synchronize()
        got = h.cpu().numpy()
        assert np.array_equal(got[:k_fit], np.full(k_fit, float(n))) and got[k_fit] == 0.0
        with pytest.raises(RuntimeError, match="elem_bytes"):
            _lib.check(lib.hmb200_krylov_dots_static(w.data_ptr(), 0, 1, w.data_ptr(), n, h.data_ptr(), 4, s))
