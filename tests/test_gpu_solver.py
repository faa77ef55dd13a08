"""Device-resident GMRES (-m gpu): the Krylov vector kernels against numpy, and `B200Solver` against the committed
fixtures produced by the REFERENCE's HPDDM GMRES (tests/golden/gmres_golden.npz, make_gmres_golden.py): iteration counts
and operator calls must be EQUAL, solutions agree to a fraction of the GMRES tolerance."""
import ctypes

import numpy as np
import pytest

import helpers
from helpers import rel_err
from test_gmres_oracle import X_TOL, check_block_solution, check_solution, gmres_block_golden, gmres_golden

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("cplx", [False, True])
@pytest.mark.parametrize("n,k,ld_pad", [(1, 1, 0), (37, 3, 5), (2048, 8, 0), (2049, 9, 3), (100003, 41, 1), (5000, 200, 0)])
def test_krylov_kernels_against_numpy(cplx, n, k, ld_pad):
    import torch

    from htool_python_b200 import _lib

    lib = _lib.lib()
    rng = np.random.default_rng(n + k)
    dt = np.complex128 if cplx else np.float64
    es = 16 if cplx else 8
    ld = n + ld_pad
    V = helpers.rand_vec(rng, k * ld, dt).reshape(k, ld)
    w = helpers.rand_vec(rng, n, dt)
    dev = torch.device("cuda", 0)
    Vd, wd = torch.from_numpy(V).to(dev), torch.from_numpy(w).to(dev)
    scratch = torch.empty(int(lib.hmb200_krylov_scratch_elems(n, k)), dtype=Vd.dtype, device=dev)
    h = torch.empty(k, dtype=Vd.dtype, device=dev)
    s = torch.cuda.current_stream().cuda_stream
    _lib.check(lib.hmb200_krylov_dots(Vd.data_ptr(), ld, k, wd.data_ptr(), n, h.data_ptr(), scratch.data_ptr(), es, s))
    h_ref = V[:, :n].conj() @ w
    assert rel_err(h.cpu().numpy(), h_ref) < 1e-13
    # bit-reproducible
    h2 = torch.empty_like(h)
    _lib.check(lib.hmb200_krylov_dots(Vd.data_ptr(), ld, k, wd.data_ptr(), n, h2.data_ptr(), scratch.data_ptr(), es, s))
    assert torch.equal(torch.view_as_real(h) if cplx else h, torch.view_as_real(h2) if cplx else h2)
    # combine: w -= sum h_j V_j, with the norm
    n2 = torch.empty(1, dtype=torch.float64, device=dev)
    _lib.check(lib.hmb200_krylov_combine(Vd.data_ptr(), ld, k, h.data_ptr(), -1.0, wd.data_ptr(), n, n2.data_ptr(), scratch.data_ptr(), es, s))
    w_ref = w - h.cpu().numpy() @ V[:, :n]
    assert rel_err(wd.cpu().numpy(), w_ref) < 1e-13
    assert abs(n2.item() - np.vdot(w_ref, w_ref).real) <= 1e-13 * max(np.vdot(w_ref, w_ref).real, 1e-300)
    # scale
    _lib.check(lib.hmb200_krylov_scale(wd.data_ptr(), n, n2.data_ptr(), es, s))
    assert rel_err(wd.cpu().numpy(), w_ref / np.sqrt(n2.item())) < 1e-15
    # axpby + norm
    x, y = helpers.rand_vec(rng, n, dt), helpers.rand_vec(rng, n, dt)
    a, b = (1.5 - 0.25j, -0.5 + 2j) if cplx else (1.5, -0.5)
    xd, yd, od = torch.from_numpy(x).to(dev), torch.from_numpy(y).to(dev), torch.empty(n, dtype=Vd.dtype, device=dev)
    ca, cb = complex(a), complex(b)
    _lib.check(lib.hmb200_krylov_axpby((ctypes.c_double * 2)(ca.real, ca.imag), xd.data_ptr(), (ctypes.c_double * 2)(cb.real, cb.imag), yd.data_ptr(),
                                       od.data_ptr(), n, n2.data_ptr(), scratch.data_ptr(), es, s))
    o_ref = a * x + b * y
    assert rel_err(od.cpu().numpy(), o_ref) < 1e-15
    assert abs(n2.item() - np.vdot(o_ref, o_ref).real) <= 1e-13 * np.vdot(o_ref, o_ref).real


@pytest.mark.parametrize("case", range(6))
def test_device_gmres_matches_reference_hpddm_counts(case):
    from htool_python_b200 import B200DistributedOperator, B200Solver

    g = gmres_golden()[case]
    flat, _ = helpers.load_golden(g["fixture"])
    op = B200DistributedOperator(flat, [(0, flat.n_rows)])
    solver = B200Solver(op)
    x = np.zeros(flat.n_rows, flat.np_dtype)
    it = solver.solve(x, g["b"], g["args"])
    assert it == g["iterations"], (it, g["iterations"])
    assert solver.last_info["operator_calls"] == g["calls"]
    check_solution(flat, g, x, case, x_tol=[10 * t for t in X_TOL])
    assert solver.space.kernel_launches > 0 and op.engine.launch_count > 0
    # a second solve from the converged solution: zero iterations are needed only if the residual is already below tol
    it2 = solver.solve(x, g["b"], g["args"])
    assert it2 <= it
    op.close()


def test_device_gmres_block_of_rhs_and_errors():
    from htool_python_b200 import B200DistributedOperator, B200Solver
    from oracle import flat_oracle as fo

    flat, _ = helpers.load_golden("d_SL_cylinder")
    op = B200DistributedOperator(flat, [(0, flat.n_rows)])
    solver = B200Solver(op)
    rng = np.random.default_rng(4)
    B = np.asfortranarray(rng.random((flat.n_rows, 2)))
    X = np.zeros_like(B, order="F")
    solver.solve(X, B, "-hpddm_tol 1e-5 -hpddm_max_it 300 -hpddm_gmres_restart 50 -hpddm_schwarz_method none")
    for c in range(2):
        assert np.linalg.norm(fo.matvec_user(flat, np.ascontiguousarray(X[:, c])) - B[:, c]) / np.linalg.norm(B[:, c]) < 2e-5
    with pytest.raises(RuntimeError, match="Wrong size"):
        solver.solve(np.zeros(flat.n_rows + 1), np.zeros(flat.n_rows + 1))
    op.close()


@pytest.mark.parametrize("case", [0, 3, 4])
def test_device_gmres_over_peer_collectives_single_rank(case):
    """the solver on the peer-memory collectives (hmb200_peer_allgather / hmb200_peer_allreduce_f64: copy kernel + epoch flags,
    one-CTA all-reduce) with one rank: same HPDDM counts; the 2-rank version runs in test_gpu_parity.py::test_distributed_operator"""
    from htool_python_b200 import B200DistributedOperator, B200Solver

    g = gmres_golden()[case]
    flat, _ = helpers.load_golden(g["fixture"])
    op = B200DistributedOperator(flat, [(0, flat.n_rows)], fused_allgather=True)
    solver = B200Solver(op)
    assert op.fused and solver.space.use_peer
    x = np.zeros(flat.n_rows, flat.np_dtype)
    it = solver.solve(x, g["b"], g["args"])
    assert it == g["iterations"] and solver.last_info["operator_calls"] == g["calls"]
    check_solution(flat, g, x, case, x_tol=[10 * t for t in X_TOL])
    assert op.peer_status() == 0
    # the all-reduce alone: in-place, rank-ordered sum (one rank: identity), real and complex
    import torch

    for t in (torch.arange(7, dtype=torch.float64, device="cuda"), torch.arange(5, dtype=torch.float64, device="cuda") * (1 + 2j)):
        ref = t.clone()
        for _ in range(3):
            op.peer_allreduce_(t)
        assert torch.equal(torch.view_as_real(t) if t.is_complex() else t, torch.view_as_real(ref) if ref.is_complex() else ref)
    op.close()


@pytest.mark.parametrize("product", ["dmma", "auto"])
@pytest.mark.parametrize("peer", [False, True])
@pytest.mark.parametrize("case", range(4))
def test_device_block_gmres_matches_reference_hpddm(case, peer, product):
    """blocks of right-hand sides: the lockstep GMRES on the multi-RHS DMMA product == the reference's HPDDM GMRES called with
    mu > 1 (tests/golden/gmres_block_golden.npz): iteration count and number of block operator calls EQUAL, every column's
    iterate within a fraction of the tolerance; `peer` routes the small reductions through the peer-memory all-reduce;
    `product`: "dmma" forces the multi-RHS kernel, "auto" lets these narrow blocks take the streamed product per column."""
    from htool_python_b200 import B200DistributedOperator, B200Solver

    g = gmres_block_golden()[case]
    flat, _ = helpers.load_golden(g["fixture"])
    op = B200DistributedOperator(flat, [(0, flat.n_rows)], fused_allgather=peer)
    solver = B200Solver(op)
    if product == "dmma":
        solver.space.mm_min_cols = 0
    X = np.zeros_like(g["B"], order="F")
    it = solver.solve(X, g["B"], g["args"])
    assert it == g["iterations"] and solver.last_info["operator_calls"] == g["calls"], (it, solver.last_info["operator_calls"])
    check_block_solution(flat, g, X, case, scale=10.0)
    if peer:
        assert op.peer_status() == 0
    op.close()
