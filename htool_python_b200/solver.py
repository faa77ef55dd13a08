"""`B200Solver`: the GMRES of the reference's DDM solver with the operator AND the Krylov basis resident on the GPU.

Reference path (paths under /root/reference/lib):
  Python `Solver.solve(x, b, hpddm_args)` (src/htool/solver/solver.hpp:23-60) -> `DDM::solve`
  (htool/include/htool/solvers/ddm.hpp:127-229): permute b to partition numbering, keep the local slab, call
  `HPDDM::IterativeMethod::solve` (:193) -> `IterativeMethod::GMRES` (hpddm/include/HPDDM_GMRES.hpp:30-196) whose operator
  callback is `HPDDMOperator::GMV` (htool/include/htool/wrappers/wrapper_hpddm.hpp:102-145) ->
  `internal_add_distributed_operator_vector_product_local_to_local`
  (htool/include/htool/distributed_operator/linalg/add_distributed_operator_vector_product_local_to_local.hpp:19-89):
  all-gather the local input slabs, then the rank's global-to-local H-matrix product; finally all-gather x and permute back.

Here every rank keeps its slab of b, x and of the Krylov basis in HBM; one iteration is
  all-gather(v_i) -> slab product (sm_100a stream kernels) -> dots / combine / scale kernels (csrc/kernels/krylov_kernels.cuh)
  with two tiny all-reduces, and ONE small device->host read (the new Hessenberg column) for the Givens update and the
  convergence test, which stay on the host exactly as HPDDM does them (HPDDM_iterative.hpp:748-765, :88-91).
The control flow (restart, `hasConverged` sentinel, max_it handling, returned iteration count) follows HPDDM_GMRES.hpp
line by line so that iteration counts match the reference's.  Scope: one right-hand side, or a block of mu right-hand sides
advanced in lockstep on the multi-right-hand-side product (`gmres_block_driver`), no preconditioner (`-hpddm_schwarz_method none`); Schwarz / GenEO
preconditioners are outside the hot path (out of scope).
"""
from __future__ import annotations

import ctypes
import math
import time
from typing import Optional

import numpy as np

from . import _lib

HPDDM_EPS = 1.0e-12  # HPDDM_define.hpp:48


def parse_hpddm_args(args: str, base: Optional[dict] = None) -> dict:
    """The subset of HPDDM's option string the GMRES path reads (HPDDM_iterative.hpp:194-215); unknown keys are kept."""
    opt = dict(base) if base else dict(tol=1.0e-6, max_it=100, gmres_restart=40, orthogonalization="cgs", variant="right",
                                       krylov_method="gmres", schwarz_method="none", verbosity=0)
    toks = (args or "").split()
    i = 0
    while i < len(toks):
        key = toks[i].lstrip("-")
        key = key[6:] if key.startswith("hpddm_") else key
        has_val = i + 1 < len(toks) and not toks[i + 1].startswith("-hpddm")
        val = toks[i + 1] if has_val else "1"
        i += 2 if has_val else 1
        if key == "tol":
            opt[key] = float(val)
        elif key in ("max_it", "gmres_restart", "verbosity"):
            opt[key] = int(float(val))
        else:
            opt[key] = val
    return opt


class _Hessenberg:
    """Host side of Arnoldi for one right-hand side: Givens rotations on the new column and the rotated right-hand side
    (HPDDM_iterative.hpp:748-765); H[k] is column k (length m+1)."""

    def __init__(self, m: int, dtype):
        self.m = m
        self.H = [np.zeros(m + 1, dtype) for _ in range(m)]
        self.s = np.zeros(m + 1, dtype)
        self.sn = np.zeros(m)

    def start(self, beta: float):
        self.s[:] = 0
        self.s[0] = beta

    def push(self, i: int, hcol, norm2: float) -> float:
        """hcol = <V_k, w> for k <= i, norm2 = ||w - sum hcol_k V_k||^2; returns |s[i+1]| (the residual estimate)."""
        H, s, sn = self.H, self.s, self.sn
        Hi = H[i]
        Hi[: i + 1] = hcol
        Hi[i + 1] = math.sqrt(norm2)
        for k in range(i):
            gamma = np.conj(H[k][k + 1]) * Hi[k] + sn[k] * Hi[k + 1]
            Hi[k + 1] = -sn[k] * Hi[k] + H[k][k + 1] * Hi[k + 1]
            Hi[k] = gamma
        delta = math.sqrt(abs(Hi[i]) ** 2 + abs(Hi[i + 1]) ** 2)
        sn[i] = np.real(Hi[i + 1]) / delta
        Hi[i + 1] = Hi[i] / delta
        Hi[i] = delta
        s[i + 1] = -sn[i] * s[i]
        s[i] *= np.conj(Hi[i + 1])
        return abs(s[i + 1])

    def solve(self, dim: int):
        """back substitution on the rotated Hessenberg (computeMin, HPDDM_iterative.hpp:283-297) -> y[:dim]"""
        y = np.array(self.s[:dim])
        for r in range(dim - 1, -1, -1):
            for c in range(r + 1, dim):
                y[r] -= self.H[c][r] * y[c]
            y[r] /= self.H[r][r]
        return y


def gmres_driver(sp, b_host, x_host, opt: dict):
    """Restarted GMRES on the vector space `sp` (see CudaKrylovSpace for the interface), control flow of
    IterativeMethod::GMRES (HPDDM_GMRES.hpp:30-196) for mu = 1.  Returns (x_host, iterations, info)."""
    tol, max_it = float(opt["tol"]), int(opt["max_it"])
    m = min(int(opt["gmres_restart"]), max_it)
    mgs = str(opt.get("orthogonalization", "cgs")).lower() in ("mgs", "1")
    if abs(tol) < np.finfo(np.float64).eps:  # HPDDM_iterative.hpp:223-231
        tol = 4 * np.finfo(np.float64).eps
    b, x = sp.vector(), sp.vector()
    sp.set_from_user(b, b_host)
    sp.set_from_user(x, x_host)
    norm_b = math.sqrt(float(sp.host([sp.norm2(b)])[0].real.reshape(-1)[0]))
    if norm_b < HPDDM_EPS:
        norm_b = 1.0
    V = sp.basis(m + 1)
    hess = _Hessenberg(m, sp.np_dtype)
    calls, residuals, unused = 0, [], 0
    speculate = bool(opt.get("speculate", True)) and hasattr(sp, "host_begin")
    if not hasattr(sp, "host_begin"):  # spaces without an asynchronous read: read synchronously
        sp.host_begin, sp.host_end = (lambda handles: handles), (lambda token: sp.host(token))
    conv = -m
    j = 1
    while j <= max_it:
        sp.matvec(x, V.row(0))
        calls += 1
        beta2_dev = sp.axpby_norm(1.0, b, -1.0, V.row(0), V.row(0))  # V_0 = b - A x
        beta2 = float(sp.host([beta2_dev])[0].real.reshape(-1)[0])
        if j == 1 and beta2 < np.finfo(np.float64).eps ** 2:
            j, conv = 0, 0
            break
        if conv > 0:
            conv = 0
        hess.start(math.sqrt(beta2))
        sp.scale(V.row(0), beta2_dev)
        i = 0
        launched = False  # the product of step i was already enqueued (speculatively) by step i-1
        while i < m and j <= max_it:
            w = V.row(i + 1)
            if not launched:
                sp.matvec(V.row(i), w)
            launched = False
            calls += 1
            if mgs:  # HPDDM_iterative.hpp:522-540
                hs = []
                n2_dev = None
                for k in range(i + 1):
                    hk = sp.dots(V, k, 1, w)
                    n2_dev = sp.combine(V, k, 1, hk, -1.0, w, want_norm=(k == i))
                    hs.append(hk)
                token = sp.host_begin(hs + [n2_dev])
            elif hasattr(sp, "cgs_step"):  # the same two kernels + two reductions on preallocated buffers, one read-back
                h_dev, n2_dev, token = sp.cgs_step(V, i + 1, w)
            else:  # classical Gram-Schmidt (:541-552)
                h_dev = sp.dots(V, 0, i + 1, w)
                n2_dev = sp.combine(V, 0, i + 1, h_dev, -1.0, w, want_norm=True)
                token = sp.host_begin([h_dev, n2_dev])
            if i < m - 1:
                sp.scale(w, n2_dev)
                if speculate and j + 1 <= max_it:
                    # enqueue the NEXT operator call before the host reads the Hessenberg column: the device computes it
                    # while the host does the Givens update and the convergence test.  If this step converges the product
                    # is simply unused (counted in info["speculative_unused"], not in operator_calls).
                    sp.matvec(w, V.row(i + 2))
                    launched = True
            got = sp.host_end(token)
            if mgs:
                hcol, n2 = np.array([g[0] for g in got[:-1]]), float(got[-1].real.reshape(-1)[0])
            else:
                hcol, n2 = got[0], float(got[1].real.reshape(-1)[0])
            res = hess.push(i, hcol, n2)
            i += 1
            residuals.append(res / norm_b)
            if conv == -m and ((tol > 0 and res / norm_b <= tol) or (tol < 0 and res <= -tol)):
                conv = i
            if conv != -m:
                i = 0
                unused += int(launched)
                break
            j += 1
        if j != max_it + 1 and i == m:  # restart
            sp.combine(V, 0, m, sp.upload(hess.solve(m)), 1.0, x, want_norm=False)
        else:
            break
    if j == max_it + 1 and m > 0:
        rem = max_it % m
        if conv < 0:
            conv = rem if rem > 0 else -conv
    if j != 0 and abs(conv) > 0:
        sp.combine(V, 0, abs(conv), sp.upload(hess.solve(abs(conv))), 1.0, x, want_norm=False)
    it = min(j, max_it)
    x_out = sp.get_to_user(x)
    return x_out, it, dict(operator_calls=calls, residuals=residuals, converged=bool(j <= max_it), norm_b=norm_b, speculative_unused=unused)


def gmres_block_driver(sp, B_host, X_host, opt: dict):
    """GMRES on a block of mu right-hand sides, control flow of IterativeMethod::GMRES with mu > 1
    (HPDDM_GMRES.hpp:30-196): mu independent recurrences advanced in LOCKSTEP -- ONE operator call on the n x mu block per
    iteration (the multi-right-hand-side DMMA kernel, as HPDDMOperator::GMV with mu > 1 calls the row-major matrix product,
    wrapper_hpddm.hpp:102-145), one Hessenberg / rotation set / `hasConverged` entry per column (HPDDM_iterative.hpp:88-91,
    :722-768); the loop ends when no column is left at the sentinel, and the solution update uses each column's own dimension
    (computeMin / addSol, :283-326).  Per iteration the host reads the mu new Hessenberg columns in one transfer; the
    reductions of all columns share one all-reduce (orthogonalization :549 reduces the k x mu block at once).
    B_host, X_host: (n, mu).  Returns (X_host_new, iterations, info)."""
    tol, max_it = float(opt["tol"]), int(opt["max_it"])
    m = min(int(opt["gmres_restart"]), max_it)
    mgs = str(opt.get("orthogonalization", "cgs")).lower() in ("mgs", "1")
    eps = np.finfo(np.float64).eps
    if abs(tol) < eps:
        tol = 4 * eps
    mu = int(B_host.shape[1])
    cols = range(mu)
    real0 = lambda a: float(np.asarray(a).real.reshape(-1)[0])  # noqa: E731
    b, x = sp.block(mu), sp.block(mu)
    sp.set_block_from_user(b, B_host)
    sp.set_block_from_user(x, X_host)
    norm_b = [math.sqrt(real0(v)) for v in sp.host(sp.allreduce_many([sp.norm2(b[c], reduce=False) for c in cols]))]
    norm_b = [1.0 if v < HPDDM_EPS else v for v in norm_b]
    VB = sp.block_basis(m + 1, mu)
    hess = [_Hessenberg(m, sp.np_dtype) for _ in cols]
    conv = [-m] * mu
    calls, residuals, unused = 0, [], 0
    speculate = bool(opt.get("speculate", True)) and hasattr(sp, "host_begin")
    if not hasattr(sp, "host_begin"):
        sp.host_begin, sp.host_end = (lambda handles: handles), (lambda token: sp.host(token))

    def update(dims):
        for c in cols:
            if dims[c]:
                sp.combine(VB.col(c), 0, dims[c], sp.upload(hess[c].solve(dims[c])), 1.0, x[c], want_norm=False)

    j = 1
    while j <= max_it:
        V0 = VB.block(0)
        sp.matmat(x, V0)
        calls += 1
        beta2_dev = sp.allreduce_many([sp.axpby_norm(1.0, b[c], -1.0, V0[c], V0[c], reduce=False) for c in cols])  # V_0 = B - A X
        beta2 = [real0(v) for v in sp.host(beta2_dev)]
        if j == 1 and any(v < eps ** 2 for v in beta2):  # HPDDM_GMRES.hpp:87-90, any column
            j, conv = 0, [0] * mu
            break
        for c in cols:
            if conv[c] > 0:
                conv[c] = 0
            hess[c].start(math.sqrt(beta2[c]))
            sp.scale(V0[c], beta2_dev[c])
        i = 0
        launched = False  # the block product of step i was already enqueued (speculatively) by step i-1
        while i < m and j <= max_it:
            W = VB.block(i + 1)
            if not launched:
                sp.matmat(VB.block(i), W)
            launched = False
            calls += 1
            if mgs:  # HPDDM_iterative.hpp:522-540: one reduction of mu values per basis vector
                hk_all = []
                for k in range(i + 1):
                    hk = sp.allreduce_many([sp.dots(VB.col(c), k, 1, W[c], reduce=False) for c in cols])
                    n2_raw = [sp.combine(VB.col(c), k, 1, hk[c], -1.0, W[c], want_norm=(k == i), reduce=False) for c in cols]
                    hk_all.append(hk)
                n2_dev = sp.allreduce_many(n2_raw)
                token = sp.host_begin([hk_all[k][c] for c in cols for k in range(i + 1)] + n2_dev)
            else:  # classical Gram-Schmidt (:541-552)
                h_dev = sp.allreduce_many([sp.dots(VB.col(c), 0, i + 1, W[c], reduce=False) for c in cols])
                n2_dev = sp.allreduce_many([sp.combine(VB.col(c), 0, i + 1, h_dev[c], -1.0, W[c], want_norm=True, reduce=False) for c in cols])
                token = sp.host_begin(h_dev + n2_dev)
            if i < m - 1:
                for c in cols:
                    sp.scale(W[c], n2_dev[c])
                if speculate and j + 1 <= max_it:  # next block product enqueued before the host reads the Hessenberg columns
                    sp.matmat(W, VB.block(i + 2))
                    launched = True
            got = sp.host_end(token)
            hcols = [np.array([got[c * (i + 1) + k][0] for k in range(i + 1)]) for c in cols] if mgs else got[:mu]
            res = [hess[c].push(i, hcols[c], real0(got[len(got) - mu + c])) for c in cols]
            i += 1
            residuals.append([res[c] / norm_b[c] for c in cols])
            for c in cols:
                if conv[c] == -m and ((tol > 0 and res[c] / norm_b[c] <= tol) or (tol < 0 and res[c] <= -tol)):
                    conv[c] = i
            if -m not in conv:
                i = 0
                unused += int(launched)
                break
            j += 1
        if j != max_it + 1 and i == m:  # restart
            update([abs(c) for c in conv])
        else:
            break
    if j == max_it + 1 and m > 0:
        rem = max_it % m
        conv = [(rem if rem > 0 else -c) if c < 0 else c for c in conv]
    if j != 0:
        update([abs(c) for c in conv])
    it = min(j, max_it)
    return sp.get_block_to_user(x), it, dict(operator_calls=calls, residuals=residuals, converged=bool(j <= max_it), norm_b=norm_b,
                                            converged_at=list(conv), speculative_unused=unused, block=mu)


class _Basis:
    def __init__(self, data, n):
        self.data, self.n = data, n  # (count, n) row-contiguous device tensor

    def row(self, i):
        return self.data[i]


class _BlockBasis:
    """(count, mu, n) storage of a block Krylov basis: HPDDM's layout v[i] + nu * n (HPDDM_GMRES.hpp:47-53), so the basis of
    ONE column is a set of vectors with leading dimension mu * n (ldv of the gemv calls, HPDDM_iterative.hpp:542)"""

    def __init__(self, data):
        self.data = data
        self.mu, self.n = int(data.shape[1]), int(data.shape[2])

    def block(self, i):
        return self.data[i]

    def col(self, c):
        return _Basis(self.data[:, c], self.mu * self.n)


class CudaKrylovSpace:
    """Device-resident vectors over one rank's row slab of a `B200DistributedOperator`; every operation is a kernel of
    libhmb200.so (csrc/kernels/krylov_kernels.cuh, stream kernels, permutation kernels) plus, for world > 1, the
    all-gather of the operator input and two small all-reduces (`torch.distributed`, NCCL)."""

    def __init__(self, op):
        import torch

        if op.n_target != op.n_source:
            raise RuntimeError("the solver needs a square operator")
        self.op = op
        self.torch = torch
        self.lib = _lib.lib()
        self.device = op.device
        self.np_dtype = op.np_dtype
        self.dtype = op.torch_dtype
        self.es = op.es
        self.world, self.rank, self.group = op.world, op.rank, op.group
        self.off, self.n_local, self.n_global = op.offs[op.rank], op.sizes[op.rank], op.n_target
        # collectives over the operator's peer-mapped regions (no NCCL call in the iteration) when they are set up
        self.use_peer = bool(getattr(op, "fused", False))
        self._xg = torch.empty(self.n_global, dtype=self.dtype, device=self.device)   # replicated operator input
        self._xu = torch.empty(self.n_global, dtype=self.dtype, device=self.device)   # user-numbering staging
        self._pin = torch.empty(self.n_global, dtype=self.dtype).pin_memory()
        self._scratch = None
        self._hpin, self._hev = None, None
        self._k_max = 0
        self._blk = {}
        self.mm_min_cols = 2  # narrowest block that goes through the multi-RHS (DMMA) product (1 = a plain matvec)
        self.kernel_launches = 0

    def _check_peers(self):
        """after a host-side synchronisation: a wait of the peer collectives that gave up raises here (the Hessenberg
        column just read would otherwise come from partly gathered / partly summed data)"""
        if self.use_peer:
            self.op.check_peers()

    # ---- storage ------------------------------------------------------------------------------------------------
    def vector(self):
        return self.torch.zeros(self.n_local, dtype=self.dtype, device=self.device)

    def basis(self, count):
        self._reserve(count)
        return _Basis(self.torch.zeros((count, self.n_local), dtype=self.dtype, device=self.device), self.n_local)

    def block(self, mu):
        return self.torch.zeros((mu, self.n_local), dtype=self.dtype, device=self.device)

    def block_basis(self, count, mu):
        self._reserve(count)
        return _BlockBasis(self.torch.zeros((count, mu, self.n_local), dtype=self.dtype, device=self.device))

    def set_block_from_user(self, blk, host):
        host = np.asarray(host)
        if host.ndim != 2 or host.shape != (self.n_global, blk.shape[0]):
            raise RuntimeError("Wrong size for right-hand side or solution")
        for c in range(blk.shape[0]):
            self.set_from_user(blk[c], host[:, c])

    def get_block_to_user(self, blk):
        out = np.empty((self.n_global, blk.shape[0]), dtype=self.np_dtype, order="F")
        for c in range(blk.shape[0]):
            out[:, c] = self.get_to_user(blk[c])
        return out

    def _reserve(self, k):
        if self._scratch is None or k > self._k_max:
            self._k_max = max(k, 1)
            self._scratch = self.torch.empty(int(self.lib.hmb200_krylov_scratch_elems(self.n_local, self._k_max)), dtype=self.dtype, device=self.device)

    def _stream(self):
        return self.torch.cuda.current_stream(self.device).cuda_stream

    def _permute(self, gather, src, dst):
        fn = self.lib.hmb200_permute_gather if gather else self.lib.hmb200_permute_scatter
        _lib.check(fn(self.op.perm_t.data_ptr(), 0, self.n_global, src.data_ptr(), dst.data_ptr(), self.es, 1, _lib.DEVICE, self._stream()))
        self.kernel_launches += 1

    def set_from_user(self, vec, host):
        """global_to_partition_numbering + local slice (ddm.hpp:166-172)"""
        host = np.ascontiguousarray(host, dtype=self.np_dtype)
        if host.shape != (self.n_global,):
            raise RuntimeError("Wrong size for right-hand side or solution")
        with self.torch.cuda.device(self.device):
            self._pin.numpy()[:] = host
            self._xu.copy_(self._pin, non_blocking=True)
            self._permute(True, self._xu, self._xg)
            vec.copy_(self._xg[self.off:self.off + self.n_local])

    def get_to_user(self, vec):
        """all-gather of the local slabs + partition_to_global_numbering (ddm.hpp:198-223)"""
        from .distributed import allgather_slabs_

        with self.torch.cuda.device(self.device):
            if self.use_peer:
                xg = self.op.peer_allgather_(vec)
            else:
                xg = self._xg
                xg[self.off:self.off + self.n_local].copy_(vec)
                allgather_slabs_(xg, self.op.offs, self.op.sizes, 1, self.group)
            self._permute(False, xg, self._xu)
            self._pin.copy_(self._xu, non_blocking=True)
            self.torch.cuda.current_stream(self.device).synchronize()
        self._check_peers()
        return self._pin.numpy().copy()

    # ---- operator -----------------------------------------------------------------------------------------------
    def matvec(self, vin, vout):
        """vout = A_p * allgather(vin): the local-to-local product (add_distributed_operator_vector_product_local_to_local.hpp:19-58)"""
        from .distributed import allgather_slabs_

        with self.torch.cuda.device(self.device):
            if self.use_peer:
                self.op.engine.add_vector_product("N", 1.0, self.op.peer_allgather_(vin), 0.0, vout)
            elif self.world == 1:
                self.op.engine.add_vector_product("N", 1.0, vin, 0.0, vout)
            else:
                self._xg[self.off:self.off + self.n_local].copy_(vin)
                allgather_slabs_(self._xg, self.op.offs, self.op.sizes, 1, self.group)
                self.op.engine.add_vector_product("N", 1.0, self._xg, 0.0, vout)

    def matmat(self, vin, vout):
        """vout = A_p * allgather(vin) for (mu, n_local) column blocks: HPDDMOperator::GMV with mu > 1 transposes to row-major,
        calls the row-major local-to-local matrix product and transposes back (wrapper_hpddm.hpp:108-116, :129-135;
        add_distributed_operator_matrix_product_row_major_local_to_local.hpp:25-95).  The product is the multi-RHS DMMA kernel."""
        from .distributed import allgather_slabs_

        mu = int(vin.shape[0])
        if mu < self.mm_min_cols:
            for c in range(mu):
                self.matvec(vin[c], vout[c])
            return
        if mu not in self._blk:
            self._blk[mu] = (self.torch.empty(self.n_global * mu, dtype=self.dtype, device=self.device),
                             self.torch.empty(self.n_local * mu, dtype=self.dtype, device=self.device))
        xr, yr = self._blk[mu]
        with self.torch.cuda.device(self.device):
            # column blocks (mu, n_local) <-> row-major slab rows: one layout kernel each way (hmb200_block_to_rows / rows_to_block)
            _lib.check(self.lib.hmb200_block_to_rows(None, self.n_local, mu, vin.data_ptr(), self.n_local, xr.data_ptr() + self.es * mu * self.off, self.es, self._stream()))
            if self.world > 1:
                allgather_slabs_(xr, self.op.offs, self.op.sizes, mu, self.group)
            self.op.engine.add_matrix_product_row_major("N", 1.0, xr, 0.0, yr, mu)
            _lib.check(self.lib.hmb200_rows_to_block(None, self.n_local, mu, yr.data_ptr(), vout.data_ptr(), self.n_local, self.es, self._stream()))
            self.kernel_launches += 2

    # ---- reductions ---------------------------------------------------------------------------------------------
    def allreduce_many(self, handles):
        """one all-reduce for a list of small device results (the k x mu block of HPDDM_iterative.hpp:549); returns views"""
        if self.world == 1 and not self.use_peer:
            return handles
        flat = self._allreduce(self.torch.cat([h.reshape(-1) for h in handles]))  # one dtype per call (all dots or all norms)
        out, p = [], 0
        for h in handles:
            out.append(flat[p:p + h.numel()])
            p += h.numel()
        return out

    def _allreduce(self, t):
        if self.use_peer and t.numel() * (2 if t.is_complex() else 1) <= self.op.PEER_ALLREDUCE_MAX:
            return self.op.peer_allreduce_(t)
        if self.world > 1:
            import torch.distributed as dist

            dist.all_reduce(self.torch.view_as_real(t) if t.is_complex() else t, group=self.group)
        return t

    def cgs_step(self, V, k, w):
        """one classical Gram-Schmidt step against the first k basis vectors (HPDDM_iterative.hpp:541-552): h = V^H w, all-reduce,
        w -= V h with ||w||^2, all-reduce, and the asynchronous read-back of [h | ||w||^2] -- h and the norm live next to each
        other in one preallocated device buffer, so the step allocates nothing and reads back with one copy."""
        torch = self.torch
        c = 2 if self.dtype == torch.complex128 else 1
        if getattr(self, "_hbuf", None) is None or self._hbuf.numel() < c * k + 1:
            self._hbuf = torch.empty(max(c * (k + 64) + 1, 256), dtype=torch.float64, device=self.device)
            self._hpin2 = torch.empty(self._hbuf.numel(), dtype=torch.float64).pin_memory()
            self._hev = self._hev or torch.cuda.Event()
        hb = self._hbuf
        h = torch.view_as_complex(hb[:2 * k].view(k, 2)) if c == 2 else hb[:k]
        n2 = hb[c * k:c * k + 1]
        self._reserve(k)
        st = self._stream()
        _lib.check(self.lib.hmb200_krylov_dots(V.row(0).data_ptr(), V.n, k, w.data_ptr(), self.n_local, h.data_ptr(), self._scratch.data_ptr(), self.es, st))
        self._allreduce(h)
        _lib.check(self.lib.hmb200_krylov_combine(V.row(0).data_ptr(), V.n, k, h.data_ptr(), -1.0, w.data_ptr(), self.n_local, n2.data_ptr(), self._scratch.data_ptr(), self.es, st))
        self._allreduce(n2)
        self.kernel_launches += 4
        self._hpin2[:c * k + 1].copy_(hb[:c * k + 1], non_blocking=True)
        self._hev.record(torch.cuda.current_stream(self.device))
        return h, n2, ("flat", c == 2, k)

    def dots(self, V, first, k, w, reduce=True):
        h = self.torch.empty(k, dtype=self.dtype, device=self.device)
        self._reserve(k)
        _lib.check(self.lib.hmb200_krylov_dots(V.row(first).data_ptr(), V.n, k, w.data_ptr(), self.n_local, h.data_ptr(),
                                               self._scratch.data_ptr(), self.es, self._stream()))
        self.kernel_launches += 2
        return self._allreduce(h) if reduce else h

    def combine(self, V, first, k, h, sign, w, want_norm, reduce=True):
        n2 = self.torch.empty(1, dtype=self.torch.float64, device=self.device) if want_norm else None
        self._reserve(k)
        _lib.check(self.lib.hmb200_krylov_combine(V.row(first).data_ptr(), V.n, k, h.data_ptr(), float(sign), w.data_ptr(), self.n_local,
                                                  n2.data_ptr() if want_norm else None, self._scratch.data_ptr(), self.es, self._stream()))
        self.kernel_launches += 2 if want_norm else 1
        return (self._allreduce(n2) if reduce else n2) if want_norm else None

    def axpby_norm(self, alpha, x, beta, y, out, reduce=True):
        n2 = self.torch.empty(1, dtype=self.torch.float64, device=self.device)
        self._reserve(1)
        a, b = complex(alpha), complex(beta)
        _lib.check(self.lib.hmb200_krylov_axpby((ctypes.c_double * 2)(a.real, a.imag), x.data_ptr(), (ctypes.c_double * 2)(b.real, b.imag), y.data_ptr(),
                                                out.data_ptr(), self.n_local, n2.data_ptr(), self._scratch.data_ptr(), self.es, self._stream()))
        self.kernel_launches += 2
        return self._allreduce(n2) if reduce else n2

    def norm2(self, x, reduce=True):
        tmp = self.torch.empty_like(x)
        return self.axpby_norm(1.0, x, 0.0, x, tmp, reduce=reduce)

    def scale(self, w, n2_dev):
        _lib.check(self.lib.hmb200_krylov_scale(w.data_ptr(), self.n_local, n2_dev.data_ptr(), self.es, self._stream()))
        self.kernel_launches += 1

    # ---- small host <-> device traffic ----------------------------------------------------------------------------
    def host(self, handles):
        """one device->host read of a few scalars (the Hessenberg column): the only sync of an iteration"""
        flat = self.torch.cat([self.torch.view_as_real(h).reshape(-1) if h.is_complex() else h.reshape(-1).to(self.torch.float64) for h in handles])
        vals = flat.cpu().numpy()
        self._check_peers()
        out, p = [], 0
        for h in handles:
            if h.is_complex():
                out.append(vals[p:p + 2 * h.numel()].view(np.complex128).copy())
                p += 2 * h.numel()
            else:
                out.append(vals[p:p + h.numel()].copy())
                p += h.numel()
        return out

    def host_begin(self, handles):
        """enqueues the device->host copy of a few scalars into page-locked memory and records an event; `host_end` waits for
        that event only, so kernels enqueued in between (the speculative next product) keep the device busy"""
        torch = self.torch
        flat = torch.cat([torch.view_as_real(h).reshape(-1) if h.is_complex() else h.reshape(-1).to(torch.float64) for h in handles])
        if self._hpin is None or self._hpin.numel() < flat.numel():
            self._hpin = torch.empty(max(2 * flat.numel(), 1024), dtype=torch.float64).pin_memory()
            self._hev = torch.cuda.Event()
        self._hpin[:flat.numel()].copy_(flat, non_blocking=True)
        self._hev.record(torch.cuda.current_stream(self.device))
        return [(h.is_complex(), h.numel()) for h in handles]

    def host_end(self, token):
        self._hev.synchronize()
        self._check_peers()
        if isinstance(token, tuple) and token and token[0] == "flat":  # cgs_step: [h (k values) | ||w||^2]
            _, is_c, k = token
            vals = self._hpin2.numpy()
            if is_c:
                return [vals[:2 * k].view(np.complex128).copy(), vals[2 * k:2 * k + 1].copy()]
            return [vals[:k].copy(), vals[k:k + 1].copy()]
        vals = self._hpin.numpy()
        out, p = [], 0
        for is_c, n in token:
            if is_c:
                out.append(vals[p:p + 2 * n].view(np.complex128).copy())
                p += 2 * n
            else:
                out.append(vals[p:p + n].copy())
                p += n
        return out

    def upload(self, y):
        return self.torch.from_numpy(np.ascontiguousarray(y, dtype=self.np_dtype)).to(self.device)


class B200Solver:
    """Drop-in for the solve surface of the reference's `Htool.Solver` / `ComplexSolver` (`solve(x, b, hpddm_args)`,
    `set_hpddm_args`, `get_information`; src/htool/solver/solver.hpp:23-64) for the unpreconditioned GMRES path, on a
    `B200DistributedOperator` (a single-process `B200HMatrix` is wrapped with `B200DistributedOperator.from_engine(h.engine, flat)`)."""

    def __init__(self, distributed_operator, space=None):
        self.op = distributed_operator
        self.space = space if space is not None else CudaKrylovSpace(distributed_operator)
        self.opt = parse_hpddm_args("")
        self.infos = {}

    def set_hpddm_args(self, hpddm_args: str):
        self.opt = parse_hpddm_args(hpddm_args, self.opt)

    def get_information(self):
        return dict(self.infos)

    def solve(self, x, b, hpddm_args: str = ""):
        """x (in/out, initial guess) and b in USER numbering; 1-D, or 2-D (n, mu) Fortran-ordered blocks, whose columns advance
        in lockstep with one block operator call per iteration like HPDDM's GMRES with mu > 1 (`-hpddm_block_rhs columns`, an
        option of this class only, solves them one after the other instead).  x is updated IN PLACE like the reference
        (py::arg("x").noconvert())."""
        self.set_hpddm_args(hpddm_args)
        if str(self.opt.get("krylov_method", "gmres")).lower() not in ("gmres", "0"):
            raise NotImplementedError("only -hpddm_krylov_method gmres is on the accelerated path")
        if str(self.opt.get("schwarz_method", "none")).lower() not in ("none",):
            raise NotImplementedError("Schwarz preconditioners are outside the accelerated path (use -hpddm_schwarz_method none)")
        x_arr, b_arr = np.asarray(x), np.asarray(b)
        if not ((b_arr.ndim == 1 and x_arr.ndim == 1) or (b_arr.ndim == 2 and x_arr.ndim == 2 and b_arr.shape[1] == x_arr.shape[1])):
            raise ValueError(f"Wrong dimension for right-hand side or solution\nright-hand side: {b_arr.shape}\nsolution: {x_arr.shape}\n")
        if not isinstance(x, np.ndarray) or x.dtype != self.space.np_dtype:
            raise TypeError("x must be a numpy array of the operator's dtype (it is updated in place)")
        t0 = time.perf_counter()
        its, calls = [], 0
        lockstep = b_arr.ndim == 2 and b_arr.shape[1] > 1 and str(self.opt.get("block_rhs", "lockstep")).lower() != "columns"
        if lockstep:  # HPDDM's GMRES with mu > 1: all columns advance together, one block operator call per iteration
            sol, it, info = gmres_block_driver(self.space, b_arr, x_arr, self.opt)
            x_arr[...] = sol
            its, calls, self.last_info = [it], info["operator_calls"], info
        else:
            cols = [(x_arr, b_arr)] if b_arr.ndim == 1 else [(x_arr[:, c], b_arr[:, c]) for c in range(b_arr.shape[1])]
            for xc, bc in cols:
                sol, it, info = gmres_driver(self.space, bc, xc, self.opt)
                xc[...] = sol
                its.append(it)
                calls += info["operator_calls"]
                self.last_info = info
        self.infos.update({"Solve": str(time.perf_counter() - t0), "Nb_it": str(max(its)), "Nb_subdomains": str(getattr(self.space, "world", 1)),
                           "Precond": "None", "nb_mat_vec_prod": str(calls)})
        return max(its)
