"""TEST-ONLY stand-in for `htool_python_b200.solver.CudaKrylovSpace` on CPU: same interface, numpy vectors, the ORACLE as
slab product, `gloo` all-gather / all-reduce.  It exists so that the HOST logic of the solver (HPDDM's GMRES control flow,
Givens update, restart, slab bookkeeping, collectives) is exercised without a GPU; it is never imported by the package."""
import numpy as np


class _Basis:
    def __init__(self, data, n):
        self.data, self.n = data, n

    def row(self, i):
        return self.data[i]


class NumpyKrylovSpace:
    def __init__(self, slab_flat, offs, sizes, rank=0, world=1):
        from oracle import flat_oracle as fo

        self.fo, self.flat = fo, slab_flat
        self.np_dtype = slab_flat.np_dtype
        self.world, self.rank = world, rank
        self.offs, self.sizes = offs, sizes
        self.off, self.n_local, self.n_global = offs[rank], sizes[rank], sum(sizes)
        self.perm = np.asarray(slab_flat.target_permutation)
        self.xg = np.zeros(self.n_global, self.np_dtype)

    def vector(self):
        return np.zeros(self.n_local, self.np_dtype)

    def basis(self, count):
        return _Basis(np.zeros((count, self.n_local), self.np_dtype), self.n_local)

    def block(self, mu):
        return np.zeros((mu, self.n_local), self.np_dtype)

    def block_basis(self, count, mu):
        from htool_python_b200.solver import _BlockBasis

        return _BlockBasis(np.zeros((count, mu, self.n_local), self.np_dtype))

    def set_block_from_user(self, blk, host):
        for c in range(blk.shape[0]):
            self.set_from_user(blk[c], np.asarray(host)[:, c])

    def get_block_to_user(self, blk):
        return np.asfortranarray(np.stack([self.get_to_user(blk[c]) for c in range(blk.shape[0])], axis=1))

    def matmat(self, vin, vout):
        mu = vin.shape[0]
        xr = np.ascontiguousarray(np.stack([self._gather(vin[c]).copy() for c in range(mu)], axis=1))  # row-major n_global x mu
        yr = np.zeros((self.n_local, mu), self.np_dtype)
        self.fo.add_matrix_product_row_major(self.flat, "N", 1.0, xr, 0.0, yr, mu)
        vout[:] = yr.T

    def allreduce_many(self, handles):
        flat = self._allreduce(np.concatenate([np.asarray(h).reshape(-1) for h in handles]))
        out, p = [], 0
        for h in handles:
            out.append(flat[p:p + np.size(h)])
            p += np.size(h)
        return out

    def set_from_user(self, vec, host):
        vec[:] = np.asarray(host, dtype=self.np_dtype)[self.perm][self.off:self.off + self.n_local]

    def _gather(self, vec):
        self.xg[self.off:self.off + self.n_local] = vec
        if self.world > 1:
            import torch

            from htool_python_b200.distributed import allgather_slabs_

            t = torch.from_numpy(self.xg)
            allgather_slabs_(t, self.offs, self.sizes, 1)
        return self.xg

    def get_to_user(self, vec):
        out = np.empty(self.n_global, self.np_dtype)
        out[self.perm] = self._gather(vec)
        return out

    def matvec(self, vin, vout):
        self.fo.add_vector_product(self.flat, "N", 1.0, np.ascontiguousarray(self._gather(vin)), 0.0, vout)

    def _allreduce(self, a):
        if self.world > 1:
            import torch
            import torch.distributed as dist

            t = torch.from_numpy(a.view(np.float64))
            dist.all_reduce(t)
        return a

    def dots(self, V, first, k, w, reduce=True):
        h = np.ascontiguousarray(V.data[first:first + k].conj() @ w)
        return self._allreduce(h) if reduce else h

    def combine(self, V, first, k, h, sign, w, want_norm, reduce=True):
        w += sign * (h @ V.data[first:first + k])
        if not want_norm:
            return None
        n2 = np.array([np.vdot(w, w).real])
        return self._allreduce(n2) if reduce else n2

    def axpby_norm(self, alpha, x, beta, y, out, reduce=True):
        out[:] = alpha * x + beta * y
        n2 = np.array([np.vdot(out, out).real])
        return self._allreduce(n2) if reduce else n2

    def norm2(self, x, reduce=True):
        n2 = np.array([np.vdot(x, x).real])
        return self._allreduce(n2) if reduce else n2

    def scale(self, w, n2):
        w /= np.sqrt(n2[0])

    def host(self, handles):
        return [np.array(h) for h in handles]

    def host_begin(self, handles):  # snapshot now (the device version copies at this point of the stream)
        return [np.array(h) for h in handles]

    def host_end(self, token):
        return token

    def upload(self, y):
        return np.asarray(y, dtype=self.np_dtype)
