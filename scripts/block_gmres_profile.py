"""Where the time of one lockstep block-GMRES iteration goes (N-point Laplace operator, 1 GPU):
  * the row-major multi-RHS product alone for mu = 1..64 (events on torch's stream, 10 products each),
  * the transposes around it, the per-column Krylov kernels of one step,
  * whole block solves (tol 1e-3) with / without the speculative next product, and column by column."""
import os
import sys
import time
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

import bench  # noqa: E402
from htool_python_b200 import B200DistributedOperator, B200HMatrix, B200Solver  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
mus = [int(v) for v in (sys.argv[2].split(",") if len(sys.argv) > 2 else "1,2,4,8,16,32,64".split(","))]
args = types.SimpleNamespace(n=n, eps=1e-3, eta=10.0, leaf=64, kernel="laplace", wavenumber=0.0)
torch.cuda.set_device(0)
tree, H, flat, bt = bench.assemble(args, 1, 0, lambda m: print("[block-profile]", m, file=sys.stderr, flush=True))
hb = B200HMatrix(flat, device=0)
op = B200DistributedOperator.from_engine(hb.engine, flat)
solver = B200Solver(op)
sp = solver.space


def timed(fn, reps=10, warm=2):
    for _ in range(warm):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


for mu in mus:
    X = torch.rand(n * mu, dtype=torch.float64, device="cuda")
    Y = torch.zeros(n * mu, dtype=torch.float64, device="cuda")
    if mu == 1:
        ms = timed(lambda: hb.engine.add_vector_product("N", 1.0, X, 0.0, Y))
    else:
        ms = timed(lambda: hb.engine.add_matrix_product_row_major("N", 1.0, X, 0.0, Y, mu))
    line = f"block-profile: product mu={mu:3d}: {ms:8.3f} ms = {ms / mu:7.3f} ms per right-hand side"
    if mu > 1:
        vin, vout = sp.block(mu), sp.block(mu)
        vin.copy_(torch.rand_like(vin))
        ms2 = timed(lambda: sp.matmat(vin, vout))
        line += f" | space.matmat (transposes + product) {ms2:8.3f} ms"
        VB = sp.block_basis(21, mu)
        VB.data.copy_(torch.rand_like(VB.data))

        def krylov_step(k=20):
            W = VB.block(k)
            h = [sp.dots(VB.col(c), 0, k, W[c], reduce=False) for c in range(mu)]
            n2 = [sp.combine(VB.col(c), 0, k, h[c], -1.0, W[c], want_norm=True, reduce=False) for c in range(mu)]
            for c in range(mu):
                sp.scale(W[c], n2[c])

        ms3 = timed(krylov_step)
        t0 = time.perf_counter()
        for _ in range(10):
            krylov_step()
        host_ms = (time.perf_counter() - t0) * 100  # enqueue time per step (no sync)
        torch.cuda.synchronize()
        line += f" | Krylov kernels of one step (k=20) {ms3:7.3f} ms on device, {host_ms:6.3f} ms to enqueue"
        del VB, vin, vout
    print(line, flush=True)
    del X, Y
    torch.cuda.empty_cache()

mu = 8
B = np.asfortranarray(op @ np.asfortranarray(np.random.default_rng(4).random((n, mu))))
X = np.zeros_like(B, order="F")
opts = "-hpddm_krylov_method gmres -hpddm_schwarz_method none -hpddm_tol 1e-3 -hpddm_gmres_restart 40 -hpddm_max_it 200"
solver.solve(X, B, opts + " -hpddm_max_it 3")
for label, extra, spec in (("lockstep, speculation on", " -hpddm_block_rhs lockstep", True), ("lockstep, speculation off", " -hpddm_block_rhs lockstep", False),
                           ("column by column", " -hpddm_block_rhs columns", True)):
    solver.opt["speculate"] = spec
    for rep in range(2):
        X[:] = 0
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        it = solver.solve(X, B, opts + extra)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
    print(f"block-profile: solve mu={mu} {label:28s}: {it} iterations, {solver.get_information()['nb_mat_vec_prod']} operator calls, {dt * 1e3:8.2f} ms", flush=True)
# the fixed cost of moving the block in and out (user numbering, host arrays)
blk = sp.block(mu)
t0 = time.perf_counter()
sp.set_block_from_user(blk, B)
torch.cuda.synchronize()
t1 = time.perf_counter()
sp.get_block_to_user(blk)
t2 = time.perf_counter()
print(f"block-profile: set_block_from_user {1e3 * (t1 - t0):.2f} ms, get_block_to_user {1e3 * (t2 - t1):.2f} ms (mu={mu})", flush=True)
