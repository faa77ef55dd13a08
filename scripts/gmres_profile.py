"""One device-resident GMRES cycle on the N-point Laplace operator (for ncu launch lists / captures of the Krylov kernels)."""
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
max_it = int(sys.argv[2]) if len(sys.argv) > 2 else 45
args = types.SimpleNamespace(n=n, eps=1e-3, eta=10.0, leaf=64, kernel="laplace", wavenumber=0.0)
torch.cuda.set_device(0)
tree, H, flat, bt = bench.assemble(args, 1, 0, lambda m: print("[gmres-profile]", m, file=sys.stderr, flush=True))
hb = B200HMatrix(flat, device=0)
op = B200DistributedOperator.from_engine(hb.engine, flat)
solver = B200Solver(op)
b = op.matvec(np.random.default_rng(3).random(n))
x = np.zeros(n)
opts = f"-hpddm_krylov_method gmres -hpddm_schwarz_method none -hpddm_tol 1e-12 -hpddm_gmres_restart 40 -hpddm_max_it {max_it}"
import ctypes  # noqa: E402

from htool_python_b200 import _lib  # noqa: E402


def run(label, extra=""):
    x[:] = 0
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    it = solver.solve(x, b, opts + extra)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print(f"gmres-profile: {label:34s} n={n} iterations={it} solve {dt * 1e3:8.2f} ms = {dt * 1e3 / it:.3f} ms/iteration", flush=True)


for k in range(3):
    run(f"solve #{k + 1} (graphs on, speculation on)")
if os.environ.get("GMRES_PROFILE_VARIANTS"):
    solver.opt["speculate"] = False
    run("speculation off")
    solver.opt["speculate"] = True
    _lib.check(_lib.lib().hmb200_set_graphs(hb.engine._h, 0))
    run("graphs off")
    run("graphs off (2nd)")
    # the product alone, back to back, for reference
    xc = torch.rand(n, dtype=torch.float64, device="cuda")
    yc = torch.zeros(n, dtype=torch.float64, device="cuda")
    for _ in range(3):
        hb.engine.add_vector_product("N", 1.0, xc, 0.0, yc)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(50):
        hb.engine.add_vector_product("N", 1.0, xc, 0.0, yc)
    torch.cuda.synchronize()
    print(f"gmres-profile: product alone {(time.perf_counter() - t0) * 20:.3f} ms", flush=True)
if os.environ.get("GMRES_PROFILE_VARIANTS"):
    # the device pipeline of one iteration without any host read: product + dots + combine (+norm) + scale on a 20-vector basis
    sp = solver.space
    V = sp.basis(41)
    V.data.normal_()
    _lib.check(_lib.lib().hmb200_set_graphs(hb.engine._h, 1))

    def pipeline(n_it, with_krylov=True):
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        ev0.record()
        for it_ in range(n_it):
            i = it_ % 39
            w = V.row(i + 1)
            sp.matvec(V.row(i), w)
            if with_krylov:
                h = sp.dots(V, 0, 20, w)
                n2 = sp.combine(V, 0, 20, h, -1.0, w, want_norm=True)
                sp.scale(w, n2)
        ev1.record()
        torch.cuda.synchronize()
        return ev0.elapsed_time(ev1) / n_it

    pipeline(80)
    print(f"gmres-profile: device pipeline without host reads: product only {pipeline(120, False):.3f} ms, product + dots/combine/scale (k=20) {pipeline(120):.3f} ms per iteration", flush=True)
