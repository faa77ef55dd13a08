"""Width sweep of the multi-RHS product on one assembled operator: per-pass kernel times (CUDA events around every kernel) and the
whole row-major product for each block width, next to the single-RHS streamed product.
    usage: python scripts/mm_width_sweep.py [N] [mu,mu,...] [engine opts k=v,k=v]"""
import os
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402
from htool_python_b200 import B200HMatrix  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
mus = [int(v) for v in (sys.argv[2] if len(sys.argv) > 2 else "1,2,4,8,16,32,64").split(",")]
opts = {k: int(v) for k, v in (kv.split("=") for kv in (sys.argv[3] if len(sys.argv) > 3 else "").split(",") if kv)}
args = types.SimpleNamespace(n=n, eps=1e-3, eta=10.0, leaf=64, kernel=os.environ.get("KERNEL", "laplace"), wavenumber=0.0)
torch.cuda.set_device(0)
tree, H, flat, bt = bench.assemble(args, 1, 0, lambda m: print("[mm-sweep]", m, file=sys.stderr, flush=True))
dmma, dfma = bench.fp64_rates(0)
hb = B200HMatrix(flat, device=0, **opts)
e = hb.engine
tdt = torch.float64 if flat.dtype == "d" else torch.complex128
W = flat.stored_coefficients
print(f"mm-sweep: N={n} {args.kernel} W={W:.4e} opts={opts}  measured DMMA {dmma} TFLOP/s, DFMA {dfma} TFLOP/s")
print("mm-sweep: | mu | phase 1 ms | reduce ms | phase 2 ms | product ms | ms per RHS | TFLOP/s | of DMMA rate |")
for mu in mus:
    X = torch.rand(n * mu, dtype=torch.float64, device="cuda").to(tdt)
    Y = torch.zeros(n * mu, dtype=tdt, device="cuda")
    fn = (lambda: e.add_vector_product("N", 1.0, X, 0.0, Y)) if mu == 1 else (lambda: e.add_matrix_product_row_major("N", 1.0, X, 0.0, Y, mu))
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        fn()
    e1.record()
    torch.cuda.synchronize()
    tot = e0.elapsed_time(e1) / 10
    e.enable_timing(True)
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    k = e.kernel_ms_avg()
    e.enable_timing(False)
    fl = (2 if flat.dtype == "d" else 8) * mu * W
    print(f"mm-sweep: | {mu} | {k['phase1']:.3f} | {k['reduce']:.3f} | {k['phase2']:.3f} | {tot:.3f} | {tot / mu:.3f} | {fl / tot / 1e9:.2f} | {fl / tot / 1e9 / dmma:.2f} |", flush=True)
    del X, Y
hb.close()
