"""multi-RHS timing on one assembled operator: python scripts/tune_mm.py N MU [engine opts ...]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
import numpy as np, torch
import bench
from htool_python_b200 import Engine

class A: pass
a = A(); a.n = int(sys.argv[1]); a.eps = 1e-3; a.eta = 10.0; a.leaf = 64; a.kernel = os.environ.get("KERNEL", "laplace"); a.wavenumber = 0.0
mu = int(sys.argv[2])
tree, H, flat, _ = bench.assemble(a, 1, 0, print)
tdt = torch.float64 if flat.dtype == "d" else torch.complex128
X = torch.rand(a.n, mu, dtype=torch.float64, device="cuda").to(tdt); Y = torch.zeros(a.n, mu, dtype=tdt, device="cuda")
W = flat.stored_coefficients
fl = (2 if flat.dtype == "d" else 8) * mu * W
for cfg in sys.argv[3:] or [""]:
    opts = {k: int(v) for k, v in (kv.split("=") for kv in cfg.split(",") if kv)}
    e = Engine(flat, **opts)
    inf = e.info
    print('   mm info:', {k: inf[k] for k in ('mm_stream_bytes','mm_tiles','mm_stages','mm_rows','n_pieces2','n_stages2','workspace_bytes')}, 'sum_r rows', int(np.maximum(flat.rank,0).sum()))
    for _ in range(2): e.add_matrix_product_row_major("N", 1.0, X, 0.0, Y, mu)
    torch.cuda.synchronize(); e.enable_timing(True)
    t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(5): e.add_matrix_product_row_major("N", 1.0, X, 0.0, Y, mu)
    t1.record(); torch.cuda.synchronize(); ms = e.kernel_ms_avg(); tot = t0.elapsed_time(t1) / 5
    # check a few columns against the single-RHS streamed product
    err = 0.0
    for j in (0, mu - 1):
        y1 = torch.zeros(a.n, dtype=tdt, device="cuda"); e.add_vector_product("N", 1.0, X[:, j].contiguous(), 0.0, y1)
        err = max(err, float((Y[:, j] - y1).norm() / y1.norm()))
    print(f"{cfg:40s} mu={mu} per-pass: p1 {ms['phase1']:.3f} red {ms['reduce']:.3f} p2 {ms['phase2']:.3f} ms | whole matmat {tot:.3f} ms = {fl/tot/1e9:.2f} TFLOP/s "
          f"({100*fl/tot/1e9/37.0:.0f}% of 37 TF DMMA), {mu/tot*1e3:.0f} RHS/s, err vs matvec {err:.1e}", flush=True)
    e.close()
if os.environ.get("CPU_REF"):
    import helpers_ref
    ro = helpers_ref.ref_ops()
    Xh = np.ascontiguousarray(X.cpu().numpy()); Yh = np.zeros_like(Xh)
    fn = getattr(ro, "internal_matrix_product_row_major_addr_" + flat.dtype)
    t = time.perf_counter(); fn(H.address, "N", 1.0, Xh, 0.0, Yh, mu, True); dt = time.perf_counter() - t
    print(f"reference OpenMP matmat mu={mu}: {dt*1e3:.0f} ms; rel err GPU vs reference {np.linalg.norm(Y.cpu().numpy()-Yh)/np.linalg.norm(Yh):.2e}")
