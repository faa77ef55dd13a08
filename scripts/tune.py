"""Option sweep on one assembled operator: python scripts/tune.py N 'k=v,k=v' 'k=v' ...  (each arg = one engine config)"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from htool_python_b200 import Engine

class A: pass
a = A(); a.n = int(sys.argv[1]); a.eps = 1e-3; a.eta = 10.0; a.leaf = 64
tree, H, flat, _ = bench.assemble(a, 1, 0, print)
w1, w2 = bench.w_split(flat)
x = torch.rand(a.n, dtype=torch.float64, device="cuda"); y = torch.zeros_like(x)
yref = None
for cfg in sys.argv[2:] or [""]:
    opts = {k: int(v) for k, v in (kv.split("=") for kv in cfg.split(",") if kv)}
    try:
        e = Engine(flat, **opts)
    except Exception as ex:
        print(f"{cfg:60s} FAILED {ex}"); continue
    inf = e.info
    for _ in range(3): e.add_vector_product("N", 1.0, x, 0.0, y)
    torch.cuda.synchronize(); e.enable_timing(True)
    for _ in range(10): e.add_vector_product("N", 1.0, x, 0.0, y)
    torch.cuda.synchronize(); ms = e.kernel_ms_avg()
    if yref is None: yref = y.clone()
    err = float((y - yref).norm() / yref.norm())
    print(f"{cfg:60s} p1 {ms['phase1']:.3f} ms ({8*w1/ms['phase1']/1e6:5.0f} alg GB/s, {inf['stream1_bytes']/ms['phase1']/1e6:5.0f} raw) red {ms['reduce']:.3f} "
          f"p2 {ms['phase2']:.3f} ms ({8*w2/ms['phase2']/1e6:5.0f} alg, {inf['stream2_bytes']/ms['phase2']/1e6:5.0f} raw) total {ms['total']:.3f} "
          f"stages {inf['n_stages1']}/{inf['n_stages2']} pieces {inf['n_pieces1']}/{inf['n_pieces2']} err {err:.1e}", flush=True)
    e.close()
