"""Per-rank cost of the P-way block-row split, measured on ONE GPU: for each rank r of a P-way partition of the N-point
operator, assemble slab r with the reference builder, pack it, and time (a) the kernels (CUDA events inside the library),
(b) back-to-back device-resident products, (c) the same product replayed from a CUDA graph.  The all-gather is not included.

    python scripts/slab_scan.py [N] [P] [ranks...]
"""
import json
import os
import sys
import time
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

import bench  # noqa: E402
from htool_python_b200 import Engine  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
P = int(sys.argv[2]) if len(sys.argv) > 2 else 8
ranks = [int(a) for a in sys.argv[3:]] or list(range(P))
args = types.SimpleNamespace(n=n, eps=1e-3, eta=10.0, leaf=64, kernel="laplace", wavenumber=0.0)
dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
rows = []
for r in ranks:
    tree, H, flat, bt = bench.assemble(args, P, r, lambda m: print("[scan]", m, file=sys.stderr, flush=True))
    e = Engine(flat, device=0)
    x = torch.rand(n, dtype=torch.float64, device=dev)
    y = torch.zeros(flat.n_rows, dtype=torch.float64, device=dev)

    def step():
        e.add_vector_product("N", 1.0, x, 0.0, y)

    for _ in range(5):
        step()
    torch.cuda.synchronize()
    e.enable_timing(True)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(50):
        step()
    ev1.record()
    torch.cuda.synchronize()
    kms = e.kernel_ms_avg()
    e.enable_timing(False)
    ms_step = ev0.elapsed_time(ev1) / 50
    y_eager = y.clone()
    # CUDA graph of one product on a side stream
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        step()
    torch.cuda.current_stream().wait_stream(s)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=s):
        step()
    for _ in range(5):
        g.replay()
    torch.cuda.synchronize()
    ev0.record()
    for _ in range(50):
        g.replay()
    ev1.record()
    torch.cuda.synchronize()
    ms_graph = ev0.elapsed_time(ev1) / 50
    same = bool(torch.equal(y, y_eager))
    row = dict(rank=r, P=P, n=n, rows=flat.n_rows, W=flat.stored_coefficients, leaves=flat.n_leaves, kernel_ms={k: round(v, 4) for k, v in kms.items()},
               ms_step=round(ms_step, 4), ms_graph=round(ms_graph, 4), graph_same=same, GBps_step=round(flat.algorithmic_bytes(1) / ms_step / 1e6, 1), **bt)
    print(json.dumps(row), flush=True)
    rows.append(row)
    e.close()
    del H, flat, e
if rows:
    print(json.dumps(dict(summary=True, P=P, n=n, max_ms_step=max(r["ms_step"] for r in rows), max_ms_graph=max(r["ms_graph"] for r in rows),
                          sum_W=sum(r["W"] for r in rows), max_W=max(r["W"] for r in rows))), flush=True)
