"""Per-rank cost of the P-way block-row split, measured on ONE GPU: for each rank r of a P-way partition of the N-point
operator, assemble slab r with the reference builder, pack it, and time (a) the kernels (CUDA events inside the library),
(b) back-to-back device-resident products replayed from the library's CUDA graphs, (c) the same products launched eagerly.  The all-gather is not included.

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
from htool_python_b200 import Engine, _lib  # noqa: E402

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

    def timed(n_steps=50):
        for _ in range(5):
            step()
        torch.cuda.synchronize()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        for _ in range(n_steps):
            step()
        ev1.record()
        torch.cuda.synchronize()
        return ev0.elapsed_time(ev1) / n_steps

    e.enable_timing(True)   # per-kernel CUDA events (forces eager launches)
    ms_events = timed()
    kms = e.kernel_ms_avg()
    e.enable_timing(False)
    ms_step = timed()       # replayed from the library's CUDA graphs
    y_graph = y.clone()
    # the same products with the library's CUDA graphs switched off (eager launches)
    _lib.check(_lib.lib().hmb200_set_graphs(e._h, 0))
    ms_eager = timed()
    same = bool(torch.equal(y, y_graph))
    ms_graph = ms_step
    row = dict(rank=r, P=P, n=n, rows=flat.n_rows, W=flat.stored_coefficients, leaves=flat.n_leaves, kernel_ms={k: round(v, 4) for k, v in kms.items()},
               ms_graph=round(ms_graph, 4), ms_eager=round(ms_eager, 4), ms_eager_with_events=round(ms_events, 4), graph_same=same, GBps_step=round(flat.algorithmic_bytes(1) / ms_step / 1e6, 1), **bt)
    print(json.dumps(row), flush=True)
    rows.append(row)
    e.close()
    del H, flat, e
if rows:
    print(json.dumps(dict(summary=True, P=P, n=n, max_ms_eager=max(r["ms_eager"] for r in rows), max_ms_graph=max(r["ms_graph"] for r in rows),
                          sum_W=sum(r["W"] for r in rows), max_W=max(r["W"] for r in rows))), flush=True)
