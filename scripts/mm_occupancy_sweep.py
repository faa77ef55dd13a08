"""Next-round experiment: does a second resident CTA per SM hide the per-piece latency of the
multi-RHS DMMA kernels for narrow passes?  Shared memory of those kernels is sized by the operand rows (ld = mu + pad), so with
`ctas = 2 x SMs` phase 1 fits twice per SM for mu <= 16 with the default packets and phase 2 once its packets are cut to 40 KB.
Times the row-major product for a few widths under each option set, and the single-RHS product (whose stream shares
`stage_bytes`) beside it.   usage: python scripts/mm_occupancy_sweep.py [N] [mu,mu,...]"""
import os
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402
from htool_python_b200 import B200HMatrix  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
mus = [int(v) for v in (sys.argv[2] if len(sys.argv) > 2 else "1,4,8,16,64").split(",")]
args = types.SimpleNamespace(n=n, eps=1e-3, eta=10.0, leaf=64, kernel="laplace", wavenumber=0.0)
torch.cuda.set_device(0)
tree, H, flat, bt = bench.assemble(args, 1, 0, lambda m: print("[mm-sweep]", m, file=sys.stderr, flush=True))
sms = torch.cuda.get_device_properties(0).multi_processor_count

VARIANTS_ALL = [
    ("default (1 CTA / SM)", {}),
    ("ctas = 2 x SMs", dict(ctas=2 * sms)),
    ("ctas = 2 x SMs, 40 KB packets", dict(ctas=2 * sms, stage_bytes=40960)),
    ("40 KB packets only", dict(stage_bytes=40960)),
    ("EXPERIMENTAL DFMA consumer (mm_variant bit 2)", dict(mm_variant=1 + 4)),
    ("EXPERIMENTAL DFMA consumer, ctas = 2 x SMs", dict(mm_variant=1 + 4, ctas=2 * sms)),
]

sel = os.environ.get("MM_SWEEP_VARIANTS")
VARIANTS = [VARIANTS_ALL[int(i)] for i in sel.split(",")] if sel else VARIANTS_ALL


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


for label, opts in VARIANTS:
    hb = B200HMatrix(flat, device=0, **opts)
    row = []
    for mu in mus:
        X = torch.rand(n * mu, dtype=torch.float64, device="cuda")
        Y = torch.zeros(n * mu, dtype=torch.float64, device="cuda")
        if mu == 1:
            ms = timed(lambda: hb.engine.add_vector_product("N", 1.0, X, 0.0, Y))
        else:
            ms = timed(lambda: hb.engine.add_matrix_product_row_major("N", 1.0, X, 0.0, Y, mu))
        row.append(f"mu={mu}: {ms:7.3f} ms")
        del X, Y
    print(f"mm-sweep: {label:32s} " + " | ".join(row), flush=True)
    hb.close()
    torch.cuda.empty_cache()
