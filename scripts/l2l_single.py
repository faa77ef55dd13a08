"""single-rank GPU check of B200DistributedOperator.matvec_local_to_local_ ('N', 'T') against the oracle on a golden fixture"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import helpers  # noqa: E402
from htool_python_b200 import B200DistributedOperator  # noqa: E402
from oracle import flat_oracle as fo  # noqa: E402

flat, _ = helpers.load_golden("d_N_cylinder")
n = flat.n_rows
op = B200DistributedOperator(flat, [(0, n)], fused_allgather=False, build_transpose=1)
rng = np.random.default_rng(0)
for tr in ("N", "T"):
    x, y0 = rng.random(n), rng.random(n)
    y = torch.from_numpy(y0.copy()).to(op.device)
    op.matvec_local_to_local_(torch.from_numpy(x).to(op.device), y, tr, 1.25, -0.5)
    ref = y0.copy()
    fo.add_vector_product(flat, tr, 1.25, x, -0.5, ref)
    print(f"l2l {tr}: rel err {np.linalg.norm(y.cpu().numpy() - ref) / np.linalg.norm(ref):.2e}", flush=True)
op.close()
