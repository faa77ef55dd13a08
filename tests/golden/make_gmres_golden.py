"""Generates tests/golden/gmres_golden.npz by running the REFERENCE's own Krylov solver -- HPDDM::IterativeMethod::solve
(GMRES, lib/hpddm/include/HPDDM_GMRES.hpp:30-196, the call of DDM::solve, lib/htool/include/htool/solvers/ddm.hpp:193),
compiled from /root/reference into oracle/_ref/libhpddm_ref.so (oracle/hpddm_ref.cpp, one rank, no preconditioner) --
on the committed golden H-matrices of this directory (assembled by the reference, see make_golden.py); the operator is
the flat restatement of the reference product (oracle/flat_oracle.c, itself pinned against the reference).

Stored per case: right-hand side, HPDDM's solution, HPDDM's iteration count and number of operator calls, the options.
The reference cannot travel to the GPU box; these vectors are what the device solver is compared with there.

Run:  python tests/golden/make_gmres_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers  # noqa: E402
from oracle import flat_oracle as fo  # noqa: E402
from oracle import helpers_ref as hr  # noqa: E402

CASES = [  # (fixture, tol, restart, max_it, orthogonalization)
    ("d_SL_cylinder", 1e-6, 30, 200, "cgs"),
    ("d_SU_sphere", 1e-8, 200, 200, "cgs"),
    ("d_SU_sphere", 1e-6, 10, 25, "cgs"),    # stops at max_it, not converged
    ("d_SL_cylinder", 1e-5, 40, 200, "mgs"),
    ("z_N_sphere", 1e-10, 40, 100, "cgs"),
    ("z_HL_cylinder", 1e-12, 10, 100, "cgs"),
]

out = {}
for k, (name, tol, restart, max_it, ortho) in enumerate(CASES):
    flat, _ = helpers.load_golden(name)
    rng = np.random.default_rng(1000 + k)
    b = helpers.rand_vec(rng, flat.n_rows, flat.np_dtype)
    args = f"-hpddm_krylov_method gmres -hpddm_tol {tol} -hpddm_gmres_restart {restart} -hpddm_max_it {max_it} -hpddm_variant right -hpddm_orthogonalization {ortho}"
    x, it, calls = hr.hpddm_solve(lambda v: fo.matvec_user(flat, v), b, args=args)
    res = np.linalg.norm(fo.matvec_user(flat, x) - b) / np.linalg.norm(b)
    print(f"{name:16s} {args}\n    iterations={it} operator_calls={calls} true residual={res:.3e}")
    out[f"{k}_fixture"] = np.array(name)
    out[f"{k}_args"] = np.array(args)
    out[f"{k}_b"] = b
    out[f"{k}_x"] = x
    out[f"{k}_iterations"] = np.array(it)
    out[f"{k}_calls"] = np.array(calls)
out["n_cases"] = np.array(len(CASES))
np.savez_compressed(os.path.join(HERE, "gmres_golden.npz"), **out)
