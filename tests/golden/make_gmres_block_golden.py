"""Generates tests/golden/gmres_block_golden.npz: the REFERENCE's HPDDM GMRES (lib/hpddm/include/HPDDM_GMRES.hpp:30-196,
compiled from /root/reference into oracle/_ref/libhpddm_ref.so) run on BLOCKS of mu > 1 right-hand sides -- the lockstep
recurrences HPDDMOperator::GMV serves with one row-major matrix product per iteration (wrapper_hpddm.hpp:102-145) -- on the
committed golden H-matrices; operator = the flat restatement of the reference product (oracle/flat_oracle.c).

Stored per case: the block of right-hand sides, HPDDM's solutions, its iteration count and its number of (block) operator calls.
Column 0 of every block is A * ones (converges early), the others are random, so the columns leave the iteration at
different steps.

Run:  python tests/golden/make_gmres_block_golden.py
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

CASES = [  # (fixture, mu, tol, restart, max_it, orthogonalization)
    ("d_SL_cylinder", 3, 1e-6, 30, 200, "cgs"),
    ("d_SU_sphere", 4, 1e-6, 10, 25, "cgs"),     # stops at max_it with columns still running
    ("d_SL_cylinder", 2, 1e-5, 40, 200, "mgs"),
    ("z_N_sphere", 2, 1e-8, 40, 100, "cgs"),
]

out = {}
for k, (name, mu, tol, restart, max_it, ortho) in enumerate(CASES):
    flat, _ = helpers.load_golden(name)
    rng = np.random.default_rng(2000 + k)
    B = np.asfortranarray(np.stack([helpers.rand_vec(rng, flat.n_rows, flat.np_dtype) for _ in range(mu)], axis=1))
    B[:, 0] = fo.matvec_user(flat, np.ones(flat.n_rows, flat.np_dtype))
    args = f"-hpddm_krylov_method gmres -hpddm_tol {tol} -hpddm_gmres_restart {restart} -hpddm_max_it {max_it} -hpddm_variant right -hpddm_orthogonalization {ortho}"
    X, it, calls = hr.hpddm_solve(lambda v: fo.matvec_user(flat, v), B, args=args, mu=mu)
    res = [np.linalg.norm(fo.matvec_user(flat, np.ascontiguousarray(X[:, c])) - B[:, c]) / np.linalg.norm(B[:, c]) for c in range(mu)]
    print(f"{name:16s} mu={mu} {args}\n    iterations={it} block operator calls={calls} true residuals={['%.2e' % r for r in res]}")
    out[f"{k}_fixture"] = np.array(name)
    out[f"{k}_args"] = np.array(args)
    out[f"{k}_B"] = B
    out[f"{k}_X"] = X
    out[f"{k}_iterations"] = np.array(it)
    out[f"{k}_calls"] = np.array(calls)
out["n_cases"] = np.array(len(CASES))
np.savez_compressed(os.path.join(HERE, "gmres_block_golden.npz"), **out)
