"""oracle/gmres_oracle.py -- TEST INFRASTRUCTURE ONLY (checker; never imported by htool_python_b200/).

Plain-numpy restatement of the Krylov solver the reference's DDM::solve runs
(lib/htool/include/htool/solvers/ddm.hpp:193 -> HPDDM::IterativeMethod::solve -> IterativeMethod::GMRES,
lib/hpddm/include/HPDDM_GMRES.hpp:30-196), for ONE right-hand side (`gmres`) and for a block of mu right-hand sides
advanced in lockstep as HPDDM does with mu > 1 (`gmres_block`), no preconditioner
(-hpddm_schwarz_method none, ddm.hpp:136-138: with M = I the left/right variants coincide), restarted GMRES(m) with the
classical (default) or modified Gram-Schmidt of IterativeMethod::orthogonalization (HPDDM_iterative.hpp:512-560) and the
Givens-rotation least-squares update of IterativeMethod::Arnoldi (HPDDM_iterative.hpp:722-768), solution update
IterativeMethod::updateSol / computeMin / addSol (HPDDM_iterative.hpp:276-330).

PARITY PINNED: tests/test_gmres_oracle.py checks this restatement against the reference's own HPDDM code compiled from
/root/reference (oracle/_ref/libhpddm_ref.so, oracle/hpddm_ref.cpp): same iteration counts, same number of operator calls,
iterates equal to ~1e-12, on real/complex dense systems (single and block right-hand sides) and on the reference's solver
fixture (lib/htool_generate_data_test/data/output_*/{matrix,rhs,sol}.bin).
"""
import numpy as np

HPDDM_EPS = 1.0e-12  # HPDDM_define.hpp:48

DEFAULTS = dict(tol=1.0e-6, max_it=100, gmres_restart=40, orthogonalization="cgs", variant="right")  # HPDDM_iterative.hpp:194-215


def parse_hpddm_args(args):
    """'-hpddm_tol 1e-6 -hpddm_max_it 200 ...' -> dict (the subset of HPDDM options that the GMRES path reads)."""
    opt = dict(DEFAULTS)
    toks = args.split()
    i = 0
    while i < len(toks):
        key = toks[i].lstrip("-")
        if key.startswith("hpddm_"):
            key = key[len("hpddm_"):]
        val = toks[i + 1] if i + 1 < len(toks) and not toks[i + 1].startswith("-hpddm") else None
        i += 1 if val is None else 2
        if key == "tol":
            opt["tol"] = float(val)
        elif key == "max_it":
            opt["max_it"] = int(val)
        elif key == "gmres_restart":
            opt["gmres_restart"] = int(val)
        elif key == "orthogonalization":
            opt["orthogonalization"] = val
        elif key == "variant":
            opt["variant"] = val
        else:
            opt[key] = val
    return opt


def _arnoldi(H, V, s, sn, i, m, ortho):
    """IterativeMethod::Arnoldi for mu = 1 (HPDDM_iterative.hpp:722-768).  H[k] is COLUMN k of the Hessenberg matrix
    (length m+1); after the call H[i][:i+1] is the rotated (triangular) column, H[i][i+1] the rotation cosine, sn[i] the sine."""
    w = V[i + 1]
    if ortho == "mgs":  # HPDDM_iterative.hpp:522-540
        for k in range(i + 1):
            H[i][k] = np.vdot(V[k], w)
            w -= H[i][k] * V[k]
    else:  # classical Gram-Schmidt: one gemv 'C', one gemv 'N' (HPDDM_iterative.hpp:541-552)
        H[i][:i + 1] = V[:i + 1].conj() @ w
        w -= H[i][:i + 1] @ V[:i + 1]
    sn[i] = np.vdot(w, w).real
    H[i][i + 1] = np.sqrt(sn[i])
    if i < m - 1:
        w /= H[i][i + 1]
    for k in range(i):  # previous rotations (HPDDM_iterative.hpp:748-754)
        gamma = np.conj(H[k][k + 1]) * H[i][k] + sn[k] * H[i][k + 1]
        H[i][k + 1] = -sn[k] * H[i][k] + H[k][k + 1] * H[i][k + 1]
        H[i][k] = gamma
    delta = np.sqrt(abs(H[i][i]) ** 2 + abs(H[i][i + 1]) ** 2)  # nrm2 of the two entries (:757)
    sn[i] = np.real(H[i][i + 1]) / delta
    H[i][i + 1] = H[i][i] / delta
    H[i][i] = delta
    s[i + 1] = -sn[i] * s[i]
    s[i] *= np.conj(H[i][i + 1])


def _update_sol(x, H, V, s, dim):
    """updateSol (HPDDM_iterative.hpp:276-281): back substitution on the rotated Hessenberg (computeMin :283-297),
    then x += V[:, :dim] y (addSol, right variant with the identity preconditioner :316-326)."""
    if dim == 0:
        return
    y = np.array(s[:dim])
    for r in range(dim - 1, -1, -1):  # trsv 'U','N','N'
        for c in range(r + 1, dim):
            y[r] -= H[c][r] * y[c]
        y[r] /= H[r][r]
    s[:dim] = y
    x += y @ V[:dim]


def gmres(matvec, b, x0=None, tol=1.0e-6, max_it=100, gmres_restart=40, orthogonalization="cgs", variant="right", **_ignored):
    """returns (x, iterations, info): `iterations` is what HPDDM returns (min(j, max_it), HPDDM_GMRES.hpp:196);
    info = dict(operator_calls=..., residuals=[|s_i| / ||b|| per iteration], converged=bool)."""
    b = np.asarray(b)
    dt = np.complex128 if np.iscomplexobj(b) else np.float64
    b = b.astype(dt)
    n = b.shape[0]
    x = np.zeros(n, dt) if x0 is None else np.array(x0, dtype=dt)
    m = min(gmres_restart, max_it)  # HPDDM_iterative.hpp:213
    H = [np.zeros(m + 1, dt) for _ in range(m)]
    V = np.zeros((m + 1, n), dt)
    s = np.zeros(m + 1, dt)
    sn = np.zeros(m)
    norm_b = np.sqrt(np.vdot(b, b).real)  # initializeNorm (HPDDM_iterative.hpp:477-492) + HPDDM_GMRES.hpp:83-86
    if norm_b < HPDDM_EPS:
        norm_b = 1.0
    calls, residuals = 0, []
    conv = -m  # hasConverged sentinel (HPDDM_GMRES.hpp:57)
    j = 1
    while j <= max_it:
        V[0] = b - matvec(x)  # HPDDM_GMRES.hpp:62-64
        calls += 1
        beta2 = np.vdot(V[0], V[0]).real
        if j == 1 and beta2 < np.finfo(np.float64).eps ** 2:  # :87-90 -> breakdown / zero residual
            j = 0
            conv = 0
            break
        if conv > 0:
            conv = 0
        s[0] = np.sqrt(beta2)
        V[0] /= s[0]
        i = 0
        while i < m and j <= max_it:
            V[i + 1] = matvec(V[i])  # right variant, identity preconditioner (:131-135)
            calls += 1
            _arnoldi(H, V, s, sn, i, m, orthogonalization)
            i += 1
            residuals.append(abs(s[i]) / norm_b)
            if conv == -m and ((tol > 0 and abs(s[i]) / norm_b <= tol) or (tol < 0 and abs(s[i]) <= -tol)):  # checkConvergence :88-91
                conv = i
            if conv != -m:
                i = 0
                break
            j += 1
        if j != max_it + 1 and i == m:  # restart (:154-161)
            _update_sol(x, H, V, s, abs(conv))
        else:
            break
    if j == max_it + 1 and m > 0:  # :165-168
        rem = max_it % m
        if conv < 0:
            conv = rem if rem > 0 else -conv
    if j != 0:
        _update_sol(x, H, V, s, abs(conv))
    return x, min(j, max_it), dict(operator_calls=calls, residuals=residuals, converged=(j <= max_it and j != 0) or j == 0)


def gmres_block(matmat, B, X0=None, tol=1.0e-6, max_it=100, gmres_restart=40, orthogonalization="cgs", variant="right", **_ignored):
    """IterativeMethod::GMRES for a block of mu right-hand sides (HPDDM_GMRES.hpp:30-196 with mu > 1): mu independent GMRES
    recurrences advanced in LOCKSTEP -- one operator call (GMV on the n x mu block) per iteration, one Hessenberg matrix,
    rotation set and `hasConverged` entry per column; the loop ends when no column is left at the sentinel (:147-150), and
    updateSol uses each column's own dimension |hasConverged[nu]| (computeMin / addSol, HPDDM_iterative.hpp:283-326).
    matmat(X) -> A X on (n, mu) arrays.  Returns (X, iterations, info); info["converged_at"] = the per-column hasConverged."""
    B = np.asarray(B)
    dt = np.complex128 if np.iscomplexobj(B) else np.float64
    B = B.astype(dt)
    n, mu = B.shape
    X = np.zeros((n, mu), dt) if X0 is None else np.array(X0, dtype=dt)
    m = min(gmres_restart, max_it)
    H = [[np.zeros(m + 1, dt) for _ in range(m)] for _ in range(mu)]
    V = np.zeros((mu, m + 1, n), dt)  # V[nu][i] = column nu of the i-th block of the basis
    s = np.zeros((mu, m + 1), dt)
    sn = np.zeros((mu, m))
    norm_b = np.sqrt(np.einsum("ij,ij->j", B.conj(), B).real)
    norm_b[norm_b < HPDDM_EPS] = 1.0
    calls = 0
    conv = [-m] * mu
    j = 1
    while j <= max_it:
        R = B - matmat(X)
        calls += 1
        beta2 = np.einsum("ij,ij->j", R.conj(), R).real
        if j == 1 and np.any(beta2 < np.finfo(np.float64).eps ** 2):  # :87-90, any column
            j = 0
            conv = [0] * mu
            break
        for nu in range(mu):
            if conv[nu] > 0:
                conv[nu] = 0
            s[nu, 0] = np.sqrt(beta2[nu])
            V[nu, 0] = R[:, nu] / s[nu, 0]
        i = 0
        while i < m and j <= max_it:
            W = matmat(np.ascontiguousarray(V[:, i, :].T))
            calls += 1
            for nu in range(mu):
                V[nu, i + 1] = W[:, nu]
                _arnoldi(H[nu], V[nu], s[nu], sn[nu], i, m, orthogonalization)
            i += 1
            for nu in range(mu):
                r = abs(s[nu, i])
                if conv[nu] == -m and ((tol > 0 and r / norm_b[nu] <= tol) or (tol < 0 and r <= -tol)):
                    conv[nu] = i
            if -m not in conv:
                i = 0
                break
            j += 1
        if j != max_it + 1 and i == m:
            for nu in range(mu):
                _update_sol(X[:, nu], H[nu], V[nu], s[nu], abs(conv[nu]))
        else:
            break
    if j == max_it + 1 and m > 0:
        rem = max_it % m
        conv = [(rem if rem > 0 else -c) if c < 0 else c for c in conv]
    if j != 0:
        for nu in range(mu):
            _update_sol(X[:, nu], H[nu], V[nu], s[nu], abs(conv[nu]))
    return X, min(j, max_it), dict(operator_calls=calls, converged_at=list(conv), converged=(j <= max_it))
