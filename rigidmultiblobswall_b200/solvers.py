"""Device-resident Krylov solvers around the blob mobility product.

gmres    -- restarted, right-preconditioned GMRES; same contract as the
            reference's general_application_utils.gmres (:514-635), which wraps
            scipy.sparse.linalg.gmres with A' = A P^-1 and returns x = P^-1 y.
lanczos  -- stochastic_forcing.stochastic_forcing_lanczos
            (stochastic_forcing/stochastic_forcing.py:112-264): factor * M^{1/2} z.
stochastic_forcing_eig / _eig_symm / _cholesky -- the reference's dense variants (:7-110) on torch tensors.

The Krylov basis, all matvecs and all dot products stay on the GPU; only the
(restart+1)-sized Hessenberg / tridiagonal data are handled on the host.  With
torch.distributed initialised and `group` given, vectors are rank-local shards
and every inner product is all-reduced (NCCL).
"""
import numpy as np
import torch


def _dots(V, w, group):
    h = V @ w
    if group is not None:
        torch.distributed.all_reduce(h, group=group)
    return h


def _norm(w, group):
    if group is None:
        return torch.linalg.norm(w)
    s = torch.dot(w, w).reshape(1)
    torch.distributed.all_reduce(s, group=group)
    return torch.sqrt(s[0])


def gmres(A, b, x0=None, tol=1e-5, restart=60, maxiter=None, M=None, callback=None, atol=0.0, group=None):
    """Solve A x = b.  A, M: callables on device vectors (M applies the RIGHT preconditioner P^-1).
    Returns (x, info, iterations); info = 0 on convergence, else the iteration count (SciPy's convention).
    Convergence: ||b - A x|| <= max(atol, tol*||b||), tested on the Givens residual estimate every
    inner iteration like scipy.sparse.linalg.gmres; callback(residual_norm/||b||) once per inner iteration.
    maxiter counts restart cycles (SciPy >= 1.12 semantics).  x0: initial guess for x (not for the preconditioned
    variable); a good guess saves iterations because the target stays max(atol, tol*||b||)."""
    n = b.numel()
    dev, dt = b.device, b.dtype
    if maxiter is None:
        maxiter = max(1, 10 * n)
    restart = min(restart, n) if group is None else restart
    bnorm = float(_norm(b, group))
    if bnorm == 0.0:
        return torch.zeros_like(b), 0, 0
    target = max(float(atol), tol * bnorm)
    apply_AM = (lambda v: A(M(v))) if M is not None else A
    if x0 is not None:
        # Initial guess, in the unpreconditioned variable: solve A dx = b - A x0 to the SAME absolute target, add.
        # (The reference's right-preconditioned wrapper drops x0, general_application_utils.py:627.)
        r0 = b - A(x0)
        scale = float(_norm(r0, group)) / bnorm
        cb = None if callback is None else (lambda rel: callback(rel * scale))     # residuals stay relative to ||b||
        dx, info, its = gmres(A, r0, tol=0.0, restart=restart, maxiter=maxiter, M=M, callback=cb, atol=target,
                              group=group)
        return x0 + dx, info, its
    y = torch.zeros_like(b)                    # solution of A P^-1 y = b
    V = torch.empty(restart + 1, n, dtype=dt, device=dev)
    its = 0
    r = b.clone()
    beta = bnorm
    for _cycle in range(maxiter):
        if beta <= target:
            break
        V[0] = r / beta
        H = np.zeros((restart + 1, restart))
        cs, sn = np.zeros(restart), np.zeros(restart)
        g = np.zeros(restart + 1)
        g[0] = beta
        k_done = 0
        res = beta
        for j in range(restart):
            w = apply_AM(V[j])
            # classical Gram-Schmidt, twice (two GEMV pairs instead of j dependent dot products)
            h = _dots(V[:j + 1], w, group)
            w = w - V[:j + 1].T @ h
            h2 = _dots(V[:j + 1], w, group)
            w = w - V[:j + 1].T @ h2
            hn = _norm(w, group)
            col = torch.cat([h + h2, hn.reshape(1)]).cpu().numpy()      # the one host sync of the iteration
            V[j + 1] = w / hn if col[-1] > 0 else w
            for i in range(j):                                          # previous Givens rotations
                t = cs[i] * col[i] + sn[i] * col[i + 1]
                col[i + 1] = -sn[i] * col[i] + cs[i] * col[i + 1]
                col[i] = t
            d = np.hypot(col[j], col[j + 1])
            cs[j], sn[j] = (col[j] / d, col[j + 1] / d) if d > 0 else (1.0, 0.0)
            col[j], col[j + 1] = d, 0.0
            H[:j + 2, j] = col
            g[j + 1] = -sn[j] * g[j]
            g[j] = cs[j] * g[j]
            its += 1
            k_done = j + 1
            res = abs(g[j + 1])
            if callback is not None:
                callback(res / bnorm)
            if res <= target:
                break
        coef = np.linalg.solve(np.triu(H[:k_done, :k_done]), g[:k_done]) if k_done else np.zeros(0)
        y = y + V[:k_done].T @ torch.from_numpy(coef).to(dev)
        if k_done and res <= target:
            # converged on the Givens residual estimate: like scipy.sparse.linalg.gmres, no extra operator application
            # to re-measure it (one product in twelve at configs 2 and 5); with two passes of Gram-Schmidt the
            # estimate tracks the true residual to rounding (tests assert the true residual)
            beta = res
            break
        r = b - apply_AM(y)
        beta = float(_norm(r, group))
    x = M(y) if M is not None else y
    info = 0 if beta <= target else its
    return x, info, its


def lanczos(factor=1.0, tolerance=1e-6, max_iter=1000, dim=None, mobility_mult=None, L_mult=None, z=None,
            print_residual=False, group=None, generator=None):
    """factor * M^{1/2} z by Lanczos (Ando et al. 2012), same arguments / return as
    stochastic_forcing.stochastic_forcing_lanczos (stochastic_forcing.py:112): (noise, iterations).
    Differences in HOW: the basis lives on the device, full reorthogonalisation is two GEMVs, and the
    noise vector is only assembled from the coefficient vector c_i = V^T-coordinates when needed."""
    if factor == 0.0:
        return torch.zeros(dim if dim is not None else z.numel(), dtype=torch.float64,
                           device=z.device if z is not None else "cuda"), 0
    if z is None:
        z = torch.randn(dim, dtype=torch.float64, device="cuda", generator=generator)
    dim = z.numel()
    dev = z.device
    cap = min(max_iter + 2, 64)
    V = torch.empty(cap, dim, dtype=torch.float64, device=dev)
    v_norm = float(_norm(z, group))
    V[0] = z.reshape(-1) / v_norm
    h_diag, h_sup = [], []
    coef_old = None
    for i in range(max_iter + 1):
        if i + 2 > V.shape[0]:                                   # grow the basis storage geometrically
            Vn = torch.empty(min(2 * V.shape[0], max_iter + 2), dim, dtype=torch.float64, device=dev)
            Vn[:V.shape[0]] = V
            V = Vn
        w = mobility_mult(V[i]).reshape(-1)
        if i > 0:
            w = w - h_sup[i - 1] * V[i - 1]
        hd = torch.dot(w, V[i]).reshape(1)
        if group is not None:
            torch.distributed.all_reduce(hd, group=group)
        w = w - hd * V[i]
        hs = _norm(w, group)
        d_s = torch.cat([hd, hs.reshape(1)]).cpu().numpy()        # one host sync per iteration
        h_diag.append(float(d_s[0]))
        h_sup.append(float(d_s[1]))
        if h_sup[i] > 0:
            w = w / h_sup[i]
        else:
            w = w.clone()
            w[0] = 1.0
        k = len(h_diag)
        h = np.diag(h_diag) + np.diag(h_sup[:k - 1], -1) + np.diag(h_sup[:k - 1], 1)
        eig_values, eig_vectors = np.linalg.eigh(h)
        sq = np.sqrt(np.maximum(eig_values, 0.0))
        coef = eig_vectors @ (v_norm * factor * sq * eig_vectors[0, :])   # noise = V^T coef (Ando eq. 16)
        # full re-orthogonalisation against the basis (two passes of classical Gram-Schmidt)
        for _ in range(2):
            w = w - V[:k].T @ _dots(V[:k], w, group)
        V[i + 1] = w
        if i > 0:
            # ||noise - noise_old|| and ||noise_old|| in the orthonormal basis: plain coefficient norms
            old = np.concatenate([coef_old, [0.0]])
            diff_norm = np.linalg.norm(coef - old)
            old_norm = np.linalg.norm(coef_old)
            if print_residual:
                if i == 1:
                    print('lanczos =  0 1')
                print('lanczos = ', i, diff_norm / old_norm)
            if diff_norm / max(old_norm, np.finfo(float).eps) < tolerance:
                noise = V[:k].T @ torch.from_numpy(coef).to(dev)
                return (noise if L_mult is None else L_mult(noise).reshape(-1)), i
        coef_old = coef
    noise = V[:len(coef_old)].T @ torch.from_numpy(coef_old).to(dev)
    return (noise if L_mult is None else L_mult(noise).reshape(-1)), max_iter


# ---- dense M^{1/2} z (stochastic_forcing/stochastic_forcing.py:7-110): for the small systems whose dense mobility
# ---- mobility.single_wall_fluid_mobility_b200 / rotne_prager_tensor_b200 assemble; same names, arguments and
# ---- conventions as the reference, on torch tensors (plain library calls: eigh / cholesky)
def _noise_vector(mobility, z, generator):
    M = torch.as_tensor(mobility, dtype=torch.float64)
    if z is None:
        z = torch.randn(M.shape[0], dtype=torch.float64, device=M.device, generator=generator)
    return M, torch.as_tensor(z, dtype=torch.float64).to(M.device).reshape(-1)


def stochastic_forcing_eig(mobility, factor=1.0, z=None, generator=None):
    """factor * V S^{1/2} z with M = V S V^T; negative eigenvalues count as 0 (stochastic_forcing.py:7-37)."""
    M, z = _noise_vector(mobility, z, generator)
    w, V = torch.linalg.eigh(M)
    return factor * (V @ (torch.sqrt(torch.clamp(w, min=0.0)) * z))


def stochastic_forcing_eig_symm(mobility, factor=1.0, z=None, generator=None):
    """factor * V S^{1/2} V^T z, the symmetric square root (stochastic_forcing.py:40-73)."""
    M, z = _noise_vector(mobility, z, generator)
    w, V = torch.linalg.eigh(M)
    return factor * (V @ (torch.sqrt(torch.clamp(w, min=0.0)) * (V.T @ z)))


def stochastic_forcing_cholesky(mobility, factor=1.0, z=None, generator=None):
    """factor * L z with M = L L^T (stochastic_forcing.py:77-110); M must be positive definite."""
    M, z = _noise_vector(mobility, z, generator)
    return factor * (torch.linalg.cholesky(M) @ z)
