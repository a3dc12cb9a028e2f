"""A small equality-constrained Newton (SQP) solver with a fixed log barrier for simple bounds -- test
infrastructure only.  It plays Ipopt's part in tests/test_gpu_problem.py: it sees a problem exclusively through the
five callbacks Ipopt uses (f, grad f, g, Jacobian values, Hessian-of-the-Lagrangian values, with their COO
structures), applies Ipopt's inertia correction (Hessian shifted until the KKT matrix has n positive and m negative
eigenvalues) and an l1 merit line search.  Dense linear algebra: meant for a few thousand variables."""
import numpy as np
from scipy.linalg import ldl


def _inertia(D):
    """(positive, negative, zero) eigenvalue counts of the block-diagonal factor of an LDL^T."""
    pos = neg = zero = 0
    i, n = 0, D.shape[0]
    while i < n:
        if i + 1 < n and D[i + 1, i] != 0.0:
            ev = np.linalg.eigvalsh(D[i:i + 2, i:i + 2])
            pos += int((ev > 0).sum()); neg += int((ev < 0).sum()); zero += int((ev == 0).sum())
            i += 2
        else:
            pos += D[i, i] > 0; neg += D[i, i] < 0; zero += D[i, i] == 0
            i += 1
    return pos, neg, zero


def solve(cb, Z0, lb, ub, max_iter=100, mu=1e-4, tol=1e-8, log=None):
    """cb: object with n, m, jac_structure() / hess_structure() (1-based rows, cols), f(Z, new_x), grad_f, g, jac_g,
    h(Z, sigma, lam, new_x).  lb == ub fixes a variable; infinite bounds are allowed.  Returns (Z, lam, history)."""
    n, m = cb.n, cb.m
    jr, jc = (a - 1 for a in cb.jac_structure())
    hr, hc = (a - 1 for a in cb.hess_structure())
    free = np.flatnonzero(ub > lb)
    pos_of = -np.ones(n, dtype=np.int64); pos_of[free] = np.arange(free.size)
    has_l, has_u = np.isfinite(lb[free]), np.isfinite(ub[free])
    Z = Z0.copy()
    Z[ub <= lb] = lb[ub <= lb]
    Z[free] = np.clip(Z[free], np.where(has_l, lb[free] + 1e-3 * (ub[free] - lb[free]).clip(max=1.0), -np.inf),
                      np.where(has_u, ub[free] - 1e-3 * (ub[free] - lb[free]).clip(max=1.0), np.inf))
    lam = np.zeros(m)
    nf = free.size
    jsel = pos_of[jc] >= 0
    hsel = (pos_of[hr] >= 0) & (pos_of[hc] >= 0)

    def barrier(z):
        zf = z[free]
        return -mu * (np.log(zf[has_l] - lb[free][has_l]).sum() + np.log(ub[free][has_u] - zf[has_u]).sum())

    def barrier_grad_hess(z):
        zf = z[free]
        g, h = np.zeros(nf), np.zeros(nf)
        sl, su = zf[has_l] - lb[free][has_l], ub[free][has_u] - zf[has_u]
        g[has_l] -= mu / sl; g[has_u] += mu / su
        h[has_l] += mu / sl ** 2; h[has_u] += mu / su ** 2
        return g, h

    hist = []
    nu = 10.0
    delta_last = 0.0
    step = 0.0
    for it in range(max_iter):
        f = cb.f(Z, True)
        gf = cb.grad_f(Z, False)[free]
        g = cb.g(Z, False)
        J = np.zeros((m, nf)); np.add.at(J, (jr[jsel], pos_of[jc[jsel]]), cb.jac_g(Z, False)[jsel])
        hv = cb.h(Z, 1.0, lam, False)
        H = np.zeros((nf, nf)); np.add.at(H, (pos_of[hr[hsel]], pos_of[hc[hsel]]), hv[hsel])
        H = H + np.triu(H, 1).T
        bg, bh = barrier_grad_hess(Z)
        grad_phi = gf + bg
        H[np.diag_indices(nf)] += bh
        kkt_res = max(np.abs(grad_phi + J.T @ lam).max(), np.abs(g).max() if m else 0.0)
        hist.append((it, f, float(np.abs(g).max()) if m else 0.0, float(kkt_res), step))
        if log:
            log("%3d f %.8e viol %.2e kkt %.2e step %.2e" % hist[-1])
        if kkt_res < tol:
            break
        # Newton step on the KKT system with inertia correction (Ipopt, section 3.1 of its implementation paper)
        delta = 0.0
        while True:
            Kmat = np.block([[H + delta * np.eye(nf), J.T], [J, -1e-9 * np.eye(m)]])
            Lf, Df, perm = ldl(Kmat, lower=True)
            p, q, z = _inertia(Df)
            if p == nf and q == m:
                break
            delta = max(1e-4, delta_last / 3) if delta == 0.0 else delta * 8.0
            if delta > 1e12:
                raise RuntimeError("inertia correction failed")
        delta_last = delta
        sol = np.linalg.solve(Kmat, -np.concatenate([grad_phi + J.T @ lam, g]))
        dz, dlam = sol[:nf], sol[nf:]
        # fraction to the boundary
        zf = Z[free]
        amax = 1.0
        neg, posd = dz < 0, dz > 0
        lo = neg & has_l
        if lo.any():
            amax = min(amax, (0.99 * (zf[lo] - lb[free][lo]) / -dz[lo]).min())
        hi = posd & has_u
        if hi.any():
            amax = min(amax, (0.99 * (ub[free][hi] - zf[hi]) / dz[hi]).min())
        # l1 merit line search
        lam_new = lam + dlam
        nu = max(nu, 2.0 * np.abs(lam_new).max() if m else nu)
        phi0 = f + barrier(Z) + nu * np.abs(g).sum()
        dphi = grad_phi @ dz - nu * np.abs(g).sum()
        a = amax
        for _ in range(30):
            Zt = Z.copy(); Zt[free] = zf + a * dz
            ft = cb.f(Zt, True)
            gt = cb.g(Zt, False)
            if np.isfinite(ft) and ft + barrier(Zt) + nu * np.abs(gt).sum() <= phi0 + 1e-4 * a * min(dphi, 0.0):
                break
            a *= 0.5
        Z = Zt
        lam = lam + a * dlam
        step = a
    return Z, lam, hist
