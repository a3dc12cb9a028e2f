"""CPU oracle for the per-knot collocation evaluation (TEST INFRASTRUCTURE ONLY).

This file restates, with numpy/scipy, the mathematics of the hot path that
QuantumCollocation.jl delegates to DirectTrajOpt.jl's ``BilinearIntegrator``:

    defect      delta_k = x_{k+1} - exp(dt_k * G(u_k)) x_k     (docs/src/index.md:99-105,
                                                                 src/quantum_integrators.jl:28-45)
    generator   G(u) = G_drift + sum_j u_j G_j                   (src/quantum_integrators.jl:31,43)
    Jacobian    d delta / d(x_k, u_k, dt_k, x_{k+1})
    Hessian     d^2 (mu_k . delta_k) / d z_k^2
    fidelities  src/quantum_objectives.jl:21-27, 88-103, 173-191
    constraints src/quantum_constraints.jl:20-35, 65-100, 106-121

PARITY STATUS.  The arithmetic of this path lives in un-vendored Julia
dependencies (DirectTrajOpt.jl 0.8, ExponentialAction.jl 0.2,
PiccoloQuantumObjects.jl 0.10 -- compat bounds in Project.toml:21-34) and no
Julia toolchain exists in this image, so the reference evaluator cannot be run
here.  What the reference tree does hold are OUTPUTS of that evaluator: solved
trajectories (docs/data/*.jld2) and the Ipopt summaries of the runs that made
them (docs/literate/examples/multilevel_transmon.jl:139-141,241-243;
two_qubit_gates.jl:261-263).  tests/test_reference_fixtures.py replays them:

  PINNED against the reference's own output
    * defect: max|delta| on three solved trajectories equals the logged
      "Constraint violation" to 6e-12, 2e-12 and 0 (last digit) relative;
    * objective terms: (subspace) unitary infidelity, quadratic regularisers,
      minimum-time term and leakage objective reproduce the three logged
      "Objective" values to <= 1e-14;
    * iso-vec / knot layout (test/test_utils.jl:52-103), CoherentKet known answers
      (src/quantum_objectives.jl:315-394), structure cardinalities of the Ipopt logs
      (multilevel_transmon.jl:105-113, two_qubit_gates.jl:135-143).
  PARITY UNPINNED (no reference output exists in the tree to compare with)
    * Jacobian and Hessian-of-Lagrangian VALUES and the DirectTrajOpt 0.8 COO
      ORDER.  They are derivatives of the pinned defect function and are checked
      by central differences (the reference's own ``test_integrator`` criterion,
      src/quantum_integrators.jl:194) and against scipy's expm_frechet; the order
      is taken from the reference at run time (qcb_set_structure).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import
this module.  The product path (the CUDA library) never does.

The oracle is deliberately *independent* of the algorithm the CUDA kernels use
(truncated Taylor + scaling-and-squaring with adjoint second-order sweeps): it
uses scipy's Pade expm, scipy's expm_frechet, and block-triangular
exponentials for the second-order Frechet terms.
"""
from __future__ import annotations

import numpy as np
from scipy.linalg import expm, expm_frechet

# --------------------------------------------------------------------------- #
# Isomorphisms (PiccoloQuantumObjects conventions as used by the reference:
# src/quantum_objectives.jl:25,177; test/test_utils.jl:88-92)
# --------------------------------------------------------------------------- #


def iso(Z: np.ndarray) -> np.ndarray:
    """Real representation of a complex matrix: [[Re, -Im], [Im, Re]]."""
    Z = np.asarray(Z, dtype=complex)
    return np.block([[Z.real, -Z.imag], [Z.imag, Z.real]])


def G_of_H(H: np.ndarray) -> np.ndarray:
    """Generator G(H) = iso(-iH) = [[Im H, Re H], [-Re H, Im H]] (skew for Hermitian H)."""
    H = np.asarray(H, dtype=complex)
    return np.block([[H.imag, H.real], [-H.real, H.imag]])


def ket_to_iso(psi: np.ndarray) -> np.ndarray:
    psi = np.asarray(psi, dtype=complex)
    return np.concatenate([psi.real, psi.imag])


def iso_to_ket(x: np.ndarray) -> np.ndarray:
    n = x.shape[0] // 2
    return x[:n] + 1j * x[n:]


def operator_to_iso_vec(U: np.ndarray) -> np.ndarray:
    """vec([Re U; Im U]): per column j the 2n numbers [Re U[:,j]; Im U[:,j]]."""
    U = np.asarray(U, dtype=complex)
    return np.concatenate([np.concatenate([U[:, j].real, U[:, j].imag]) for j in range(U.shape[1])])


def iso_vec_to_operator(x: np.ndarray) -> np.ndarray:
    n = int(round(np.sqrt(x.shape[0] / 2)))
    P = x.reshape(n, 2 * n).T  # column-major reshape to (2n, n)
    return P[:n, :] + 1j * P[n:, :]


# --------------------------------------------------------------------------- #
# Per-unit evaluation (one knot k of one integrator)
# --------------------------------------------------------------------------- #


def second_frechet(A: np.ndarray, B: np.ndarray, C: np.ndarray) -> np.ndarray:
    """D^2 exp(A)[B, C] via exp of a 3x3 block upper-triangular matrix (both orderings)."""
    d = A.shape[0]
    Zr = np.zeros((d, d))

    def ordered(P, Q):
        T = np.block([[A, P, Zr], [Zr, A, Q], [Zr, Zr, A]])
        return expm(T)[:d, 2 * d:]

    return ordered(B, C) + ordered(C, B)


def unit_eval(G0, Gd, X, Xn, M, u, dt, want_hessian=True):
    """All canonical blocks for one unit.

    G0 (d,d), Gd (m,d,d); X, Xn, M are (d,c) panels (column-major reshape of the
    x_dim-vectors); u (m,), dt scalar.  Signs are those of delta = Xn - E X.
    Returns dict: E, delta (d,c), J_u (m,d,c), J_dt (d,c), and (if want_hessian)
    H_xu (m,d,c), H_xdt (d,c), H_uu (m,m), H_udt (m,), H_dtdt (scalar).
    """
    m = Gd.shape[0]
    G = G0 + np.tensordot(u, Gd, axes=1) if m else G0.copy()
    A = dt * G
    E = expm(A)
    out = {"E": E, "delta": Xn - E @ X}
    L = [expm_frechet(A, dt * Gd[j], compute_expm=False) for j in range(m)]
    GE = G @ E
    out["J_u"] = np.array([-(L[j] @ X) for j in range(m)]).reshape(m, *X.shape)
    out["J_dt"] = -(GE @ X)
    if want_hessian:
        W = M @ X.T
        out["H_xu"] = np.array([-(L[j].T @ M) for j in range(m)]).reshape(m, *X.shape)
        out["H_xdt"] = -(GE.T @ M)
        Huu = np.zeros((m, m))
        for i in range(m):
            for j in range(i, m):
                Q = second_frechet(A, dt * Gd[i], dt * Gd[j])
                Huu[i, j] = Huu[j, i] = -np.sum(W * Q)
        out["H_uu"] = Huu
        out["H_udt"] = np.array([-np.sum(W * (Gd[j] @ E + G @ L[j])) for j in range(m)])
        out["H_dtdt"] = -np.sum(W * (G @ GE))
    return out


# --------------------------------------------------------------------------- #
# Problem layout + native sparsity structure
# --------------------------------------------------------------------------- #


class Layout:
    """Knot-major flat Z layout for one (possibly sampled) bilinear integrator set.

    d: generator dim, c: state columns (x_dim = d*c), m: drives, N: knots,
    S: systems; x_off[s], u_off, dt_off: 0-based offsets inside a knot of width z_dim.
    jac_mode: 0 = dense x-block (x_dim^2 per knot, the structure the docs-era Ipopt
    logs imply), 1 = block-diagonal (c blocks of d^2).
    """

    def __init__(self, d, c, m, N, S, z_dim, x_off, u_off, dt_off, jac_mode=0):
        self.d, self.c, self.m, self.N, self.S = d, c, m, N, S
        self.z_dim, self.x_off, self.u_off, self.dt_off = z_dim, list(x_off), u_off, dt_off
        self.jac_mode = jac_mode
        self.x_dim = d * c
        self.K = N - 1

    @property
    def jac_nnz_unit(self):
        xb = self.x_dim ** 2 if self.jac_mode == 0 else self.c * self.d ** 2
        return xb + self.x_dim * (self.m + 2)

    @property
    def hess_nnz_unit(self):
        m = self.m
        return self.x_dim * (m + 1) + m * (m + 1) // 2 + m + 1

    def unit_inputs(self, Z, mu, s, k):
        d, c, zd = self.d, self.c, self.z_dim
        zk = Z[k * zd:(k + 1) * zd]
        zn = Z[(k + 1) * zd:(k + 2) * zd]
        X = zk[self.x_off[s]:self.x_off[s] + self.x_dim].reshape(c, d).T
        Xn = zn[self.x_off[s]:self.x_off[s] + self.x_dim].reshape(c, d).T
        u = zk[self.u_off:self.u_off + self.m]
        dt = zk[self.dt_off]
        M = None
        if mu is not None:
            M = mu[(s * self.K + k) * self.x_dim:(s * self.K + k + 1) * self.x_dim].reshape(c, d).T
        return X, Xn, M, u, dt

    # ---- native structure: per unit, entries sorted by (column, row) like Julia's findnz ----
    def jac_unit_pattern(self, s):
        """List of (local_row, local_col_in_[z_k;z_{k+1}], source) sorted by (col,row).

        source = ('E', r, q) -> -E[r,q]; ('Z',) -> 0; ('U', j, i) -> J_u[j] flat i;
        ('T', i) -> J_dt flat i; ('I',) -> +1.
        """
        d, c, m, xd, zd = self.d, self.c, self.m, self.x_dim, self.z_dim
        ent = []
        for q in range(xd):
            b, qq = divmod(q, d)
            col = self.x_off[s] + q
            if self.jac_mode == 0:
                for r in range(xd):
                    a, rr = divmod(r, d)
                    ent.append((r, col, ("E", rr, qq) if a == b else ("Z",)))
            else:
                for rr in range(d):
                    ent.append((b * d + rr, col, ("E", rr, qq)))
        for j in range(m):
            for r in range(xd):
                ent.append((r, self.u_off + j, ("U", j, r)))
        for r in range(xd):
            ent.append((r, self.dt_off, ("T", r)))
        for r in range(xd):
            ent.append((r, zd + self.x_off[s] + r, ("I",)))
        ent.sort(key=lambda e: (e[1], e[0]))
        return ent

    def hess_unit_pattern(self, s):
        """Upper-triangular (row<=col in z_k indices) entries sorted by (col,row).

        source = ('XU', j, i), ('XT', i), ('UU', i, j), ('UT', j), ('TT',).
        """
        m, xd = self.m, self.x_dim
        ent = []

        def add(a, b, src):
            ent.append((min(a, b), max(a, b), src))

        for j in range(m):
            for i in range(xd):
                add(self.x_off[s] + i, self.u_off + j, ("XU", j, i))
        for i in range(xd):
            add(self.x_off[s] + i, self.dt_off, ("XT", i))
        for i in range(m):
            for j in range(i, m):
                add(self.u_off + i, self.u_off + j, ("UU", i, j))
        for j in range(m):
            add(self.u_off + j, self.dt_off, ("UT", j))
        add(self.dt_off, self.dt_off, ("TT",))
        ent.sort(key=lambda e: (e[1], e[0]))
        return ent

    def jacobian_structure(self, row_base=0):
        """1-based (rows, cols) of the whole integrator set, order [s][k][unit pattern]."""
        rows, cols = [], []
        for s in range(self.S):
            pat = self.jac_unit_pattern(s)
            lr = np.array([e[0] for e in pat], dtype=np.int64)
            lc = np.array([e[1] for e in pat], dtype=np.int64)
            for k in range(self.K):
                rows.append(row_base + (s * self.K + k) * self.x_dim + lr + 1)
                cols.append(k * self.z_dim + lc + 1)
        return np.concatenate(rows), np.concatenate(cols)

    def hessian_structure(self):
        rows, cols = [], []
        for s in range(self.S):
            pat = self.hess_unit_pattern(s)
            lr = np.array([e[0] for e in pat], dtype=np.int64)
            lc = np.array([e[1] for e in pat], dtype=np.int64)
            for k in range(self.K):
                rows.append(k * self.z_dim + lr + 1)
                cols.append(k * self.z_dim + lc + 1)
        return np.concatenate(rows), np.concatenate(cols)


def _jac_value(src, ev):
    t = src[0]
    if t == "E":
        return -ev["E"][src[1], src[2]]
    if t == "Z":
        return 0.0
    if t == "U":
        return ev["J_u"][src[1]].T.reshape(-1)[src[2]]
    if t == "T":
        return ev["J_dt"].T.reshape(-1)[src[1]]
    return 1.0


def _hess_value(src, ev):
    t = src[0]
    if t == "XU":
        return ev["H_xu"][src[1]].T.reshape(-1)[src[2]]
    if t == "XT":
        return ev["H_xdt"].T.reshape(-1)[src[1]]
    if t == "UU":
        return ev["H_uu"][src[1], src[2]]
    if t == "UT":
        return ev["H_udt"][src[1]]
    return ev["H_dtdt"]


def assemble_unit(layout: Layout, s: int, ev: dict):
    """COO value vectors (delta, jac, hess) of ONE unit of system s in the native order, from the
    dictionary of blocks `unit_eval` returns (or one loaded from a golden file).  hess is None when
    the dictionary holds no second-order blocks."""
    delta = np.asarray(ev["delta"]).T.reshape(-1)
    jac = np.array([_jac_value(e[2], ev) for e in layout.jac_unit_pattern(s)])
    hess = None
    if "H_uu" in ev:
        hess = np.array([_hess_value(e[2], ev) for e in layout.hess_unit_pattern(s)])
    return delta, jac, hess


def evaluate(layout: Layout, G0, Gd, Z, mu=None, want_jac=True, want_hess=True,
             k_range=None):
    """Whole-trajectory oracle evaluation in the native structure order.

    G0: (S,d,d), Gd: (S,m,d,d).  Returns (delta, jac_vals, hess_vals); delta is
    ordered [s][k][x_dim] (src/problem_templates/smooth_pulse_problem.jl:513 --
    ``integrator.dim`` = x_dim*(N-1) per integrator).
    """
    L = layout
    ks = range(L.K) if k_range is None else range(*k_range)
    nk = len(ks)
    delta = np.zeros(L.S * nk * L.x_dim)
    jac = np.zeros(L.S * nk * L.jac_nnz_unit) if want_jac else None
    hess = np.zeros(L.S * nk * L.hess_nnz_unit) if (want_hess and mu is not None) else None
    for s in range(L.S):
        jp = L.jac_unit_pattern(s) if want_jac else None
        hp = L.hess_unit_pattern(s) if hess is not None else None
        for ik, k in enumerate(ks):
            X, Xn, M, u, dt = L.unit_inputs(Z, mu, s, k)
            ev = unit_eval(G0[s], Gd[s], X, Xn, M, u, dt, want_hessian=hess is not None)
            o = (s * nk + ik)
            delta[o * L.x_dim:(o + 1) * L.x_dim] = ev["delta"].T.reshape(-1)
            if want_jac:
                jac[o * L.jac_nnz_unit:(o + 1) * L.jac_nnz_unit] = [_jac_value(e[2], ev) for e in jp]
            if hess is not None:
                hess[o * L.hess_nnz_unit:(o + 1) * L.hess_nnz_unit] = [_hess_value(e[2], ev) for e in hp]
    return delta, jac, hess


def defect_only(layout: Layout, G0, Gd, Z):
    L = layout
    delta = np.zeros(L.S * L.K * L.x_dim)
    for s in range(L.S):
        for k in range(L.K):
            X, Xn, _, u, dt = L.unit_inputs(Z, None, s, k)
            G = G0[s] + np.tensordot(u, Gd[s], axes=1)
            o = s * L.K + k
            delta[o * L.x_dim:(o + 1) * L.x_dim] = (Xn - expm(dt * G) @ X).T.reshape(-1)
    return delta


# --------------------------------------------------------------------------- #
# Terminal fidelities (src/quantum_objectives.jl) -- value, gradient, Hessian
# --------------------------------------------------------------------------- #

KIND_UNITARY = 0        # unitary_fidelity_loss, src/quantum_objectives.jl:173-180
KIND_KET = 1            # ket_fidelity_loss, :21-27
KIND_COHERENT_KET = 2   # coherent_ket_fidelity, :88-103
KIND_UNITARY_SUBSPACE = 3  # EmbeddedOperator variant, :182-191


def unitary_fidelity(x, U_goal):
    U = iso_vec_to_operator(x)
    n = U.shape[0]
    return abs(np.trace(U_goal.conj().T @ U)) ** 2 / n ** 2


def unitary_subspace_fidelity(x, U_goal_sub, subspace):
    U = iso_vec_to_operator(x)[np.ix_(subspace, subspace)]
    n = len(subspace)
    Mm = U_goal_sub.conj().T @ U
    return (abs(np.trace(Mm.conj().T @ Mm)) + abs(np.trace(Mm)) ** 2) / (n * (n + 1))


def ket_fidelity(x, psi_goal):
    return abs(np.vdot(psi_goal, iso_to_ket(x))) ** 2


def coherent_ket_fidelity(xs, goals):
    n = len(xs)
    return abs(sum(np.vdot(goals[i], iso_to_ket(xs[i])) for i in range(n)) / n) ** 2


def fidelity_quadratic_form(kind, n_or_dim, goal, subspace=None):
    """Every fidelity above is F(x) = sum_r (g_r . x)^2 * w for fixed real vectors g_r.

    Returns (Gmat [R, x_dim], w).  unitary/ket/coherent: R = 2 (a, b); subspace: many rows.
    """
    if kind == KIND_UNITARY:
        U_goal = np.asarray(goal, dtype=complex)
        n = U_goal.shape[0]
        ga = operator_to_iso_vec(U_goal)                       # Re tr(Ug^† U)
        gb = operator_to_iso_vec(1j * U_goal)                  # Im tr(Ug^† U) = Re tr((iUg)^† U)
        return np.array([ga, gb]), 1.0 / n ** 2
    if kind == KIND_KET:
        psi = np.asarray(goal, dtype=complex)
        return np.array([ket_to_iso(psi), ket_to_iso(1j * psi)]), 1.0
    if kind == KIND_COHERENT_KET:
        goals = [np.asarray(g, dtype=complex) for g in goal]
        n = len(goals)
        ga = np.concatenate([ket_to_iso(g) for g in goals])
        gb = np.concatenate([ket_to_iso(1j * g) for g in goals])
        return np.array([ga, gb]), 1.0 / n ** 2
    if kind == KIND_UNITARY_SUBSPACE:
        Ug = np.asarray(goal, dtype=complex)
        n_full = n_or_dim
        ns = len(subspace)
        rows = []
        # |tr M|^2 with M = Ug^† U_sub: two rows; tr(M^†M) = ||U_sub||_F^2 for unitary Ug:
        # in general ||Ug^† U_sub||_F^2 = sum over entries (Re^2 + Im^2) of M.
        def embed(Vsub):
            V = np.zeros((n_full, n_full), dtype=complex)
            V[np.ix_(subspace, subspace)] = Vsub
            return operator_to_iso_vec(V)
        rows.append(embed(Ug))
        rows.append(embed(1j * Ug))
        # M[a,b] = sum_r conj(Ug[r,a]) U_sub[r,b]  -> Re and Im parts as linear forms
        for a in range(ns):
            for b in range(ns):
                V = np.zeros((ns, ns), dtype=complex)
                V[:, b] = Ug[:, a]
                rows.append(embed(V))
                rows.append(embed(1j * V))
        return np.array(rows), 1.0 / (ns * (ns + 1))
    raise ValueError(kind)


def fidelity_vgh(kind, x, goal, n_full=None, subspace=None):
    """F, grad F, Hess F (dense) from the quadratic form."""
    Gm, w = fidelity_quadratic_form(kind, n_full, goal, subspace)
    p = Gm @ x
    F = w * float(p @ p)
    g = 2 * w * (Gm.T @ p)
    H = 2 * w * (Gm.T @ Gm)
    return F, g, H


# --------------------------------------------------------------------------- #
# Finite-difference self-consistency (the reference's test_integrator criterion,
# src/quantum_integrators.jl:194)
# --------------------------------------------------------------------------- #


def fd_check_unit(G0, Gd, X, Xn, M, u, dt, h=1e-6):
    """max abs error of analytic Jacobian / Hessian blocks vs central differences."""
    ev = unit_eval(G0, Gd, X, Xn, M, u, dt)
    m = len(u)

    def defect(uu, tt, XX):
        G = G0 + np.tensordot(uu, Gd, axes=1)
        return Xn - expm(tt * G) @ XX

    errs = {}
    ju = []
    for j in range(m):
        e = np.zeros(m); e[j] = h
        ju.append((defect(u + e, dt, X) - defect(u - e, dt, X)) / (2 * h))
    errs["J_u"] = max((np.abs(np.array(ju) - ev["J_u"]).max(),)) if m else 0.0
    jt = (defect(u, dt + h, X) - defect(u, dt - h, X)) / (2 * h)
    errs["J_dt"] = np.abs(jt - ev["J_dt"]).max()

    def lag(uu, tt):
        return np.sum(M * defect(uu, tt, X))

    hh = 1e-4
    Huu = np.zeros((m, m))
    for i in range(m):
        for j in range(m):
            ei = np.zeros(m); ei[i] = hh
            ej = np.zeros(m); ej[j] = hh
            Huu[i, j] = (lag(u + ei + ej, dt) - lag(u + ei - ej, dt) - lag(u - ei + ej, dt) + lag(u - ei - ej, dt)) / (4 * hh * hh)
    errs["H_uu"] = np.abs(Huu - ev["H_uu"]).max() if m else 0.0
    Hut = np.zeros(m)
    for j in range(m):
        ej = np.zeros(m); ej[j] = hh
        Hut[j] = (lag(u + ej, dt + hh) - lag(u + ej, dt - hh) - lag(u - ej, dt + hh) + lag(u - ej, dt - hh)) / (4 * hh * hh)
    errs["H_udt"] = np.abs(Hut - ev["H_udt"]).max() if m else 0.0
    Htt = (lag(u, dt + hh) - 2 * lag(u, dt) + lag(u, dt - hh)) / (hh * hh)
    errs["H_dtdt"] = abs(Htt - ev["H_dtdt"])
    # d/dX of (mu . J_u[j]) = H_xu[j]
    return errs


# --------------------------------------------------------------------------- #
# Auxiliary knot-point terms
# --------------------------------------------------------------------------- #


def derivative_integrator(Z, N, z_dim, dt_off, pairs, mu=None):
    """delta_k = x_{k+1} - x_k - dt_k dx_k for every (x_off, dx_off, dim) pair
    (src/problem_templates/smooth_pulse_problem.jl:150-152, spline_pulse_problem.jl:176).
    Returns delta [K, rows], jac [K, rows, 4] in the order (x_k, dx_k, dt_k, x_{k+1}), hess [K, rows]."""
    Zm = np.asarray(Z).reshape(N, z_dim)
    rx = np.concatenate([np.arange(p[0], p[0] + p[2]) for p in pairs])
    rdx = np.concatenate([np.arange(p[1], p[1] + p[2]) for p in pairs])
    dt = Zm[:-1, dt_off][:, None]
    delta = Zm[1:, rx] - Zm[:-1, rx] - dt * Zm[:-1, rdx]
    jac = np.stack([-np.ones_like(delta), -dt * np.ones_like(delta), -Zm[:-1, rdx], np.ones_like(delta)], axis=-1)
    hess = None if mu is None else -np.asarray(mu).reshape(delta.shape)
    return delta, jac, hess


def leakage(Z, N, z_dim, x_off, idx, Qs=None, value=0.0, mu=None):
    """Objective sum_t Q_t sum(abs2, x_t[idx]) / len(idx) (src/quantum_objectives.jl:298) and constraint
    abs2.(x_t[idx]) .- value (src/quantum_constraints.jl:163) with their derivatives."""
    Zm = np.asarray(Z).reshape(N, z_dim)
    v = Zm[:, x_off + np.asarray(idx)]
    Q = np.ones(N) if Qs is None else np.asarray(Qs)
    obj = float(np.sum(Q * np.sum(v * v, axis=1) / len(idx)))
    grad = 2 * Q[:, None] * v / len(idx)
    con = v * v - value
    cjac = 2 * v
    chess = None if mu is None else 2 * np.asarray(mu).reshape(v.shape)
    return obj, grad, con, cjac, chess


def quadratic_regularizer(Z, N, z_dim, dt_off, terms, min_time_D=0.0):
    """Objective terms around the integrator, in the form the reference's solver logs pin
    (tests/test_reference_fixtures.py; call sites src/problem_templates/smooth_pulse_problem.jl:139-152,
    minimum_time_problem.jl:84-131):
        J = sum_terms 1/2 sum_{t=1..N} sum_i R_i (dt_t v_{t,i})^2 + D sum_{t=1..N-1} dt_t
    terms: [(offset, dim, R)].  Returns (J, dense gradient [N, z_dim], dense Hessian [N, z_dim, z_dim])."""
    Zm = np.asarray(Z, dtype=np.float64).reshape(N, z_dim)
    dt = Zm[:, dt_off]
    J = float(min_time_D * dt[:-1].sum())
    grad = np.zeros((N, z_dim))
    hess = np.zeros((N, z_dim, z_dim))
    grad[:-1, dt_off] += min_time_D
    for off, dim, R in terms:
        R = np.broadcast_to(np.asarray(R, dtype=np.float64), (dim,))
        for i in range(dim):
            c = off + i
            v = Zm[:, c]
            J += float(0.5 * np.sum(R[i] * (dt * v) ** 2))
            grad[:, c] += R[i] * dt ** 2 * v
            grad[:, dt_off] += R[i] * dt * v ** 2
            hess[:, c, c] += R[i] * dt ** 2
            hess[:, c, dt_off] += 2 * R[i] * dt * v
            hess[:, dt_off, c] += 2 * R[i] * dt * v
            hess[:, dt_off, dt_off] += R[i] * v ** 2
    return J, grad, hess


def timesteps_all_equal(Z, N, z_dim, dt_off):
    """TimeStepsAllEqualConstraint ("timesteps all equal constraint" in the reference's logs, 2 nnz per
    row): rows dt_t - dt_N for t < N."""
    Zm = np.asarray(Z, dtype=np.float64).reshape(N, z_dim)
    return Zm[:-1, dt_off] - Zm[-1, dt_off]


def leakage_indices(subspace, levels):
    """0-based iso-vec indices of the operator entries that connect the subspace with its complement
    (row outside & column inside, and row inside & column outside; real and imaginary parts).  This
    set reproduces the logged objective of the reference's leakage example to 1e-18
    (docs/literate/examples/multilevel_transmon.jl:241, tests/test_reference_fixtures.py)."""
    n = levels
    sub = set(int(s) for s in subspace)
    idx = []
    for j in range(n):
        for part in (0, n):
            for i in range(n):
                if (i in sub) != (j in sub):
                    idx.append(j * 2 * n + part + i)
    return np.asarray(idx, dtype=np.int64)


# --------------------------------------------------------------------------- #
# open-system and variational generators, free-phase, sensitivity and density
# losses.  Restated from src/quantum_integrators.jl:52-56,126-169 and
# src/quantum_objectives.jl:203-279, src/quantum_constraints.jl:123-144.  The iso-vec of a density
# matrix ([vec(Re rho); vec(Im rho)]), the Lindbladian `sys.𝒢` and `Isomorphisms.var_G` live in
# PiccoloQuantumObjects 0.10 (not in the tree): their forms here are restated from memory of that
# package [memory] -- parity for them is unpinned, like every value-level item of this path.
# --------------------------------------------------------------------------- #


def density_to_iso_vec(rho):
    rho = np.asarray(rho, dtype=complex)
    return np.concatenate([rho.real.T.reshape(-1), rho.imag.T.reshape(-1)])     # column-major vec


def iso_vec_to_density(x):
    n = int(round(np.sqrt(len(x) / 2)))
    return (x[:n * n] + 1j * x[n * n:]).reshape(n, n).T


def lindblad_generators(H_drift, H_drives, dissipators=()):
    """(G_drift, [G_j]) of d/dt rho_iso = (G_drift + sum_j u_j G_j) rho_iso for
    rho' = -i[H, rho] + sum_L (L rho L^† - 1/2 {L^†L, rho}), vec column-major:
    vec(A rho B) = (B^T kron A) vec(rho); iso(M) = [[Re M, -Im M], [Im M, Re M]]."""
    n = np.asarray(H_drift).shape[0]
    I = np.eye(n)

    def ad(H):
        H = np.asarray(H, dtype=complex)
        return -1j * (np.kron(I, H) - np.kron(H.T, I))
    D = np.zeros((n * n, n * n), dtype=complex)
    for Lop in dissipators:
        Lop = np.asarray(Lop, dtype=complex)
        LdL = Lop.conj().T @ Lop
        D += np.kron(Lop.conj(), Lop) - 0.5 * np.kron(I, LdL) - 0.5 * np.kron(LdL.T, I)
    return iso(ad(H_drift) + D), [iso(ad(h)) for h in H_drives]


def var_G(G, G_vars):
    """Isomorphisms.var_G: block lower-triangular generator of [x; dx/da_1; ...]:
    d/dt x_v = G x_v + G_v x  (src/quantum_integrators.jl:149,166)."""
    d, nv = G.shape[0], len(G_vars)
    out = np.zeros(((1 + nv) * d, (1 + nv) * d))
    for b in range(1 + nv):
        out[b * d:(b + 1) * d, b * d:(b + 1) * d] = G
    for v, Gv in enumerate(G_vars):
        out[(v + 1) * d:(v + 2) * d, :d] = Gv
    return out


def density_pure_state_fidelity(x, psi):
    """src/quantum_objectives.jl:232-239: real(psi' * rho * psi)."""
    psi = np.asarray(psi, dtype=complex)
    return float(np.real(psi.conj() @ iso_vec_to_density(x) @ psi))


def free_phase_goal(U_goal, phase_map, theta, side=0):
    D = np.diag(np.exp(1j * (np.asarray(phase_map) @ np.asarray(theta))))
    return D @ U_goal if side == 0 else U_goal @ D


def free_phase_fidelity(z, U_goal, phase_map, side=0):
    """l(z) of src/quantum_objectives.jl:211-214 without the |1 - .|: z = [x; theta],
    unitary_fidelity_loss(x, U_goal(theta)) with U_goal(theta) = diag(e^{i P theta}) U_goal."""
    nt = np.asarray(phase_map).shape[1]
    x, theta = (z[:-nt], z[-nt:]) if nt else (z, np.zeros(0))
    return unitary_fidelity(x, free_phase_goal(U_goal, phase_map, theta, side))


def free_phase_vgh(z, U_goal, phase_map, side=0):
    """F, gradient and dense Hessian over z = [x; theta], analytically: with A(theta) the 2 x x_dim
    matrix of the linear forms (Re, Im) of tr(U_goal(theta)' U), F = |A x|^2 / n^2; the theta
    derivatives of A come from d/dtheta_k of the goal, -i P_k-weighted."""
    U_goal = np.asarray(U_goal, dtype=complex)
    P = np.asarray(phase_map, dtype=float)
    n, nt = U_goal.shape[0], P.shape[1]
    x, theta = (z[:-nt], z[-nt:]) if nt else (z, np.zeros(0))
    xd = len(x)

    def forms(Ug):    # rows (Re, Im) of tr(Ug' U) as linear forms of x
        return np.array([operator_to_iso_vec(Ug), operator_to_iso_vec(1j * Ug)])
    Ug = free_phase_goal(U_goal, P, theta, side)
    dUg, d2Ug = [], [[None] * nt for _ in range(nt)]
    for k in range(nt):
        Dk = np.diag(1j * P[:, k])
        dUg.append(Dk @ Ug if side == 0 else Ug @ Dk)
        for l in range(nt):
            Dkl = np.diag(-(P[:, k] * P[:, l]))
            d2Ug[k][l] = Dkl @ Ug if side == 0 else Ug @ Dkl
    A = forms(Ug)
    s = A @ x                                       # (Re S, Im S)
    w = 1.0 / n ** 2
    dA = [forms(D) for D in dUg]
    ds = np.array([a @ x for a in dA]).reshape(nt, 2)
    g = np.concatenate([2 * w * (A.T @ s), 2 * w * (ds @ s)])
    H = np.zeros((xd + nt, xd + nt))
    H[:xd, :xd] = 2 * w * (A.T @ A)
    for k in range(nt):
        H[:xd, xd + k] = H[xd + k, :xd] = 2 * w * (A.T @ ds[k] + dA[k].T @ s)
        for l in range(nt):
            H[xd + k, xd + l] = 2 * w * (ds[k] @ ds[l] + s @ (forms(d2Ug[k][l]) @ x))
    return w * float(s @ s), g, H


def unitary_self_overlap(x):
    """unitary_fidelity_loss(Ũ⃗) of src/quantum_objectives.jl:255-261: |tr(U'U)|^2 / n^2."""
    U = iso_vec_to_operator(x)
    return abs(np.trace(U.conj().T @ U)) ** 2 / U.shape[0] ** 2


def sensitivity_objective(Z, N, z_dim, x_off, x_dim, times, Qs=None, scale=1.0):
    """UnitarySensitivityObjective (:263-279): sum_t Q_t scale^4 |tr(U_t'U_t)|^2/n^2, times 1-based.
    Returns (J, grad [len(times), x_dim], hess [len(times), x_dim, x_dim])."""
    Zk = np.asarray(Z).reshape(N, z_dim)
    n = int(round(np.sqrt(x_dim / 2)))
    Qs = np.ones(len(times)) if Qs is None else np.asarray(Qs, dtype=float)
    J, g, Hs = 0.0, [], []
    for q, t in zip(Qs, times):
        x = Zk[t - 1, x_off:x_off + x_dim]
        s = float(x @ x)
        J += q * scale ** 4 * unitary_self_overlap(x)
        g.append(q * scale ** 4 * 4 * s * x / n ** 2)
        Hs.append(q * scale ** 4 * (4 * s * np.eye(x_dim) + 8 * np.outer(x, x)) / n ** 2)
    return J, np.array(g), np.array(Hs)
