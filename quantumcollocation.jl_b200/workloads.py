"""Synthetic workloads of BASELINE.json's configs, seed-deterministic.

Each workload is (layout, G_drift [S,d,d], G_drives [S,m,d,d], Z, mu, goal info).  States are
Haar-random per knot (the evaluator's cost does not depend on the values), controls
0.1*randn clipped, dt constant +-10 % jitter, multipliers randn.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

import numpy as np

from .isomorphisms import G_of_H, density_to_iso_vec, ket_to_iso, lindblad_generators, operator_to_iso_vec
from .layout import JAC_BLOCK_DIAG, JAC_DENSE_X, KnotLayout

X2 = np.array([[0, 1], [1, 0]], dtype=complex)
Y2 = np.array([[0, -1j], [1j, 0]], dtype=complex)
Z2 = np.array([[1, 0], [0, -1]], dtype=complex)
I2 = np.eye(2, dtype=complex)
H_GATE = np.array([[1, 1], [1, -1]], dtype=complex) / np.sqrt(2)


def annihilate(levels: int) -> np.ndarray:
    return np.diag(np.sqrt(np.arange(1, levels)), 1).astype(complex)


def kron_all(ops):
    out = np.eye(1, dtype=complex)
    for o in ops:
        out = np.kron(out, o)
    return out


def embed(op, site, nsites, levels=2):
    ops = [np.eye(levels, dtype=complex)] * nsites
    ops = list(ops)
    ops[site] = op
    return kron_all(ops)


def haar_unitary(n, rng):
    A = rng.normal(size=(n, n)) + 1j * rng.normal(size=(n, n))
    Q, R = np.linalg.qr(A)
    return Q * (np.diag(R) / np.abs(np.diag(R)))


@dataclass
class Workload:
    name: str
    layout: KnotLayout
    G_drift: np.ndarray       # [S, d, d]
    G_drives: np.ndarray      # [S, m, d, d]
    Z: np.ndarray             # [N * z_dim]
    mu: np.ndarray            # [S * K * x_dim]
    U_goal: Optional[np.ndarray] = None
    kind: str = "unitary"     # or "ket"

    @property
    def units(self) -> int:
        return self.layout.S * self.layout.K


def _systems(name):
    """Returns (list of (H_drift, [H_drives])) per system, dt, N, goal, columns-kind."""
    if name == "cfg1":      # README example: qubit, X/Y drives, T=50, dt=0.2, Hadamard
        return [(0.0 * Z2, [X2, Y2])], 0.2, 50, H_GATE, "unitary"
    if name == "cfg1_drift":  # the in-source tests' system: 0.1 Z drift (test/test_utils.jl:109)
        return [(0.1 * Z2, [X2, Y2])], 0.2, 50, H_GATE, "unitary"
    if name == "cfg2":      # 3-level transmon, delta = 0.2 (docs/literate/examples/multilevel_transmon.jl:6,43)
        a = annihilate(3)
        n = a.conj().T @ a
        Hd = -0.2 / 2 * n @ (n - np.eye(3))
        goal = np.eye(3, dtype=complex)
        goal[:2, :2] = X2
        return [(Hd, [a + a.conj().T, 1j * (a - a.conj().T)])], 10.0 / 200, 200, goal, "unitary"
    if name in ("cfg3", "cfg3_ket"):   # two transmon qutrits, CZ
        a = annihilate(3)
        n = a.conj().T @ a
        I3 = np.eye(3)
        a1, a2 = np.kron(a, I3), np.kron(I3, a)
        Hd = sum(-0.2 / 2 * (np.kron(n @ (n - I3), I3), np.kron(I3, n @ (n - I3)))[i] for i in range(2))
        Hd = Hd + 0.05 * (a1.conj().T @ a2 + a2.conj().T @ a1)
        drives = [a1 + a1.conj().T, 1j * (a1 - a1.conj().T), a2 + a2.conj().T, 1j * (a2 - a2.conj().T)]
        goal = np.eye(9, dtype=complex)
        goal[4, 4] = -1  # |11> -> -|11> in the qubit subspace (levels {0,1} of each qutrit)
        return [(2 * np.pi * Hd, [2 * np.pi * h for h in drives])], 40.0 / 500, 500, goal, \
            ("unitary" if name == "cfg3" else "ket")
    if name == "cfg4":      # robust qubit, 256 detuning/amplitude-error samples, T=400
        rng = np.random.default_rng(1234)
        eps = rng.uniform(-0.15, 0.15, 256)     # docs/literate/examples/robust_control.jl:170-171
        eta = rng.uniform(-0.05, 0.05, 256)
        return [((1 + e) * 0.1 * Z2, [(1 + h) * X2, (1 + h) * Y2]) for e, h in zip(eps, eta)], 0.2, 400, H_GATE, "unitary"
    if name == "cfg5":      # six coupled 2-level transmons, 12 drives, T=2000
        nq = 6
        rng = np.random.default_rng(5)
        omega = 0.05 + 0.02 * rng.normal(size=nq)   # rotating-frame detunings, GHz
        J = 0.005
        Hd = sum(omega[i] * embed(Z2, i, nq) / 2 for i in range(nq))
        for i in range(nq - 1):
            Hd = Hd + J * (embed(X2, i, nq) @ embed(X2, i + 1, nq) + embed(Y2, i, nq) @ embed(Y2, i + 1, nq)) / 2
        drives = []
        for i in range(nq):
            drives += [embed(X2, i, nq) / 2, embed(Y2, i, nq) / 2]
        goal = haar_unitary(2 ** nq, np.random.default_rng(6))
        return [(2 * np.pi * Hd, [2 * np.pi * h for h in drives])], 200.0 / 2000, 2000, goal, "unitary"
    raise KeyError(name)


def make_lindblad(N: Optional[int] = None, seed: int = 0, u_scale: float = 0.1) -> Workload:
    """Open 3-level transmon (cfg2's Hamiltonian) with relaxation a and dephasing a'a: the
    DensityTrajectory / `sys.𝒢` path of src/quantum_integrators.jl:52-56 -- dense, non-skew
    18 x 18 generator on the density iso-vec, c = 1.  States: random density matrices."""
    a = annihilate(3)
    nop = a.conj().T @ a
    Hd = -0.2 / 2 * nop @ (nop - np.eye(3))
    drives = [a + a.conj().T, 1j * (a - a.conj().T)]
    diss = [np.sqrt(0.02) * a, np.sqrt(0.01) * nop]
    G0, Gd = lindblad_generators(Hd, drives, diss)
    N = N or 100
    n, m = 3, 2
    L = KnotLayout.smooth_pulse(2 * n * n, 1, m, N, 1, JAC_DENSE_X)
    rng = np.random.default_rng(seed)
    Z = np.zeros((N, L.z_dim))
    for k in range(N):
        A = rng.normal(size=(n, n)) + 1j * rng.normal(size=(n, n))
        rho = A @ A.conj().T
        Z[k, :L.x_dim] = density_to_iso_vec(rho / np.trace(rho).real)
    Z[:, L.u_off:L.u_off + m] = np.clip(u_scale * rng.normal(size=(N, m)), -1, 1)
    Z[:, L.u_off + m:] = 0.01 * rng.normal(size=(N, L.z_dim - L.u_off - m))
    Z[:, L.dt_off] = 0.1 * (1 + 0.1 * rng.uniform(-1, 1, N))
    Z[:, L.dt_off - 1] = np.concatenate([[0], np.cumsum(Z[:-1, L.dt_off])])
    mu = rng.normal(size=(N - 1) * L.x_dim)
    return Workload("lindblad3", L, G0[None], Gd[None], Z.reshape(-1), mu, None, "density")


def make(name: str, N: Optional[int] = None, S: Optional[int] = None, seed: int = 0,
         jac_mode: Optional[int] = None, u_scale: float = 0.1) -> Workload:
    if name == "lindblad3":
        return make_lindblad(N, seed, u_scale)
    systems, dt, N0, goal, kind = _systems(name)
    if S is not None:
        systems = systems[:S]
    N = N or N0
    S = len(systems)
    n = systems[0][0].shape[0]
    d = 2 * n
    m = len(systems[0][1])
    c = n if kind == "unitary" else 1
    if jac_mode is None:
        jac_mode = JAC_BLOCK_DIAG if d * c > 512 else JAC_DENSE_X
    L = KnotLayout.smooth_pulse(d, c, m, N, S, jac_mode, with_derivatives=(S == 1))
    G0 = np.array([G_of_H(h) for h, _ in systems])
    Gd = np.array([[G_of_H(hj) for hj in hs] for _, hs in systems])
    rng = np.random.default_rng(seed)
    Z = np.zeros((N, L.z_dim))
    for k in range(N):
        for s in range(S):
            if kind == "unitary":
                xs = operator_to_iso_vec(haar_unitary(n, rng))
            else:
                v = rng.normal(size=n) + 1j * rng.normal(size=n)
                xs = ket_to_iso(v / np.linalg.norm(v))
            Z[k, L.x_off[s]:L.x_off[s] + L.x_dim] = xs
    Z[:, L.u_off:L.u_off + m] = np.clip(u_scale * rng.normal(size=(N, m)), -1, 1)
    if L.z_dim > L.u_off + m:
        Z[:, L.u_off + m:] = 0.01 * rng.normal(size=(N, L.z_dim - L.u_off - m))
    Z[:, L.dt_off] = dt * (1 + 0.1 * rng.uniform(-1, 1, N))
    Z[:, L.dt_off - 1] = np.concatenate([[0], np.cumsum(Z[:-1, L.dt_off])])
    mu = rng.normal(size=S * (N - 1) * L.x_dim)
    return Workload(name, L, G0, Gd, Z.reshape(-1), mu, goal, kind)
