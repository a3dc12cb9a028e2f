"""Real-isomorphic representations (PiccoloQuantumObjects conventions used by the reference:
src/quantum_objectives.jl:25,177,187; layout fixture test/test_utils.jl:52-103)."""
from __future__ import annotations

import numpy as np


def G_of_H(H) -> np.ndarray:
    """iso(-iH) = [[Im H, Re H], [-Re H, Im H]]; real skew-symmetric for Hermitian H."""
    H = np.asarray(H, dtype=complex)
    return np.block([[H.imag, H.real], [-H.real, H.imag]])


def ket_to_iso(psi) -> np.ndarray:
    psi = np.asarray(psi, dtype=complex)
    return np.concatenate([psi.real, psi.imag])


def iso_to_ket(x) -> np.ndarray:
    n = len(x) // 2
    return np.asarray(x[:n]) + 1j * np.asarray(x[n:])


def operator_to_iso_vec(U) -> np.ndarray:
    U = np.asarray(U, dtype=complex)
    return np.vstack([U.real, U.imag]).T.reshape(-1)


def iso_vec_to_operator(x) -> np.ndarray:
    x = np.asarray(x)
    n = int(round(np.sqrt(len(x) / 2)))
    P = x.reshape(n, 2 * n).T
    return P[:n] + 1j * P[n:]


def iso(M) -> np.ndarray:
    """Real representation [[Re, -Im], [Im, Re]] of a complex matrix."""
    M = np.asarray(M, dtype=complex)
    return np.block([[M.real, -M.imag], [M.imag, M.real]])


def density_to_iso_vec(rho) -> np.ndarray:
    """[vec(Re rho); vec(Im rho)], column-major vec (PiccoloQuantumObjects' density iso-vec [memory])."""
    rho = np.asarray(rho, dtype=complex)
    return np.concatenate([rho.real.T.reshape(-1), rho.imag.T.reshape(-1)])


def lindblad_generators(H_drift, H_drives, dissipators=()):
    """(G_drift, G_drives [m, 2n^2, 2n^2]) of the Lindbladian on the density iso-vec -- what
    `sys.𝒢` of an OpenQuantumSystem evaluates to (src/quantum_integrators.jl:52-56,126-134):
    rho' = -i[H(u), rho] + sum_L (L rho L' - 1/2 {L'L, rho}).  Dense and not skew-symmetric."""
    H_drift = np.asarray(H_drift, dtype=complex)
    n = H_drift.shape[0]
    I = np.eye(n)
    ad = lambda H: -1j * (np.kron(I, H) - np.kron(np.asarray(H, dtype=complex).T, I))
    D = np.zeros((n * n, n * n), dtype=complex)
    for L in dissipators:
        L = np.asarray(L, dtype=complex)
        LdL = L.conj().T @ L
        D += np.kron(L.conj(), L) - 0.5 * np.kron(I, LdL) - 0.5 * np.kron(LdL.T, I)
    return iso(ad(H_drift) + D), np.array([iso(ad(h)) for h in H_drives]).reshape(len(H_drives), 2 * n * n, 2 * n * n)


def var_G(G, G_vars) -> np.ndarray:
    """Block generator of [x; variations...]: diagonal blocks G, first block column G_v
    (`Isomorphisms.var_G` as used at src/quantum_integrators.jl:149,166 [memory])."""
    G = np.asarray(G, dtype=float)
    d, nv = G.shape[0], len(G_vars)
    out = np.zeros(((1 + nv) * d, (1 + nv) * d))
    for b in range(1 + nv):
        out[b * d:(b + 1) * d, b * d:(b + 1) * d] = G
    for v, Gv in enumerate(G_vars):
        out[(v + 1) * d:(v + 2) * d, :d] = Gv
    return out
