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
