"""Knot-major trajectory layout consumed by the C ABI ("Iso-vec layout").

Flat ``Z`` is knot-major with ``z_dim`` doubles per knot; inside a knot the component order is
whatever the caller's ``NamedTrajectory`` says (reference: the flat datavec has length
``traj.dim * traj.N``, src/quantum_objectives.jl:353) -- offsets are therefore run-time data.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Optional

JAC_DENSE_X = 0   # x_k block of the Jacobian stored as a dense x_dim^2 block (docs-era structure)
JAC_BLOCK_DIAG = 1  # x_k block stored as c diagonal blocks of d^2 (only feasible choice for d=128)


@dataclass
class KnotLayout:
    d: int                 # real-iso dimension of the generator
    c: int                 # columns of the state panel (n for a unitary, 1 for a ket); x_dim = d*c
    m: int                 # number of drives
    N: int                 # number of knots (K = N-1 intervals)
    S: int = 1             # number of systems sharing the controls (SamplingTrajectory)
    z_dim: int = 0
    x_off: List[int] = field(default_factory=list)
    u_off: int = 0
    dt_off: int = 0
    jac_mode: int = JAC_DENSE_X
    # optional state index map (include/qcb200.h, qcb_desc.x_index / row_index): the state is gathered
    # from several named components -- [S][x_dim] knot offsets of the panel entries, and the position
    # of every panel entry among the unit's defect rows (None: contiguous state at x_off, identity rows)
    x_index: Optional[List[List[int]]] = None
    row_index: Optional[List[int]] = None

    @property
    def x_dim(self) -> int:
        return self.d * self.c

    @property
    def K(self) -> int:
        return self.N - 1

    @property
    def jac_nnz_unit(self) -> int:
        xb = self.x_dim ** 2 if self.jac_mode == JAC_DENSE_X else self.c * self.d ** 2
        return xb + self.x_dim * (self.m + 2)

    @property
    def hess_nnz_unit(self) -> int:
        m = self.m
        return self.x_dim * (m + 1) + m * (m + 1) // 2 + m + 1

    @staticmethod
    def smooth_pulse(d, c, m, N, S=1, jac_mode=JAC_DENSE_X, with_derivatives=True) -> "KnotLayout":
        """Component order the reference's templates produce: states, t, dt, u [, du, ddu]
        (src/quantum_control_problem.jl:344-346, src/problem_templates/smooth_pulse_problem.jl:95-100)."""
        x_dim = d * c
        x_off = [s * x_dim for s in range(S)]
        t_off = S * x_dim
        dt_off = t_off + 1
        u_off = dt_off + 1
        z_dim = u_off + m * (3 if with_derivatives else 1)
        return KnotLayout(d, c, m, N, S, z_dim, x_off, u_off, dt_off, jac_mode)
