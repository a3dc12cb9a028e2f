"""GPU tests of the auxiliary knot-point terms (derivative integrators, leakage) against the oracle."""
import numpy as np
import pytest

import qcb200
from qcb200 import Aux, piccolo, workloads
import qc_oracle as O
from conftest import rel_err

pytestmark = pytest.mark.gpu


def test_derivative_integrators_match_oracle_and_finite_differences(built_libs):
    w = workloads.make("cfg2", N=9)                 # smooth-pulse layout: x, t, dt, u, du, ddu
    L = w.layout
    m = L.m
    pairs = [(L.u_off, L.u_off + m, m), (L.u_off + m, L.u_off + 2 * m, m)]     # (u, du), (du, ddu)
    aux = Aux(L.N, L.z_dim, L.dt_off, derivatives=pairs)
    rng = np.random.default_rng(0)
    mu = rng.normal(size=aux.derivative_dim)
    d1, j1, h1 = aux.derivative_eval(w.Z, mu)
    d0, j0, h0 = O.derivative_integrator(w.Z, L.N, L.z_dim, L.dt_off, pairs, mu)
    assert np.array_equal(d1, d0.reshape(-1)) and np.array_equal(j1, j0.reshape(-1)) and np.array_equal(h1, h0.reshape(-1))
    (jr, jc), (hr, hc) = aux.derivative_structure()
    assert len(jr) == 4 * aux.derivative_dim and aux.derivative_dim == (L.N - 1) * 2 * m   # 4 nnz per row
    assert np.all(hr <= hc)
    # Jacobian vs central differences of the defect
    h = 1e-6
    for v in rng.choice(np.unique(jc) - 1, size=12, replace=False):
        Zp, Zm = w.Z.copy(), w.Z.copy()
        Zp[v] += h; Zm[v] -= h
        fd = (aux.derivative_eval(Zp)[0] - aux.derivative_eval(Zm)[0]) / (2 * h)
        col = np.zeros(aux.derivative_dim)
        sel = jc == v + 1
        np.add.at(col, jr[sel] - 1, j1[sel])
        assert np.abs(fd - col).max() < 1e-8


def test_leakage_objective_and_constraint(built_libs):
    w = workloads.make("cfg2", N=7)
    L = w.layout
    # leakage indices: every iso-vec entry that touches level 2 of the 3-level transmon
    U = np.arange(18).reshape(3, 6).T             # column-major (6 x 3) index map of the iso-vec
    idx = sorted(set(U[[2, 5], :].reshape(-1)) | set(U[:, 2]))
    rng = np.random.default_rng(1)
    Qs = rng.uniform(0.5, 2.0, L.N)
    aux = Aux(L.N, L.z_dim, L.dt_off, leak_x_off=L.x_off[0], leak_idx=idx)
    mu = rng.normal(size=aux.leakage_dim)
    o1 = aux.leakage_eval(w.Z, Qs, 1e-2, mu)
    o0 = O.leakage(w.Z, L.N, L.z_dim, L.x_off[0], idx, Qs, 1e-2, mu)
    assert abs(o1[0] - o0[0]) < 1e-13 * max(1.0, abs(o0[0]))
    for a, b in zip(o1[1:], o0[1:]):
        assert rel_err(a, b.reshape(-1)) < 1e-14
    # reference definitions through the mirror objects
    traj = piccolo.NamedTrajectory({"U_iso": w.Z.reshape(L.N, L.z_dim)[:, :18].T, "t": np.zeros((1, L.N)),
                                    "dt": np.full((1, L.N), 0.1), "u": np.zeros((2, L.N))}, timestep="dt")
    obj = piccolo.LeakageObjective(idx, "U_iso", traj)
    x = traj["U_iso"]
    assert abs(obj(traj) - np.sum(x[idx, :] ** 2) / len(idx)) < 1e-12     # sum(abs2, x[indices]) / length(indices), Q = 1
    con = piccolo.LeakageConstraint(1e-2, idx, "U_iso", traj)
    assert rel_err(con.values(traj), (x[idx, :] ** 2 - 1e-2).T.reshape(-1)) < 1e-14
    g = obj.gradient(np.zeros(traj.dim * traj.N), traj)
    assert np.count_nonzero(g) <= len(idx) * L.N and np.abs(g).max() > 0
    di = piccolo.DerivativeIntegrator("u", "u", traj)     # degenerate but legal: x = dx = u
    assert di.dim == (L.N - 1) * 2
