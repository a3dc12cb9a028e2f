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


def test_regularizer_min_time_and_equal_timesteps(built_libs):
    """Quadratic regularisers + minimum-time term and the timesteps-all-equal rows on random data:
    value / gradient / Hessian against the oracle restatement and central differences; overlapping
    terms (the same component regularised twice) accumulate; argument errors are reported."""
    w = workloads.make("cfg2", N=8)
    L = w.layout
    m = L.m
    rng = np.random.default_rng(4)
    Z = w.Z.copy().reshape(L.N, L.z_dim)
    Z[:, L.dt_off] = rng.uniform(0.1, 0.4, L.N)
    Z = Z.reshape(-1)
    terms = [(L.u_off, m, 1e-2), (L.u_off + m, m, [0.3, 0.7]), (L.u_off, m, 0.05)]     # u twice
    aux = Aux(L.N, L.z_dim, L.dt_off)
    J, gv, hv = aux.regularizer_eval(Z, terms, min_time_D=7.5)
    J0, g0, H0 = O.quadratic_regularizer(Z, L.N, L.z_dim, L.dt_off, terms, 7.5)
    gc, (hr, hc) = aux.regularizer_structure(terms)
    assert len(gv) == len(gc) == L.N * (3 * m + 1)          # compact: one value per regularised entry + dt
    g = np.zeros(L.N * L.z_dim)
    np.add.at(g, gc - 1, gv)                                 # u is named by two terms: its entries add up
    assert np.array_equal(g, aux.regularizer_gradient(Z, terms, 7.5))
    assert abs(J - J0) <= 1e-14 * abs(J0) and rel_err(g, g0.reshape(-1)) < 1e-14
    assert len(hr) == len(hv) == L.N * (2 * 3 * m + 1) and np.all(hr <= hc)
    dense = np.zeros((L.N * L.z_dim,) * 2)
    np.add.at(dense, (hr - 1, hc - 1), hv)
    full = np.zeros_like(dense)
    for t in range(L.N):
        full[t * L.z_dim:(t + 1) * L.z_dim, t * L.z_dim:(t + 1) * L.z_dim] = np.triu(H0[t])
    assert rel_err(dense, full) < 1e-14
    h = 1e-6
    for v in rng.choice(np.flatnonzero(g0.reshape(-1)), size=10, replace=False):
        Zp, Zm = Z.copy(), Z.copy()
        Zp[v] += h; Zm[v] -= h
        fd = (aux.regularizer_eval(Zp, terms, 7.5, want_hessian=False)[0] - aux.regularizer_eval(Zm, terms, 7.5, want_hessian=False)[0]) / (2 * h)
        assert abs(fd - g[v]) < 1e-7 * max(1.0, abs(g[v]))
    # minimum time alone: D on the first N-1 timesteps
    Jt, gt, _ = aux.regularizer_eval(Z, [], 100.0)
    dts = Z.reshape(L.N, L.z_dim)[:, L.dt_off]
    assert abs(Jt - 100.0 * dts[:-1].sum()) < 1e-12 and np.array_equal(gt, np.r_[np.full(L.N - 1, 100.0), 0.0])
    # equal timesteps
    delta, (jr, jc), vals = aux.timesteps_equal(Z, row_base=11)
    assert np.array_equal(delta, O.timesteps_all_equal(Z, L.N, L.z_dim, L.dt_off))
    assert np.array_equal(jr, 11 + 1 + np.repeat(np.arange(L.N - 1), 2)) and np.array_equal(vals, np.tile([1.0, -1.0], L.N - 1))
    assert np.array_equal(jc[0::2], np.arange(L.N - 1) * L.z_dim + L.dt_off + 1) and np.all(jc[1::2] == (L.N - 1) * L.z_dim + L.dt_off + 1)
    with pytest.raises(qcb200._lib.QcbError):
        aux.regularizer_eval(Z, [(L.z_dim - 1, 2, 1.0)])          # component leaves the knot
    aux.close()


def test_mirror_regularizer_classes(built_libs):
    """piccolo mirror: QuadraticRegularizer / MinimumTimeObjective / TimeStepsAllEqualConstraint."""
    rng = np.random.default_rng(2)
    N = 6
    traj = piccolo.NamedTrajectory({"x": rng.normal(size=(4, N)), "u": rng.normal(size=(2, N)),
                                    "du": rng.normal(size=(2, N)), "dt": rng.uniform(0.1, 0.3, size=(1, N))}, timestep="dt")
    reg = piccolo.QuadraticRegularizer("du", traj, [1e-2, 2e-2])
    ref = 0.5 * np.sum(np.array([[1e-2], [2e-2]]) * (traj["dt"] * traj["du"]) ** 2)
    assert abs(reg(traj) - ref) < 1e-15
    grad = np.empty(traj.dim * traj.N)
    reg.gradient(grad, traj)
    assert np.count_nonzero(grad) == 3 * N
    mt = piccolo.MinimumTimeObjective(traj, D=10.0)
    assert abs(mt(traj) - 10.0 * traj["dt"][0, :-1].sum()) < 1e-13
    con = piccolo.TimeStepsAllEqualConstraint(traj)
    assert np.allclose(con.values(traj), traj["dt"][0, :-1] - traj["dt"][0, -1], atol=0, rtol=0)
    (rows, cols), vals = con.jacobian(traj)
    assert len(rows) == 2 * (N - 1) and set(vals) == {1.0, -1.0}
