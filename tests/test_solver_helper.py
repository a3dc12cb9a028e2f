"""CPU test of the test infrastructure a solver-driven parity check needs: the barrier-SQP helper converges on the
README problem (qubit, X / Y drives, Hadamard gate; cfg1 at a short horizon) when it is driven by the oracle's
restatement of the five Ipopt callbacks, and that restatement is consistent with its own finite differences."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "helpers"))
import barrier_sqp
from oracle_problem import OracleSmoothPulse, initial_guess_and_bounds
from conftest import oracle_layout
from qcb200 import workloads


def make_case(N=12):
    w = workloads.make("cfg1", N=N)
    L = w.layout
    m = L.m
    pairs = [(L.u_off, L.u_off + m, m), (L.u_off + m, L.u_off + 2 * m, m)]
    terms = [(L.u_off + i * m, m, 1e-2) for i in range(3)]
    return w, L, pairs, terms


def test_oracle_problem_derivatives_are_consistent():
    w, L, pairs, terms = make_case(N=6)
    cb = OracleSmoothPulse(oracle_layout(L), w.G_drift, w.G_drives, w.U_goal, pairs, terms)
    rng = np.random.default_rng(1)
    Z = w.Z.copy()
    lam = rng.normal(size=cb.m)
    jr, jc = (a - 1 for a in cb.jac_structure())
    hr, hc = (a - 1 for a in cb.hess_structure())
    J = np.zeros((cb.m, cb.n)); np.add.at(J, (jr, jc), cb.jac_g(Z))
    H = np.zeros((cb.n, cb.n)); np.add.at(H, (hr, hc), cb.h(Z, 0.8, lam))
    H = H + np.triu(H, 1).T
    grad = cb.grad_f(Z)

    def lagr_grad(Zv):
        Jv = np.zeros((cb.m, cb.n)); np.add.at(Jv, (jr, jc), cb.jac_g(Zv))
        return 0.8 * cb.grad_f(Zv) + Jv.T @ lam

    h = 1e-6
    for v in rng.choice(cb.n, size=8, replace=False):
        Zp, Zm = Z.copy(), Z.copy()
        Zp[v] += h; Zm[v] -= h
        assert abs((cb.f(Zp) - cb.f(Zm)) / (2 * h) - grad[v]) < 1e-6 * max(1.0, abs(grad[v]))
        assert np.abs((cb.g(Zp) - cb.g(Zm)) / (2 * h) - J[:, v]).max() < 1e-7 * max(1.0, np.abs(J[:, v]).max())
        assert np.abs((lagr_grad(Zp) - lagr_grad(Zm)) / (2 * h) - H[:, v]).max() < 2e-6 * max(1.0, np.abs(H[:, v]).max())


def test_solver_reaches_the_gate_on_the_oracle_problem():
    w, L, pairs, terms = make_case(N=12)
    cb = OracleSmoothPulse(oracle_layout(L), w.G_drift, w.G_drives, w.U_goal, pairs, terms)
    Z0, lb, ub = initial_guess_and_bounds(L, w.G_drift, w.G_drives)
    assert np.abs(cb.g(Z0)).max() < 1e-12 and cb.fidelity(Z0) < 0.9          # feasible start, far from the gate
    Z, lam, hist = barrier_sqp.solve(cb, Z0, lb, ub, max_iter=70)
    assert cb.fidelity(Z) > 0.9999 and np.abs(cb.g(Z)).max() < 1e-5 and hist[-1][1] < 0.1
    assert np.all(Z >= lb - 1e-12) and np.all(Z <= ub + 1e-12)
