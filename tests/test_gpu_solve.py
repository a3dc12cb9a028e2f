"""GPU test: a solver that sees the problem only through qcb_problem_*'s five Ipopt callbacks solves the README
problem (qubit, X / Y drives, Hadamard gate; `UnitarySmoothPulseProblem` + `solve!`,
src/quantum_control_problem.jl:163-164) -- and walks the same iterates as the same solver driven by the oracle."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "helpers"))
import barrier_sqp
from oracle_problem import OracleSmoothPulse, initial_guess_and_bounds
from conftest import oracle_layout
import qc_oracle as O
from qcb200 import Aux, Evaluator, Fidelity, Problem, workloads

pytestmark = pytest.mark.gpu


class ProblemCallbacks:
    """The five callbacks as Ipopt would bind them (new_x passed through)."""

    def __init__(self, prob):
        self.p, self.n, self.m = prob, prob.n, prob.m
        self.calls = 0

    def jac_structure(self):
        return self.p.jacobian_structure()

    def hess_structure(self):
        return self.p.hessian_structure()

    def f(self, Z, new_x):
        self.calls += 1
        return self.p.eval_f(Z, new_x)

    def grad_f(self, Z, new_x):
        return self.p.eval_grad_f(Z, new_x)

    def g(self, Z, new_x):
        return self.p.eval_g(Z, new_x)

    def jac_g(self, Z, new_x):
        return self.p.eval_jac_g(Z, new_x)

    def h(self, Z, sigma, lam, new_x):
        return self.p.eval_h(Z, sigma, lam, new_x)


def test_solver_driven_by_the_library_reaches_the_gate_and_follows_the_oracle(built_libs):
    w = workloads.make("cfg1", N=12)
    L = w.layout
    m = L.m
    pairs = [(L.u_off, L.u_off + m, m), (L.u_off + m, L.u_off + 2 * m, m)]
    terms = [(L.u_off + i * m, m, 1e-2) for i in range(3)]
    ev = Evaluator(L, w.G_drift, w.G_drives)
    aux = Aux(L.N, L.z_dim, L.dt_off, derivatives=pairs)
    fid = Fidelity(Fidelity.UNITARY, w.U_goal)
    prob = Problem(L.N, L.z_dim, [ev], aux, derivative_rows=True, dt_equal_rows=True, fidelity=fid,
                   fid_index=np.arange(L.x_off[0], L.x_off[0] + L.x_dim), Q=100.0, reg_terms=terms)
    Z0, lb, ub = initial_guess_and_bounds(L, w.G_drift, w.G_drives)
    cb = ProblemCallbacks(prob)
    Z, lam, hist = barrier_sqp.solve(cb, Z0, lb, ub, max_iter=70)
    xN = Z.reshape(L.N, L.z_dim)[-1, L.x_off[0]:L.x_off[0] + L.x_dim]
    F = fid.eval(xN, want_grad=False)[0]
    assert abs(F - O.unitary_fidelity(xN, w.U_goal)) < 1e-12
    assert F > 0.9999 and np.abs(prob.eval_g(Z)).max() < 1e-5 and hist[-1][1] < 0.1
    assert np.all(Z >= lb - 1e-12) and np.all(Z <= ub + 1e-12)
    # the controls alone reproduce the gate: rollout of the solved pulse (what the reference reports as
    # "rollout fidelity", src/quantum_control_problem.jl:137-150)
    states = ev.rollout(Z)
    assert O.unitary_fidelity(np.asarray(states).reshape(-1, L.N, L.x_dim)[0, -1], w.U_goal) > 0.999
    # same solver, same start, callbacks from the oracle: the first iterates coincide
    ocb = OracleSmoothPulse(oracle_layout(L), w.G_drift, w.G_drives, w.U_goal, pairs, terms)
    _, _, ohist = barrier_sqp.solve(ocb, Z0, lb, ub, max_iter=6)
    for a, b in zip(hist[:6], ohist):
        assert abs(a[1] - b[1]) <= 1e-6 * max(1.0, abs(b[1])) and abs(a[2] - b[2]) <= 1e-7, (a, b)
    prob.close()
    for p in (ev, aux, fid):
        p.close()
