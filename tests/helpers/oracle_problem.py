"""CPU restatement of the smooth-pulse problem's five Ipopt callbacks from the oracle's pieces (test
infrastructure): the checker for qcb_problem_* when a solver drives it (tests/test_gpu_solve.py), and the subject
of the CPU test of the solver helper itself (tests/test_solver_helper.py).  Reference: the lists assembled at
src/problem_templates/smooth_pulse_problem.jl:139-165 and _problem_templates.jl:38-67."""
import numpy as np
from scipy.linalg import expm

import qc_oracle as O


class OracleSmoothPulse:
    """Rows: [integrator defects][derivative rows][dt_t - dt_N]; objective Q |1 - F(x_N)| + regularisers."""

    def __init__(self, OL, G_drift, G_drives, U_goal, pairs, terms, Q=100.0):
        self.OL, self.G0, self.Gd, self.U_goal, self.pairs, self.terms, self.Q = OL, G_drift, G_drives, U_goal, pairs, terms, Q
        L = OL
        self.N, self.zd, self.dt_off = L.N, L.z_dim, L.dt_off
        self.xd = L.d * L.c
        self.x0 = L.x_off[0]
        K = L.N - 1
        self.n = L.N * L.z_dim
        self.rx = np.concatenate([np.arange(p[0], p[0] + p[2]) for p in pairs])
        self.rdx = np.concatenate([np.arange(p[1], p[1] + p[2]) for p in pairs])
        R = len(self.rx)
        self.dd0, self.nd = K * self.xd, K * R
        self.m = self.dd0 + self.nd + K
        ir, ic = L.jacobian_structure()
        k, r = np.divmod(np.arange(self.nd), R)
        zb = k * self.zd
        dr = np.repeat(self.dd0 + np.arange(self.nd), 4)
        dc = np.stack([zb + self.rx[r], zb + self.rdx[r], zb + self.dt_off, zb + self.zd + self.rx[r]], axis=1).reshape(-1)
        t = np.arange(K)
        tr = np.repeat(self.dd0 + self.nd + t, 2)
        tc = np.stack([t * self.zd + self.dt_off, np.full(K, K * self.zd + self.dt_off)], axis=1).reshape(-1)
        self.jr = np.concatenate([ir - 1, dr, tr])
        self.jc = np.concatenate([ic - 1, dc, tc])
        self.hr0, self.hc0 = (a - 1 for a in L.hessian_structure())
        a_, b_ = zb + self.rdx[r], zb + self.dt_off
        self.dhr, self.dhc = np.minimum(a_, b_), np.maximum(a_, b_)
        self.b = (L.N - 1) * self.zd + self.x0
        self._triu = np.triu_indices(self.n)

    def jac_structure(self):
        return self.jr + 1, self.jc + 1

    def hess_structure(self):
        return self._triu[0] + 1, self._triu[1] + 1

    def fidelity(self, Z):
        return O.fidelity_vgh(O.KIND_UNITARY, Z[self.b:self.b + self.xd], self.U_goal)[0]

    def f(self, Z, new_x=True):
        J = O.quadratic_regularizer(Z, self.N, self.zd, self.dt_off, self.terms)[0]
        return self.Q * abs(1 - self.fidelity(Z)) + J

    def grad_f(self, Z, new_x=True):
        F, gF, _ = O.fidelity_vgh(O.KIND_UNITARY, Z[self.b:self.b + self.xd], self.U_goal)
        g = O.quadratic_regularizer(Z, self.N, self.zd, self.dt_off, self.terms)[1].reshape(-1).copy()
        g[self.b:self.b + self.xd] += -self.Q * np.sign(1 - F) * gF
        return g

    def g(self, Z, new_x=True):
        d0 = O.defect_only(self.OL, self.G0, self.Gd, Z)
        dd = O.derivative_integrator(Z, self.N, self.zd, self.dt_off, self.pairs)[0]
        return np.concatenate([np.asarray(d0).reshape(-1), dd.reshape(-1), O.timesteps_all_equal(Z, self.N, self.zd, self.dt_off)])

    def jac_g(self, Z, new_x=True):
        jv = O.evaluate(self.OL, self.G0, self.Gd, Z, None, want_hess=False)[1]
        dj = O.derivative_integrator(Z, self.N, self.zd, self.dt_off, self.pairs)[1].reshape(-1)
        return np.concatenate([jv, dj, np.tile([1.0, -1.0], self.N - 1)])

    def h(self, Z, sigma, lam, new_x=True):
        H = np.zeros((self.n, self.n))
        hv = O.evaluate(self.OL, self.G0, self.Gd, Z, lam[:self.dd0])[2]
        np.add.at(H, (self.hr0, self.hc0), hv)
        dh = O.derivative_integrator(Z, self.N, self.zd, self.dt_off, self.pairs, lam[self.dd0:self.dd0 + self.nd])[2].reshape(-1)
        np.add.at(H, (self.dhr, self.dhc), dh)
        F, _, HF = O.fidelity_vgh(O.KIND_UNITARY, Z[self.b:self.b + self.xd], self.U_goal)
        HR = O.quadratic_regularizer(Z, self.N, self.zd, self.dt_off, self.terms)[2]
        H[self.b:self.b + self.xd, self.b:self.b + self.xd] += np.triu(sigma * -self.Q * np.sign(1 - F) * HF)
        for t in range(self.N):
            s = slice(t * self.zd, (t + 1) * self.zd)
            H[s, s] += sigma * np.triu(HR[t])
        return H[self._triu]


def initial_guess_and_bounds(L, G_drift, G_drives, dt=0.2, a_bound=1.0, dda_bound=5.0):
    """A feasible start the way the reference's templates make one (rollout of smooth initial controls,
    src/quantum_control_problem.jl:137-150) and the bounds they put on the trajectory: x_1 = identity, u_1 = u_N = 0,
    |u| <= a_bound, |ddu| <= dda_bound, dt in [dt/2, 3 dt/2], t fixed."""
    N, zd, m, xd = L.N, L.z_dim, L.m, L.d * L.c
    Zm = np.zeros((N, zd))
    Zm[:, L.dt_off] = dt
    Zm[:, L.dt_off - 1] = np.arange(N) * dt
    tt = np.linspace(0, 1, N)[:, None]
    Zm[:, L.u_off:L.u_off + m] = 0.2 * np.sin(np.pi * tt) * np.resize([1.0, -0.6], m)[None, :]
    n = L.d // 2
    x = O.operator_to_iso_vec(np.eye(n, dtype=complex))
    ident = x.copy()
    for k in range(N):
        Zm[k, L.x_off[0]:L.x_off[0] + xd] = x
        if k < N - 1:
            G = G_drift[0] + sum(Zm[k, L.u_off + j] * G_drives[0][j] for j in range(m))
            x = (expm(dt * G) @ x.reshape(L.c, L.d).T).T.reshape(-1)
    u, du, ddu = (slice(L.u_off + i * m, L.u_off + (i + 1) * m) for i in range(3))
    Zm[:-1, du] = (Zm[1:, u] - Zm[:-1, u]) / dt
    Zm[:-1, ddu] = (Zm[1:, du] - Zm[:-1, du]) / dt
    lb, ub = np.full((N, zd), -np.inf), np.full((N, zd), np.inf)
    xs = slice(L.x_off[0], L.x_off[0] + xd)
    lb[:, xs], ub[:, xs] = -1.5, 1.5
    lb[0, xs] = ub[0, xs] = ident
    lb[:, L.dt_off - 1] = ub[:, L.dt_off - 1] = Zm[:, L.dt_off - 1]
    lb[:, L.dt_off], ub[:, L.dt_off] = 0.5 * dt, 1.5 * dt
    lb[:, u], ub[:, u] = -a_bound, a_bound
    lb[0, u] = ub[0, u] = 0.0
    lb[-1, u] = ub[-1, u] = 0.0
    lb[:, ddu], ub[:, ddu] = -dda_bound, dda_bound
    return Zm.reshape(-1), lb.reshape(-1), ub.reshape(-1)
