"""GPU tests of qcb_problem_* (one C call per Ipopt callback): the composite of the smooth-pulse
template (src/problem_templates/smooth_pulse_problem.jl:139-165) against the oracle's pieces laid out by hand,
against central differences of its own callbacks, and through Ipopt's new_x protocol."""
import numpy as np
import pytest

import qcb200
from qcb200 import Aux, Evaluator, Fidelity, Problem, workloads
import qc_oracle as O
from conftest import oracle_layout, rel_err

pytestmark = pytest.mark.gpu


def _case(N=7, seed=5):
    w = workloads.make("cfg2", N=N)                 # 3-level transmon: d = 6, c = 3, m = 2; x, t, dt, u, du, ddu
    L = w.layout
    rng = np.random.default_rng(seed)
    Z = w.Z.copy().reshape(L.N, L.z_dim)
    Z[:, L.dt_off] = rng.uniform(0.1, 0.3, L.N)
    Z[:, L.u_off:L.u_off + 3 * L.m] = 0.3 * rng.normal(size=(L.N, 3 * L.m))
    Z = np.ascontiguousarray(Z.reshape(-1))
    U_goal = workloads.haar_unitary(3, rng)
    m = L.m
    pairs = [(L.u_off, L.u_off + m, m), (L.u_off + m, L.u_off + 2 * m, m)]
    terms = [(L.u_off, m, 1e-2), (L.u_off + m, m, [0.3, 0.7]), (L.u_off + 2 * m, m, 0.05)]
    return w, L, Z, U_goal, pairs, terms, rng


def _build(w, L, U_goal, pairs, terms, Q=100.0, D=0.0, fid_hessian=True):
    ev = Evaluator(L, w.G_drift, w.G_drives)
    aux = Aux(L.N, L.z_dim, L.dt_off, derivatives=pairs)
    fid = Fidelity(Fidelity.UNITARY, U_goal)
    fid_index = np.arange(L.x_off[0], L.x_off[0] + L.x_dim)
    prob = Problem(L.N, L.z_dim, [ev], aux, derivative_rows=True, dt_equal_rows=True, fidelity=fid, fid_index=fid_index,
                   Q=Q, fid_hessian=fid_hessian, reg_terms=terms, min_time_D=D)
    return prob, (ev, aux, fid)


def _dense(rows, cols, vals, shape):
    M = np.zeros(shape)
    np.add.at(M, (rows - 1, cols - 1), vals)
    return M


def test_problem_matches_the_oracle_pieces(built_libs):
    w, L, Z, U_goal, pairs, terms, rng = _case()
    Q, D = 100.0, 3.0
    prob, pieces = _build(w, L, U_goal, pairs, terms, Q, D)
    ev, aux, fid = pieces
    n, K = L.N * L.z_dim, L.N - 1
    nd = K * sum(p[2] for p in pairs)
    assert (prob.n, prob.m) == (n, ev.defect_dim + nd + K)
    assert prob.jacobian_nnz == ev.jacobian_nnz + 4 * nd + 2 * K
    xd = L.x_dim
    rows_reg = sum(t[1] for t in terms)
    assert prob.hessian_nnz == ev.hessian_nnz + nd + xd * (xd + 1) // 2 + L.N * (2 * rows_reg + 1)

    OL = oracle_layout(L)
    lam = rng.normal(size=prob.m)
    sigma = 0.7
    d0, j0, h0 = O.evaluate(OL, w.G_drift, w.G_drives, Z, lam[:ev.defect_dim])
    dd, dj, dh = O.derivative_integrator(Z, L.N, L.z_dim, L.dt_off, pairs, lam[ev.defect_dim:ev.defect_dim + nd])
    xN = Z.reshape(L.N, L.z_dim)[-1, L.x_off[0]:L.x_off[0] + xd]
    F0, gF0, HF0 = O.fidelity_vgh(O.KIND_UNITARY, xN, U_goal)
    J0, gR0, HR0 = O.quadratic_regularizer(Z, L.N, L.z_dim, L.dt_off, terms, D)

    # objective and its gradient
    f = prob.eval_f(Z)
    assert abs(f - (Q * abs(1 - F0) + J0)) <= 1e-13 * abs(f)
    g_want = gR0.reshape(-1).copy()
    g_want[(L.N - 1) * L.z_dim + L.x_off[0]:(L.N - 1) * L.z_dim + L.x_off[0] + xd] += -Q * np.sign(1 - F0) * gF0
    assert rel_err(prob.eval_grad_f(Z), g_want) < 1e-13

    # constraint rows and their Jacobian
    g = prob.eval_g(Z)
    g_rows = np.concatenate([d0, dd.reshape(-1), O.timesteps_all_equal(Z, L.N, L.z_dim, L.dt_off)])
    assert rel_err(g[:ev.defect_dim], g_rows[:ev.defect_dim]) < 1e-10 and rel_err(g[ev.defect_dim:], g_rows[ev.defect_dim:]) < 1e-14
    jr, jc = prob.jacobian_structure()
    assert jr.min() >= 1 and jr.max() == prob.m and jc.min() >= 1 and jc.max() <= n
    jv = prob.eval_jac_g(Z)
    ir, ic = OL.jacobian_structure()
    Jd = _dense(jr, jc, jv, (prob.m, n))
    J_want = np.zeros((prob.m, n))
    np.add.at(J_want, (ir - 1, ic - 1), j0)
    (ar, ac), (ahr, ahc) = aux.derivative_structure(row_base=ev.defect_dim)
    np.add.at(J_want, (ar - 1, ac - 1), dj.reshape(-1))
    for t in range(K):
        J_want[ev.defect_dim + nd + t, t * L.z_dim + L.dt_off] += 1.0
        J_want[ev.defect_dim + nd + t, K * L.z_dim + L.dt_off] -= 1.0
    assert rel_err(Jd, J_want) < 1e-10

    # Hessian of the Lagrangian
    hr, hc = prob.hessian_structure()
    assert np.all(hr <= hc) and hr.min() >= 1 and hc.max() <= n
    hv = prob.eval_h(Z, sigma, lam)
    Hd = _dense(hr, hc, hv, (n, n))
    H_want = np.zeros((n, n))
    r0, c0 = OL.hessian_structure()
    np.add.at(H_want, (r0 - 1, c0 - 1), h0)
    np.add.at(H_want, (ahr - 1, ahc - 1), dh.reshape(-1))
    b = (L.N - 1) * L.z_dim + L.x_off[0]
    H_want[b:b + xd, b:b + xd] += np.triu(sigma * -Q * np.sign(1 - F0) * HF0)
    for t in range(L.N):
        H_want[t * L.z_dim:(t + 1) * L.z_dim, t * L.z_dim:(t + 1) * L.z_dim] += sigma * np.triu(HR0[t])
    assert rel_err(Hd, H_want) < 1e-10
    prob.close()
    for p in pieces:
        p.close()


def test_problem_callbacks_are_consistent_with_central_differences(built_libs):
    """grad f, the Jacobian columns and the Lagrangian Hessian columns against central differences of the
    problem's own eval_f / eval_g / (grad f + J' lambda): offsets and structures of all pieces at once."""
    w, L, Z, U_goal, pairs, terms, rng = _case(N=5, seed=9)
    prob, pieces = _build(w, L, U_goal, pairs, terms, Q=50.0, D=2.0)
    n = prob.n
    lam = rng.normal(size=prob.m)
    sigma = 1.3
    jr, jc = prob.jacobian_structure()
    hr, hc = prob.hessian_structure()
    grad = prob.eval_grad_f(Z).copy()
    Jd = _dense(jr, jc, prob.eval_jac_g(Z), (prob.m, n))
    Hu = _dense(hr, hc, prob.eval_h(Z, sigma, lam), (n, n))
    Hs = Hu + np.triu(Hu, 1).T

    def lagr_grad(Zv):
        J = _dense(jr, jc, prob.eval_jac_g(Zv), (prob.m, n))
        return sigma * prob.eval_grad_f(Zv) + J.T @ lam

    h = 1e-6
    cols = np.concatenate([rng.choice(n, size=10, replace=False),
                           [(L.N - 1) * L.z_dim + L.x_off[0] + 3, 2 * L.z_dim + L.dt_off, L.z_dim + L.u_off + 1]])
    for v in cols:
        Zp, Zm = Z.copy(), Z.copy()
        Zp[v] += h; Zm[v] -= h
        assert abs((prob.eval_f(Zp) - prob.eval_f(Zm)) / (2 * h) - grad[v]) < 1e-6 * max(1.0, abs(grad[v]))
        fd_g = (prob.eval_g(Zp) - prob.eval_g(Zm)) / (2 * h)
        assert np.abs(fd_g - Jd[:, v]).max() < 1e-7 * max(1.0, np.abs(Jd[:, v]).max())
        fd_h = (lagr_grad(Zp) - lagr_grad(Zm)) / (2 * h)
        assert np.abs(fd_h - Hs[:, v]).max() < 2e-6 * max(1.0, np.abs(Hs[:, v]).max())
    prob.close()
    for p in pieces:
        p.close()


def test_problem_new_x_protocol_and_errors(built_libs):
    """eval_f(new_x = 1) -> eval_g(0) -> eval_jac_g(0) -> eval_h(0) gives bit-identical results to calls that all
    say new_x = 1, on two different iterates in a row (the integrator must not keep the first one)."""
    w, L, Z, U_goal, pairs, terms, rng = _case(N=6, seed=2)
    prob, pieces = _build(w, L, U_goal, pairs, terms)
    prob2, pieces2 = _build(w, L, U_goal, pairs, terms)
    lam = rng.normal(size=prob.m)
    for it in range(2):
        Zi = Z + 0.01 * it * rng.normal(size=Z.size)
        a = (prob.eval_f(Zi, True), prob.eval_grad_f(Zi, False).copy(), prob.eval_g(Zi, False).copy(),
             prob.eval_jac_g(Zi, False).copy(), prob.eval_h(Zi, 1.0, lam, False).copy())
        b = (prob2.eval_f(Zi, True), prob2.eval_grad_f(Zi, True).copy(), prob2.eval_g(Zi, True).copy(),
             prob2.eval_jac_g(Zi, True).copy(), prob2.eval_h(Zi, 1.0, lam, True).copy())
        assert a[0] == b[0]
        for x, y in zip(a[1:], b[1:]):
            assert np.array_equal(x, y)
        if it == 0:
            first = a
    assert not np.array_equal(first[2], a[2])       # the second iterate was really evaluated
    # no dense terminal block on request; a linear (density) loss has none either
    prob3, pieces3 = _build(w, L, U_goal, pairs, terms, fid_hessian=False)
    assert prob3.hessian_nnz == prob.hessian_nnz - L.x_dim * (L.x_dim + 1) // 2
    # argument errors surface with the piece's message
    with pytest.raises(qcb200._lib.QcbError) as ei:
        Problem(L.N, L.z_dim, [pieces[0]], None, derivative_rows=True)
    assert ei.value.status == 1 and "qcb_aux" in str(ei.value)
    with pytest.raises(qcb200._lib.QcbError):
        Problem(L.N + 1, L.z_dim, [pieces[0]])                                       # another trajectory shape
    slab = Evaluator(L, w.G_drift, w.G_drives, k_begin=0, k_end=2)
    with pytest.raises(qcb200._lib.QcbError) as ei:
        Problem(L.N, L.z_dim, [slab])
    assert "whole trajectory" in str(ei.value)
    slab.close()
    for pr, ps in ((prob, pieces), (prob2, pieces2), (prob3, pieces3)):
        pr.close()
        for p in ps:
            p.close()


def test_problem_on_the_large_path_reuses_the_integrator_state(built_libs):
    """cfg5 shape (d = 128) at a short horizon: the composite's integrator blocks are bit-identical to the
    evaluator's own fused call, and after eval_g the Jacobian and Hessian callbacks do not upload Z again."""
    w = workloads.make("cfg5", N=5)
    L = w.layout
    rng = np.random.default_rng(17)
    m = L.m
    pairs = [(L.u_off, L.u_off + m, m), (L.u_off + m, L.u_off + 2 * m, m)]
    terms = [(L.u_off + i * m, m, 1e-2) for i in range(3)]
    ev, ev_ref = Evaluator(L, w.G_drift, w.G_drives), Evaluator(L, w.G_drift, w.G_drives)
    aux = Aux(L.N, L.z_dim, L.dt_off, derivatives=pairs)
    fid = Fidelity(Fidelity.UNITARY, workloads.haar_unitary(64, rng))
    prob = Problem(L.N, L.z_dim, [ev], aux, derivative_rows=True, dt_equal_rows=True, fidelity=fid,
                   fid_index=np.arange(L.x_off[0], L.x_off[0] + L.x_dim), fid_hessian=False, reg_terms=terms)
    lam = rng.normal(size=prob.m)
    nd, nj, nh = ev.defect_dim, ev.jacobian_nnz, ev.hessian_nnz
    f = prob.eval_f(w.Z, True)
    xN = w.Z.reshape(L.N, L.z_dim)[-1, L.x_off[0]:L.x_off[0] + L.x_dim]
    J0 = O.quadratic_regularizer(w.Z, L.N, L.z_dim, L.dt_off, terms)[0]
    assert abs(f - (100.0 * abs(1 - fid.eval(xN, want_grad=False)[0]) + J0)) <= 1e-12 * abs(f)
    g = prob.eval_g(w.Z, False)
    z_bytes = 8 * L.N * L.z_dim
    assert ev.transfer_bytes()[0] >= z_bytes                       # first constraint callback of the iterate: Z goes up
    jv = prob.eval_jac_g(w.Z, False)
    assert ev.transfer_bytes()[0] == 0
    hv = prob.eval_h(w.Z, 1.0, lam, False)
    assert ev.transfer_bytes()[0] == 8 * nd                        # the multipliers only
    d1, j1, h1 = ev_ref.eval_all(w.Z, lam[:nd])
    assert np.array_equal(g[:nd], d1) and np.array_equal(jv[:nj], j1) and np.array_equal(hv[:nh], h1)
    grad = prob.eval_grad_f(w.Z, False)
    assert grad.size == prob.n and np.count_nonzero(grad) <= L.x_dim + L.N * (3 * m + 1)
    prob.close()
    for p in (ev, ev_ref, aux, fid):
        p.close()


def test_problem_with_leakage_rows_and_objective(built_libs):
    """LeakageConstraint rows (x^2 - value <= 0) and the LeakageObjective term of a 3-level transmon inside the
    composite: values against the oracle, derivatives against central differences of the composite itself."""
    w, L, Z, U_goal, pairs, terms, rng = _case(N=5, seed=21)
    U = np.arange(18).reshape(3, 6).T                     # column-major (6 x 3) index map of the iso-vec
    idx = sorted(set(U[[2, 5], :].reshape(-1)) | set(U[:, 2]))
    ev = Evaluator(L, w.G_drift, w.G_drives)
    aux = Aux(L.N, L.z_dim, L.dt_off, derivatives=pairs, leak_x_off=L.x_off[0], leak_idx=idx)
    fid = Fidelity(Fidelity.UNITARY, U_goal)
    base = dict(derivative_rows=True, dt_equal_rows=True, fidelity=fid, fid_index=np.arange(L.x_off[0], L.x_off[0] + L.x_dim),
                Q=100.0, reg_terms=terms)
    plain = Problem(L.N, L.z_dim, [ev], aux, **base)
    prob = Problem(L.N, L.z_dim, [ev], aux, leakage_rows=True, leakage_value=1e-2, leakage_Q=3.0, **base)
    nl = L.N * len(idx)
    assert (prob.m, prob.jacobian_nnz, prob.hessian_nnz) == (plain.m + nl, plain.jacobian_nnz + nl, plain.hessian_nnz + 2 * nl)
    obj0, grad0, con0, cjac0, _ = O.leakage(Z, L.N, L.z_dim, L.x_off[0], idx, None, 1e-2, None)
    assert abs(prob.eval_f(Z) - (plain.eval_f(Z) + 3.0 * obj0)) < 1e-12 * abs(plain.eval_f(Z))
    g = prob.eval_g(Z)
    assert np.array_equal(g[:plain.m], plain.eval_g(Z)) and rel_err(g[plain.m:], con0.reshape(-1)) < 1e-14
    n = prob.n
    jr, jc = prob.jacobian_structure()
    hr, hc = prob.hessian_structure()
    assert np.all(hr <= hc) and jr.max() == prob.m
    lam = rng.normal(size=prob.m)
    grad = prob.eval_grad_f(Z).copy()
    Jd = _dense(jr, jc, prob.eval_jac_g(Z), (prob.m, n))
    Hu = _dense(hr, hc, prob.eval_h(Z, 0.9, lam), (n, n))
    Hs = Hu + np.triu(Hu, 1).T

    def lagr_grad(Zv):
        return 0.9 * prob.eval_grad_f(Zv) + _dense(jr, jc, prob.eval_jac_g(Zv), (prob.m, n)).T @ lam

    h = 1e-6
    leak_cols = [t * L.z_dim + L.x_off[0] + i for t in (0, 3) for i in idx[:3]]
    for v in list(rng.choice(n, size=6, replace=False)) + leak_cols:
        Zp, Zm = Z.copy(), Z.copy()
        Zp[v] += h; Zm[v] -= h
        assert abs((prob.eval_f(Zp) - prob.eval_f(Zm)) / (2 * h) - grad[v]) < 1e-6 * max(1.0, abs(grad[v]))
        assert np.abs((prob.eval_g(Zp) - prob.eval_g(Zm)) / (2 * h) - Jd[:, v]).max() < 1e-7 * max(1.0, np.abs(Jd[:, v]).max())
        assert np.abs((lagr_grad(Zp) - lagr_grad(Zm)) / (2 * h) - Hs[:, v]).max() < 2e-6 * max(1.0, np.abs(Hs[:, v]).max())
    with pytest.raises(qcb200._lib.QcbError) as ei:        # an aux without leakage indices cannot serve leakage rows
        Problem(L.N, L.z_dim, [ev], Aux(L.N, L.z_dim, L.dt_off, derivatives=pairs), leakage_rows=True, **base)
    assert "leakage" in str(ei.value)
    for p in (prob, plain, ev, aux, fid):
        p.close()


def test_sampling_problem_sums_the_terminal_term_over_the_systems(built_libs):
    """UnitarySamplingProblem shape (robust qubit, three systems sharing one pulse, sampling_problem.jl:139-142): one
    integrator handle with S = 3, the terminal infidelity summed over the three final states with weights."""
    w = workloads.make("cfg4", N=6, S=3)
    L = w.layout
    assert L.S == 3 and L.z_dim == L.u_off + L.m
    rng = np.random.default_rng(4)
    Z = w.Z.copy()
    wts = np.array([1.0, 0.5, 2.0])
    terms = [(L.u_off, L.m, 0.1)]
    ev = Evaluator(L, w.G_drift, w.G_drives)
    aux = Aux(L.N, L.z_dim, L.dt_off)
    fid = Fidelity(Fidelity.UNITARY, w.U_goal)
    fid_index = np.array([np.arange(L.x_off[s], L.x_off[s] + L.x_dim) for s in range(3)])
    prob = Problem(L.N, L.z_dim, [ev], aux, dt_equal_rows=True, fidelity=fid, fid_index=fid_index, fid_weights=wts,
                   Q=100.0, reg_terms=terms)
    K, xd, n = L.N - 1, L.x_dim, L.N * L.z_dim
    assert prob.m == 3 * K * xd + K and prob.hessian_nnz == ev.hessian_nnz + 3 * xd * (xd + 1) // 2 + L.N * (2 * L.m + 1)
    last = Z.reshape(L.N, L.z_dim)[-1]
    Fs = [O.unitary_fidelity(last[L.x_off[s]:L.x_off[s] + xd], w.U_goal) for s in range(3)]
    J0, gR0, _ = O.quadratic_regularizer(Z, L.N, L.z_dim, L.dt_off, terms)
    f = prob.eval_f(Z)
    assert abs(f - (sum(wt * 100.0 * abs(1 - F) for wt, F in zip(wts, Fs)) + J0)) < 1e-12 * abs(f)
    g_want = gR0.reshape(-1).copy()
    for s in range(3):
        _, gF, _ = O.fidelity_vgh(O.KIND_UNITARY, last[L.x_off[s]:L.x_off[s] + xd], w.U_goal)
        b = (L.N - 1) * L.z_dim + L.x_off[s]
        g_want[b:b + xd] += -100.0 * wts[s] * np.sign(1 - Fs[s]) * gF
    grad = prob.eval_grad_f(Z).copy()
    assert rel_err(grad, g_want) < 1e-13
    d0, j0, _ = O.evaluate(oracle_layout(L), w.G_drift, w.G_drives, Z, None, want_hess=False)
    g = prob.eval_g(Z)
    assert rel_err(g[:ev.defect_dim], d0) < 1e-10
    jr, jc = prob.jacobian_structure()
    hr, hc = prob.hessian_structure()
    lam = rng.normal(size=prob.m)
    Hu = _dense(hr, hc, prob.eval_h(Z, 1.1, lam), (n, n))
    Hs = Hu + np.triu(Hu, 1).T

    def lagr_grad(Zv):
        return 1.1 * prob.eval_grad_f(Zv) + _dense(jr, jc, prob.eval_jac_g(Zv), (prob.m, n)).T @ lam

    h = 1e-6
    cols = list(rng.choice(n, size=5, replace=False)) + [(L.N - 1) * L.z_dim + L.x_off[s] + 2 for s in range(3)]
    for v in cols:
        Zp, Zm = Z.copy(), Z.copy()
        Zp[v] += h; Zm[v] -= h
        assert abs((prob.eval_f(Zp) - prob.eval_f(Zm)) / (2 * h) - grad[v]) < 1e-6 * max(1.0, abs(grad[v]))
        assert np.abs((lagr_grad(Zp) - lagr_grad(Zm)) / (2 * h) - Hs[:, v]).max() < 2e-6 * max(1.0, np.abs(Hs[:, v]).max())
    prob.close()
    for p in (ev, aux, fid):
        p.close()
