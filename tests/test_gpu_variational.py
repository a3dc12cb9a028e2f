"""GPU tests of SURVEY 8f-2 and row a3: variational integrators (state gathered from several named
components), the Lindblad / density path, and the free-phase, sensitivity and density-matrix losses,
all against the oracle (1e-10 block-normwise for the integrator, 1e-12 for the closed-form terms)."""
import numpy as np
import pytest

import qcb200
from qcb200 import Evaluator, Fidelity, FreePhaseFidelity, KnotLayout, piccolo, workloads, isomorphisms as iso
import qc_oracle as O
from conftest import oracle_layout, rel_err

pytestmark = pytest.mark.gpu
TOL = 1e-10


def coo_dict(rows, cols, vals):
    out = {}
    for r, c, v in zip(rows.tolist(), cols.tolist(), vals.tolist()):
        out[(r, c)] = out.get((r, c), 0.0) + v
    return out


def assert_same_coo(a, b, tol, what):
    assert a.keys() == b.keys(), what
    ref = max(abs(v) for v in b.values())
    err = max(abs(a[k] - b[k]) for k in b)
    assert err <= tol * ref, (what, err / ref)


# ---- row a3: Lindblad generator (dense, non-skew, c = 1) ---------------------------------------------
@pytest.mark.parametrize("path", [0, 2])
def test_lindblad_density_integrator_against_oracle(path, built_libs):
    w = workloads.make("lindblad3", N=7)
    L = w.layout
    assert L.d == 18 and L.c == 1 and np.abs(w.G_drift[0] + w.G_drift[0].T).max() > 1e-3      # not skew
    G0o, Gdo = O.lindblad_generators(*_transmon_open())
    assert np.allclose(G0o, w.G_drift[0], atol=1e-15) and np.allclose(np.array(Gdo), w.G_drives[0], atol=1e-15)
    ev = Evaluator(L, w.G_drift, w.G_drives, force_path=path)
    d0, j0, h0 = O.evaluate(oracle_layout(L), w.G_drift, w.G_drives, w.Z, w.mu)
    d1, j1, h1 = ev.eval_all(w.Z, w.mu)
    assert rel_err(d1, d0) < TOL and rel_err(j1, j0) < TOL and rel_err(h1, h0) < TOL
    ev.close()


def _transmon_open():
    a = workloads.annihilate(3)
    nop = a.conj().T @ a
    return (-0.2 / 2 * nop @ (nop - np.eye(3)), [a + a.conj().T, 1j * (a - a.conj().T)],
            [np.sqrt(0.02) * a, np.sqrt(0.01) * nop])


def test_density_trajectory_dispatch_and_trace_preservation(built_libs):
    """BilinearIntegrator(qtraj::DensityTrajectory) (src/quantum_integrators.jl:52-56): an exact Lindblad
    step has zero defect, and the propagated state keeps trace 1."""
    from scipy.linalg import expm
    Hd, drives, diss = _transmon_open()
    sys_ = piccolo.OpenQuantumSystem(Hd, drives, diss)
    N, n = 6, 3
    rng = np.random.default_rng(3)
    u = 0.1 * rng.normal(size=(2, N))
    dt = 0.1 * (1 + 0.1 * rng.uniform(-1, 1, (1, N)))
    rho = np.zeros((n, n), dtype=complex); rho[0, 0] = 1.0
    X = np.zeros((2 * n * n, N))
    X[:, 0] = iso.density_to_iso_vec(rho)
    for k in range(N - 1):
        X[:, k + 1] = expm(dt[0, k] * sys_.G(u[:, k])) @ X[:, k]
    traj = piccolo.NamedTrajectory({"rho_iso": X, "t": np.zeros((1, N)), "dt": dt, "u": u}, timestep="dt")
    qtraj = piccolo.DensityTrajectory(sys_, rho, rho)
    integ = piccolo.BilinearIntegrator(qtraj, traj)
    delta = integ.evaluate(np.zeros(integ.dim), traj)
    assert np.abs(delta).max() < 1e-13
    assert abs(O.iso_vec_to_density(X[:, -1]).trace() - 1.0) < 1e-12
    assert piccolo.test_integrator(integ, traj, atol=1e-3)["jacobian"] < 1e-6
    # pure-state infidelity objective on the terminal density matrix (src/quantum_objectives.jl:232-249)
    psi = np.array([1, 0, 0], dtype=complex)
    obj = piccolo.DensityMatrixPureStateInfidelityObjective("rho_iso", psi, traj, Q=100.0)
    J = piccolo.objective_value(obj, traj)
    assert abs(J - 100 * abs(1 - O.density_pure_state_fidelity(X[:, -1], psi))) < 1e-11


def test_density_fidelity_kind_linear(built_libs):
    rng = np.random.default_rng(8)
    n = 4
    psi = rng.normal(size=n) + 1j * rng.normal(size=n); psi /= np.linalg.norm(psi)
    x = rng.normal(size=2 * n * n)
    f = Fidelity(Fidelity.DENSITY_PURE, psi)
    F, g, H = f.eval(x, want_hess=True)
    assert abs(F - O.density_pure_state_fidelity(x, psi)) < 1e-13
    h = 1e-6
    gfd = np.array([(O.density_pure_state_fidelity(x + h * e, psi) - O.density_pure_state_fidelity(x - h * e, psi)) / (2 * h)
                    for e in np.eye(len(x))])
    assert np.abs(g - gfd).max() < 1e-9 and np.all(H == 0)
    assert abs(f.eval(iso.density_to_iso_vec(np.outer(psi, psi.conj())), want_grad=False)[0] - 1.0) < 1e-13


# ---- variational integrators: state = vcat(component, variations...) ---------------------------------
def _variational_case(kind, rng, N=5):
    n = 2
    Hvar = [workloads.Z2, workloads.X2]
    sys_ = piccolo.VariationalQuantumSystem(0.1 * workloads.Z2, [workloads.X2, workloads.Y2], Hvar,
                                            H_var_drives=[[0.3 * workloads.X2, 0 * workloads.X2], [0 * workloads.X2, 0.2 * workloads.Y2]])
    c = n if kind == "unitary" else 1
    d0 = 2 * n
    # interleave the named components with the controls so that the state is NOT contiguous in a knot
    comps = {"x": rng.normal(size=(d0 * c, N)), "dt": 0.2 * (1 + 0.1 * rng.uniform(-1, 1, (1, N))),
             "xv1": rng.normal(size=(d0 * c, N)), "u": 0.3 * rng.normal(size=(2, N)), "xv2": rng.normal(size=(d0 * c, N))}
    traj = piccolo.NamedTrajectory(comps, timestep="dt")
    scales = [2.0, 0.5]
    if kind == "unitary":
        integ = piccolo.VariationalUnitaryIntegrator(sys_, traj, "x", ["xv1", "xv2"], "u", scales=scales)
    else:
        sys1 = piccolo.VariationalQuantumSystem(0.1 * workloads.Z2, [workloads.X2, workloads.Y2], Hvar,
                                                H_var_drives=[[0.3 * workloads.X2, 0 * workloads.X2], [0 * workloads.X2, 0.2 * workloads.Y2]])
        integ = piccolo.VariationalKetIntegrator(sys1, traj, "x", ["xv1", "xv2"], "u", scale=2.0)
        scales = [2.0, 2.0]
    return sys_, traj, integ, scales, c, d0


@pytest.mark.parametrize("kind", ["ket", "unitary"])
def test_variational_integrator_against_oracle(kind, built_libs):
    """The reference's object: BilinearIntegrator(u -> var_G(...), vcat(names...), :u, traj)
    (src/quantum_integrators.jl:140-169) on x = vcat of the named components.  Oracle: the same
    generator on a trajectory whose state IS that concatenation, contiguous in the knot; Jacobian and
    Hessian entries are compared through their (row, column) positions after mapping the columns."""
    rng = np.random.default_rng(11 if kind == "ket" else 12)
    sys_, traj, integ, scales, c, d0 = _variational_case(kind, rng)
    N, nb = traj.N, 3
    names = ["x", "xv1", "xv2"]
    xcat = np.vstack([traj[nm] for nm in names])                       # (nb*d0*c, N): the reference's x
    # the oracle's generator acts on x directly: var_G(I (x) G0, [I (x) G_v]) for unitaries
    kron = (lambda G: np.kron(np.eye(c), G)) if c > 1 else (lambda G: G)
    G0 = O.var_G(kron(sys_.base.G_drift), [kron(g) / s for g, s in zip(sys_.Gv_drift, scales)])
    Gd = np.array([O.var_G(kron(sys_.base.G_drives[j]), [kron(sys_.Gv_drives[v][j]) / scales[v] for v in range(2)]) for j in range(2)])
    D = nb * d0 * c
    Lo = KnotLayout(d=D, c=1, m=2, N=N, S=1, z_dim=D + 3, x_off=[0], u_off=D, dt_off=D + 2)
    Zo = np.vstack([xcat, traj["u"], traj["dt"]]).T.reshape(-1)
    mu = rng.normal(size=(N - 1) * D)
    OL = oracle_layout(Lo)
    d0_, j0, h0 = O.evaluate(OL, G0[None], Gd[None], Zo, mu)
    jr0, jc0 = OL.jacobian_structure()
    hr0, hc0 = OL.hessian_structure()
    # column map: oracle knot position -> position in the caller's knot
    pos = np.concatenate([np.concatenate([traj.components[nm] for nm in names]), traj.components["u"], traj.components["dt"]])
    def colmap(cols):
        k, o = np.divmod(cols - 1, Lo.z_dim)
        return k * traj.dim + pos[o] + 1
    ev = integ.evaluator
    d1, j1, h1 = ev.eval_all(traj.datavec, mu)
    assert rel_err(d1, d0_) < TOL                                   # delta in the order of vcat(components)
    jr1, jc1 = ev.jacobian_structure()
    hr1, hc1 = ev.hessian_structure()
    ref_j = {k: v for k, v in coo_dict(jr0, colmap(jc0), j0).items() if v != 0.0}
    got_j = {k: v for k, v in coo_dict(jr1, jc1, j1).items() if abs(v) > 0.0}
    refmax = max(abs(v) for v in ref_j.values())
    for k_, v in ref_j.items():
        assert abs(got_j.get(k_, 0.0) - v) <= TOL * refmax, ("jacobian", k_)
    for k_, v in got_j.items():
        assert abs(ref_j.get(k_, 0.0) - v) <= TOL * refmax, ("jacobian extra", k_)
    lo = lambda r, c_: (min(r, c_), max(r, c_))
    ref_h, got_h = {}, {}
    for r, c_, v in zip(colmap(hr0).tolist(), colmap(hc0).tolist(), h0.tolist()):
        ref_h[lo(r, c_)] = ref_h.get(lo(r, c_), 0.0) + v
    for r, c_, v in zip(hr1.tolist(), hc1.tolist(), h1.tolist()):
        got_h[lo(r, c_)] = got_h.get(lo(r, c_), 0.0) + v
    refmax = max(abs(v) for v in ref_h.values())
    for k_ in set(ref_h) | set(got_h):
        assert abs(got_h.get(k_, 0.0) - ref_h.get(k_, 0.0)) <= TOL * refmax, ("hessian", k_)
    # the reference's own criterion on the drop-in object
    errs = piccolo.test_integrator(integ, traj, atol=1e-3)
    assert errs["jacobian"] < 1e-6 and errs["hessian"] < 1e-4
    # device-resident route
    import torch
    dZ, dmu = torch.from_numpy(traj.datavec).cuda(), torch.from_numpy(mu).cuda()
    dd = torch.zeros(ev.defect_dim, dtype=torch.float64, device="cuda")
    dj = torch.zeros(ev.jacobian_nnz, dtype=torch.float64, device="cuda")
    dh = torch.zeros(ev.hessian_nnz, dtype=torch.float64, device="cuda")
    ev.eval_all_device(dZ, dmu, dd, dj, dh)
    assert np.array_equal(dd.cpu().numpy(), d1) and np.array_equal(dj.cpu().numpy(), j1) and np.array_equal(dh.cpu().numpy(), h1)


def test_variational_integrator_large_path_and_adopted_structure(built_libs):
    rng = np.random.default_rng(13)
    sys_, traj, integ, scales, c, d0 = _variational_case("unitary", rng)
    mu = rng.normal(size=integ.dim)
    ref = integ.evaluator.eval_all(traj.datavec, mu)
    L = integ.layout
    ev2 = Evaluator(L, integ.evaluator._g0.swapaxes(-1, -2), integ.evaluator._gd.swapaxes(-1, -2), force_path=2)
    got = ev2.eval_all(traj.datavec, mu)
    for a, b in zip(got, ref):
        assert rel_err(a, b) < TOL
    # adopt a permuted structure through the maps
    jr, jc = ev2.jacobian_structure()
    nj = len(jr) // L.K
    p = rng.permutation(nj)
    jr2 = np.concatenate([jr[k * nj:(k + 1) * nj][p] for k in range(L.K)])
    jc2 = np.concatenate([jc[k * nj:(k + 1) * nj][p] for k in range(L.K)])
    ev2.set_structure((jr2, jc2), None)
    j2 = ev2.eval_jacobian(traj.datavec)
    assert np.array_equal(j2.reshape(L.K, nj), got[1].reshape(L.K, nj)[:, p])
    ev2.close()


def test_state_map_rejects_bad_maps(built_libs):
    L = KnotLayout(d=4, c=1, m=1, N=4, S=1, z_dim=8, x_off=[0], u_off=4, dt_off=5, x_index=[[0, 1, 2, 2]])
    with pytest.raises(qcb200.QcbError) as ei:
        Evaluator(L, np.zeros((1, 4, 4)), np.zeros((1, 1, 4, 4)))
    assert ei.value.status == 1
    L.x_index = [[0, 1, 2, 4]]        # names the control
    with pytest.raises(qcb200.QcbError):
        Evaluator(L, np.zeros((1, 4, 4)), np.zeros((1, 1, 4, 4)))
    L.x_index, L.row_index = [[0, 1, 2, 3]], [0, 1, 1, 3]
    with pytest.raises(qcb200.QcbError):
        Evaluator(L, np.zeros((1, 4, 4)), np.zeros((1, 1, 4, 4)))


# ---- free-phase fidelity, theta constraint --------------------------------------------------------------
@pytest.mark.parametrize("n,nt,side", [(2, 1, 0), (4, 2, 0), (4, 2, 1), (9, 4, 0), (8, 0, 0)])
def test_free_phase_fidelity_against_oracle(n, nt, side, built_libs):
    rng = np.random.default_rng(100 + n + nt + side)
    Ug = workloads.haar_unitary(n, rng)
    P = rng.integers(-1, 2, size=(n, nt)).astype(float) + (0.5 if nt else 0.0)
    x, th = rng.normal(size=2 * n * n), rng.normal(size=nt)
    f = FreePhaseFidelity(Ug, P, side)
    F, g, H = f.eval(x, th, want_hess=True)
    z = np.concatenate([x, th])
    F0, g0, H0 = O.free_phase_vgh(z, Ug, P, side)
    assert abs(F - O.free_phase_fidelity(z, Ug, P, side)) < 1e-12 * max(1.0, abs(F0))     # the reference's formula
    assert rel_err(g, g0) < 1e-12 and rel_err(H, H0) < 1e-12
    # with theta = 0 it is the plain unitary fidelity; a phase the map can absorb gives F = 1
    assert abs(f.eval(x, np.zeros(nt), want_grad=False)[0] - O.unitary_fidelity(x, Ug)) < 1e-12
    if nt:
        th2 = rng.normal(size=nt)
        U = O.free_phase_goal(Ug, P, th2, side)
        assert abs(f.eval(iso.operator_to_iso_vec(U), th2, want_grad=False)[0] - 1.0) < 1e-12


def test_free_phase_objective_and_theta_constraint(built_libs):
    rng = np.random.default_rng(21)
    N, n = 5, 4
    Ug = workloads.haar_unitary(n, rng)
    P = np.array([[0, 0], [1, 0], [0, 1], [1, 1]], dtype=float)          # Z-type phases on two qubits
    U = np.stack([iso.operator_to_iso_vec(workloads.haar_unitary(n, rng)) for _ in range(N)], axis=1)
    traj = piccolo.NamedTrajectory({"U_iso": U, "dt": np.full((1, N), 0.1), "u": np.zeros((2, N))}, timestep="dt",
                                   global_data={"phi1": [0.3], "phi2": [-0.7]})
    assert traj.global_dim == 2
    obj = piccolo.UnitaryFreePhaseInfidelityObjective(Ug, P, "U_iso", ["phi1", "phi2"], traj, Q=100.0)
    z = np.concatenate([U[:, -1], [0.3, -0.7]])
    F0, g0, H0 = O.free_phase_vgh(z, Ug, P)
    assert abs(obj(traj) - 100 * abs(1 - F0)) < 1e-10
    grad = obj.gradient(np.zeros(traj.dim * N + traj.global_dim), traj)
    assert np.all(grad[:(N - 1) * traj.dim] == 0)
    assert rel_err(grad[(N - 1) * traj.dim + traj.components["U_iso"]], -100 * g0[:2 * n * n]) < 1e-12
    assert rel_err(grad[N * traj.dim:], -100 * g0[2 * n * n:]) < 1e-12
    assert rel_err(obj.hessian_values(traj), -100 * H0) < 1e-12
    con = piccolo.FinalUnitaryFidelityConstraint(Ug, "U_iso", 0.99, traj, theta_names=["phi1", "phi2"], phase_map=P)
    assert con.equality is False and con.times == [N]                  # src/quantum_constraints.jl:136-143
    assert abs(con.value(traj)[0] - (0.99 - F0)) < 1e-12
    cols, vals = con.jacobian(traj)
    assert cols[-2:].tolist() == [N * traj.dim + 1, N * traj.dim + 2] and rel_err(vals, -g0) < 1e-12
    assert rel_err(con.hessian_of_lagrangian(traj, 1.5), -1.5 * H0) < 1e-12


# ---- sensitivity objective ---------------------------------------------------------------------------
def test_unitary_sensitivity_objective(built_libs):
    rng = np.random.default_rng(31)
    N, n = 7, 2
    comps = {"U_iso": rng.normal(size=(8, N)), "Uv": rng.normal(size=(8, N)), "u": rng.normal(size=(2, N)),
             "dt": np.full((1, N), 0.1)}
    traj = piccolo.NamedTrajectory(comps, timestep="dt")
    times, Qs, scale = [N, 3], [2.0, 0.5], 1.7
    obj = piccolo.UnitarySensitivityObjective("Uv", traj, times, Qs=Qs, scale=scale)
    J0, g0, H0 = O.sensitivity_objective(traj.datavec, N, traj.dim, int(traj.components["Uv"][0]), 8, times, Qs, scale)
    assert abs(obj(traj) - J0) < 1e-12 * abs(J0)
    # the reference's loss, restated directly: scale^4 |tr(U'U)|^2 / n^2 per knot
    direct = sum(q * scale ** 4 * O.unitary_self_overlap(traj["Uv"][:, t - 1]) for q, t in zip(Qs, times))
    assert abs(J0 - direct) < 1e-12 * abs(direct)
    grad = obj.gradient(np.zeros(traj.dim * N), traj)
    for i, t in enumerate(times):
        assert rel_err(grad[(t - 1) * traj.dim + traj.components["Uv"]], g0[i]) < 1e-12
    assert np.count_nonzero(grad) == 16
    assert rel_err(obj.hessian_values(traj), H0) < 1e-12
    # finite differences of the objective
    h = 1e-6
    e = np.zeros_like(traj.datavec); col = (N - 1) * traj.dim + int(traj.components["Uv"][2]); e[col] = h
    Zp, Zm = traj.datavec + e, traj.datavec - e
    fd = (O.sensitivity_objective(Zp, N, traj.dim, 8, 8, times, Qs, scale)[0] - O.sensitivity_objective(Zm, N, traj.dim, 8, 8, times, Qs, scale)[0]) / (2 * h)
    assert abs(fd - grad[col]) < 1e-6 * max(1.0, abs(fd))
    with pytest.raises(qcb200.QcbError):
        piccolo.UnitarySensitivityObjective("Uv", traj, [N + 1])(traj)
