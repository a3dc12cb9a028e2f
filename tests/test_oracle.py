"""CPU tests: pin the oracle against everything the reference tree holds for this path, and the
C port against the oracle.  (The reference's evaluator itself cannot run here; its logged outputs are replayed in
tests/test_reference_fixtures.py.)"""
import json
import os

import numpy as np
import pytest

import qcb200
from qcb200 import workloads
import qc_oracle as O
import qc_port as P
from conftest import ROOT, oracle_layout, rel_err

GOLD = os.path.join(ROOT, "tests", "golden")


# ---- reference fixture: iso-vec layout (test/test_utils.jl:52-103) -----------------------------
def test_layout_fixture_hadamard_geodesic():
    fx = json.load(open(os.path.join(GOLD, "hadamard_geodesic.json")))
    U_iso = np.array(fx["U_iso"])
    H = np.array([[1, 1], [1, -1]]) / np.sqrt(2)
    X = np.array([[0, 1], [1, 0]])
    assert np.allclose(O.operator_to_iso_vec(np.eye(2)), fx["initial"])
    assert np.allclose(O.operator_to_iso_vec(X), fx["goal_X"])
    assert np.allclose(O.operator_to_iso_vec(H), U_iso[:, 4], atol=5e-7)
    # every column decodes to a unitary on the geodesic exp(k/4 log H) (print rounding 5e-7)
    from scipy.linalg import logm, expm
    lg = logm(H.astype(complex))
    for k in range(5):
        U = O.iso_vec_to_operator(U_iso[:, k])
        assert np.abs(U.conj().T @ U - np.eye(2)).max() < 5e-6
        assert np.abs(U - expm(k / 4 * lg)).max() < 5e-6
    # the product-side helpers use the same layout
    from qcb200 import isomorphisms as iso
    assert np.allclose(iso.operator_to_iso_vec(H), O.operator_to_iso_vec(H))
    assert np.allclose(iso.iso_vec_to_operator(U_iso[:, 2]), O.iso_vec_to_operator(U_iso[:, 2]))


# ---- reference known answers: CoherentKetInfidelityObjective (src/quantum_objectives.jl:315-394) --
def test_coherent_ket_known_answers():
    psi0, psi1 = np.array([1, 0], dtype=complex), np.array([0, 1], dtype=complex)
    goals = [psi1, psi0]
    rng = np.random.default_rng(0)
    xs = [v / np.linalg.norm(v) for v in rng.normal(size=(2, 4))]
    J = 100 * abs(1 - O.coherent_ket_fidelity(xs, goals))
    assert 0.0 <= J <= 100.0
    perfect = [O.ket_to_iso(psi1), O.ket_to_iso(psi0)]
    assert 100 * abs(1 - O.coherent_ket_fidelity(perfect, goals)) < 1e-10
    phase = [O.ket_to_iso(psi1), O.ket_to_iso(-psi0)]
    assert 100 * abs(1 - O.coherent_ket_fidelity(phase, goals)) > 50.0
    # quadratic-form restatement agrees with the direct formula, and has a non-zero gradient
    x = np.concatenate(xs)
    F, g, H = O.fidelity_vgh(O.KIND_COHERENT_KET, x, goals)
    assert abs(F - O.coherent_ket_fidelity(xs, goals)) < 1e-15
    assert not np.all(g == 0)


@pytest.mark.parametrize("kind", ["unitary", "ket", "subspace"])
def test_fidelity_quadratic_forms(kind):
    rng = np.random.default_rng(3)
    if kind == "unitary":
        Ug = workloads.haar_unitary(3, rng)
        x = rng.normal(size=18)
        F, g, H = O.fidelity_vgh(O.KIND_UNITARY, x, Ug)
        direct = lambda y: O.unitary_fidelity(y, Ug)
    elif kind == "ket":
        pg = rng.normal(size=3) + 1j * rng.normal(size=3)
        pg /= np.linalg.norm(pg)
        x = rng.normal(size=6)
        F, g, H = O.fidelity_vgh(O.KIND_KET, x, pg)
        direct = lambda y: O.ket_fidelity(y, pg)
    else:
        Ug = workloads.haar_unitary(2, rng)
        x = rng.normal(size=18)
        F, g, H = O.fidelity_vgh(O.KIND_UNITARY_SUBSPACE, x, Ug, n_full=3, subspace=[0, 1])
        direct = lambda y: O.unitary_subspace_fidelity(y, Ug, [0, 1])
    assert abs(F - direct(x)) < 1e-13
    h = 1e-6
    fd = np.array([(direct(x + h * e) - direct(x - h * e)) / (2 * h) for e in np.eye(len(x))])
    assert np.abs(fd - g).max() < 1e-8
    assert np.abs(H - H.T).max() < 1e-15


# ---- reference structure cardinalities (docs Ipopt logs) ---------------------
def test_structure_cardinalities_match_docs_logs():
    # 5-level transmon, N=50: x_dim 50, m 2 (docs/literate/examples/multilevel_transmon.jl:105)
    L = O.Layout(10, 5, 2, 50, 1, 57, [0], 52, 51, jac_mode=0)
    assert L.jac_nnz_unit == 2500 + 50 + 100 + 50
    assert 49 * L.jac_nnz_unit - 2500 - 100 + (49 * 16 - 4) + 98 == 130578
    # two-qubit, N=100: x_dim 32, m 4 (docs/literate/examples/two_qubit_gates.jl:135)
    L = O.Layout(8, 4, 4, 100, 1, 45, [0], 34, 33, jac_mode=0)
    assert 99 * L.jac_nnz_unit - 1024 - 128 + (99 * 32 - 8) + 198 == 122590
    r, c = L.jacobian_structure()
    assert len(r) == 99 * L.jac_nnz_unit and r.min() == 1 and r.max() == 99 * 32
    hr, hc = L.hessian_structure()
    assert np.all(hr <= hc) and len(hr) == 99 * L.hess_nnz_unit


# ---- physics closure: exact steps give zero defect (the reference's post-solve criterion) --------
def test_exact_step_trajectory_has_zero_defect():
    from scipy.linalg import expm
    w = workloads.make("cfg2", N=8)
    L = w.layout
    Z = w.Z.reshape(L.N, L.z_dim).copy()
    n = L.d // 2
    U = np.eye(n, dtype=complex)
    Hd = -0.1 * np.diag([0, 0, 2.0])
    a = workloads.annihilate(3)
    Hs = [a + a.conj().T, 1j * (a - a.conj().T)]
    for k in range(L.N):
        Z[k, :L.x_dim] = O.operator_to_iso_vec(U)
        Hk = Hd + sum(Z[k, L.u_off + j] * Hs[j] for j in range(2))
        U = expm(-1j * Z[k, L.dt_off] * Hk) @ U
    delta = O.defect_only(oracle_layout(L), w.G_drift, w.G_drives, Z.reshape(-1))
    assert np.abs(delta).max() < 1e-14
    d2, _, _ = P.evaluate(oracle_layout(L), w.G_drift, w.G_drives, Z.reshape(-1), None, False, False)
    assert np.abs(d2).max() < 1e-14


# ---- derivative self-consistency (the reference's test_integrator criterion) ---------------------
@pytest.mark.parametrize("n,m,dt", [(2, 2, 0.2), (3, 2, 1.0), (4, 3, 2.0)])
def test_oracle_finite_difference_consistency(n, m, dt):
    rng = np.random.default_rng(n)
    rh = lambda: (lambda A: (A + A.conj().T) / 2)(rng.normal(size=(n, n)) + 1j * rng.normal(size=(n, n)))
    G0, Gd = O.G_of_H(rh()), np.array([O.G_of_H(rh()) for _ in range(m)])
    X, Xn, M = (rng.normal(size=(2 * n, n)) for _ in range(3))
    errs = O.fd_check_unit(G0, Gd, X, Xn, M, 0.3 * rng.normal(size=m), dt)
    assert errs["J_u"] < 1e-7 and errs["J_dt"] < 1e-7
    assert errs["H_uu"] < 1e-5 and errs["H_udt"] < 1e-5 and errs["H_dtdt"] < 1e-5


def test_oracle_vs_mpmath_expm():
    mp = pytest.importorskip("mpmath")
    rng = np.random.default_rng(11)
    A = rng.normal(size=(4, 4)); A = A - A.T
    from scipy.linalg import expm
    ref = np.array(mp.expm(mp.matrix(A.tolist())).tolist(), dtype=float)
    assert np.abs(expm(A) - ref).max() < 5e-15


# ---- the C port (same matrix program as the CUDA library) against the independent oracle ---------
CASES = [("cfg1_drift", 6, None), ("cfg1", 5, None), ("cfg2", 5, None), ("cfg3", 3, None),
         ("cfg3_ket", 4, None), ("cfg4", 4, 3)]


@pytest.mark.parametrize("name,N,S", CASES)
def test_port_matches_oracle(name, N, S, built_libs):
    w = workloads.make(name, N=N, S=S)
    OL = oracle_layout(w.layout)
    d0, j0, h0 = O.evaluate(OL, w.G_drift, w.G_drives, w.Z, w.mu)
    d1, j1, h1 = P.evaluate(OL, w.G_drift, w.G_drives, w.Z, w.mu)
    assert rel_err(d1, d0) < 1e-13 and rel_err(j1, j0) < 1e-13 and rel_err(h1, h0) < 1e-12


def test_port_many_squarings_and_range_error(built_libs):
    w = workloads.make("cfg2", N=4)
    L = w.layout
    Z = w.Z.reshape(L.N, L.z_dim).copy()
    Z[:, L.dt_off] = 9.0          # ||dt G||_1 ~ 10: six squarings
    OL = oracle_layout(L)
    d0, j0, h0 = O.evaluate(OL, w.G_drift, w.G_drives, Z.reshape(-1), w.mu)
    d1, j1, h1 = P.evaluate(OL, w.G_drift, w.G_drives, Z.reshape(-1), w.mu)
    assert rel_err(d1, d0) < 1e-12 and rel_err(j1, j0) < 1e-12 and rel_err(h1, h0) < 1e-11
    Z[1, L.dt_off] = np.nan
    with pytest.raises(FloatingPointError):
        P.evaluate(OL, w.G_drift, w.G_drives, Z.reshape(-1), w.mu)


def test_port_d128_spot_check(built_libs):
    """cfg5 (d = 128): one unit of the port against scipy on E, one Jacobian panel and one
    second-order Hessian scalar (the full block-triangular oracle would take minutes)."""
    from scipy.linalg import expm, expm_frechet
    w = workloads.make("cfg5", N=3)
    L = w.layout
    OL = oracle_layout(L)
    _, ja, ha = P.evaluate_arenas(OL, w.G_drift, w.G_drives, w.Z, w.mu, k_range=(0, 1))
    X, Xn, M, u, dt = OL.unit_inputs(w.Z, w.mu, 0, 0)
    G = w.G_drift[0] + np.tensordot(u, w.G_drives[0], axes=1)
    d, dc = L.d, L.d * L.c
    assert np.abs(ja[0, :d * d].reshape(d, d).T - expm(dt * G)).max() < 1e-13
    Lj = expm_frechet(dt * G, dt * w.G_drives[0][5], compute_expm=False)
    assert rel_err(ja[0, d * d + 5 * dc:d * d + 6 * dc].reshape(L.c, d).T, -(Lj @ X)) < 1e-12
    Q = O.second_frechet(dt * G, dt * w.G_drives[0][2], dt * w.G_drives[0][9])
    assert abs(ha[0, (L.m + 1) * dc + 2 * L.m + 9] + np.sum((M @ X.T) * Q)) < 1e-11 * np.abs(M @ X.T).sum()


# ---- committed golden vectors (frozen scipy-oracle outputs; tests/golden/make_golden.py) ----------
@pytest.mark.parametrize("name,N,S", [("cfg1_drift", 5, None), ("cfg2", 4, None), ("cfg3_ket", 4, None),
                                      ("cfg3", 3, None), ("cfg4", 4, 3)])
def test_golden_vectors(name, N, S, built_libs):
    g = np.load(os.path.join(GOLD, "oracle_small.npz"))
    w = workloads.make(name, N=N, S=S, seed=7)
    assert np.array_equal(w.Z, g[f"{name}.Z"]) and np.array_equal(w.mu, g[f"{name}.mu"])
    OL = oracle_layout(w.layout)
    r, c = OL.jacobian_structure()
    assert np.array_equal(r, g[f"{name}.jac_rows"]) and np.array_equal(c, g[f"{name}.jac_cols"])
    r, c = OL.hessian_structure()
    assert np.array_equal(r, g[f"{name}.hess_rows"]) and np.array_equal(c, g[f"{name}.hess_cols"])
    d1, j1, h1 = P.evaluate(OL, w.G_drift, w.G_drives, w.Z, w.mu)
    assert rel_err(d1, g[f"{name}.delta"]) < 1e-13
    assert rel_err(j1, g[f"{name}.jac"]) < 1e-13
    assert rel_err(h1, g[f"{name}.hess"]) < 1e-12


def test_port_coo_delivery_matches_arena_route(built_libs):
    """The timed CPU baseline (qc_port_eval_coo: COO value vectors assembled inside the C port)
    gives exactly what the arena route + numpy indexing gives."""
    for name, N, S in (("cfg2", 5, None), ("cfg4", 4, 3), ("cfg3_ket", 4, None)):
        w = workloads.make(name, N=N, S=S)
        OL = oracle_layout(w.layout)
        a = P.evaluate(OL, w.G_drift, w.G_drives, w.Z, w.mu)
        b = P.evaluate_coo(OL, w.G_drift, w.G_drives, w.Z, w.mu)
        assert all(np.array_equal(x, y) for x, y in zip(a, b))


def test_aux_oracle_definitions():
    """Derivative integrator and leakage restatements: definitions from the reference tree
    (spline_pulse_problem.jl:176 'du[k] = (u[k+1] - u[k]) / dt'; quantum_objectives.jl:298;
    quantum_constraints.jl:163) and finite-difference consistency."""
    rng = np.random.default_rng(0)
    N, zd, dt_off = 6, 9, 2
    Z = rng.normal(size=N * zd)
    pairs = [(3, 5, 2), (5, 7, 2)]
    mu = rng.normal(size=(N - 1) * 4)
    d0, j0, h0 = O.derivative_integrator(Z, N, zd, dt_off, pairs, mu)
    Zm = Z.reshape(N, zd)
    k, r = 2, 1
    assert abs(d0[k, r] - (Zm[k + 1, 4] - Zm[k, 4] - Zm[k, dt_off] * Zm[k, 6])) < 1e-15
    lag = lambda z: float(np.sum(mu * O.derivative_integrator(z, N, zd, dt_off, pairs)[0].reshape(-1)))
    h = 1e-2     # mu.delta is bilinear in (dx, dt): the mixed difference is exact up to rounding
    ei = np.zeros(N * zd); ei[k * zd + 6] = h          # dx component
    ej = np.zeros(N * zd); ej[k * zd + dt_off] = h     # dt
    fd = (lag(Z + ei + ej) - lag(Z + ei - ej) - lag(Z - ei + ej) + lag(Z - ei - ej)) / (4 * h * h)
    assert abs(fd - h0[k, r]) < 1e-9
    obj, grad, con, cjac, chess = O.leakage(Z, N, zd, 3, [0, 2, 3], Qs=np.full(N, 2.0), value=0.1, mu=np.ones(N * 3))
    assert abs(obj - 2.0 * np.sum(Zm[:, [3, 5, 6]] ** 2) / 3) < 1e-13
    assert np.allclose(con, Zm[:, [3, 5, 6]] ** 2 - 0.1) and np.allclose(cjac, 2 * Zm[:, [3, 5, 6]])


# ---- cfg5 goldens: every block of two full d = 128 units from the scipy oracle (make_golden_cfg5.py) ----
def cfg5_golden():
    """(workload, Z, mu, [(delta, jac, hess) per unit]) from tests/golden/oracle_cfg5.npz."""
    g = np.load(os.path.join(GOLD, "oracle_cfg5.npz"))
    w = workloads.make("cfg5", N=3, seed=11)
    OL = oracle_layout(w.layout)
    units = []
    for k in range(w.layout.K):
        ev = {key[len(f"u{k}."):]: g[key] for key in g.files if key.startswith(f"u{k}.")}
        units.append(O.assemble_unit(OL, 0, ev))
    return w, OL, g, units


def test_cfg5_golden_inputs_and_port(built_libs):
    """The committed cfg5 goldens belong to the seeded workload (same mu, same Z apart from the two
    scaled timesteps), cover s = 0 and s = 3, and the C port reproduces every block of both units."""
    w, OL, g, units = cfg5_golden()
    L = w.layout
    assert np.array_equal(w.mu, g["mu"])
    Zg, Zw = g["Z"].reshape(L.N, L.z_dim), w.Z.reshape(L.N, L.z_dim)
    mask = np.ones(L.z_dim, dtype=bool)
    mask[L.dt_off] = False
    assert np.array_equal(Zg[:, mask], Zw[:, mask])
    assert O_squarings(float(g["u0.norm1"])) == 0 and O_squarings(float(g["u1.norm1"])) == 3
    d1, j1, h1 = P.evaluate(OL, w.G_drift, w.G_drives, g["Z"], g["mu"])
    nj, nh, xd = L.jac_nnz_unit, L.hess_nnz_unit, L.x_dim
    for k, (d0, j0, h0) in enumerate(units):
        assert rel_err(d1[k * xd:(k + 1) * xd], d0) < 1e-12, k
        assert rel_err(j1[k * nj:(k + 1) * nj], j0) < 1e-11, k
        assert rel_err(h1[k * nh:(k + 1) * nh], h0) < 1e-10, k


def O_squarings(norm1, theta=0.5):
    s, t = 0, theta
    while norm1 > t:
        t *= 2.0
        s += 1
    return s


# ---- SURVEY 8f-2 / a3 restatements ----------------------------------------------------------------
def test_lindblad_generator_matches_the_master_equation():
    rng = np.random.default_rng(40)
    n = 3
    H = rng.normal(size=(n, n)) + 1j * rng.normal(size=(n, n)); H = H + H.conj().T
    Hj = rng.normal(size=(n, n)) + 1j * rng.normal(size=(n, n)); Hj = Hj + Hj.conj().T
    Ls = [rng.normal(size=(n, n)) + 1j * rng.normal(size=(n, n)) for _ in range(2)]
    G0, Gd = O.lindblad_generators(H, [Hj], Ls)
    r = rng.normal(size=(n, n)) + 1j * rng.normal(size=(n, n)); r = r @ r.conj().T; r /= np.trace(r)
    u = 0.37
    Hu = H + u * Hj
    rhs = -1j * (Hu @ r - r @ Hu) + sum(L @ r @ L.conj().T - 0.5 * (L.conj().T @ L @ r + r @ L.conj().T @ L) for L in Ls)
    got = (G0 + u * Gd[0]) @ O.density_to_iso_vec(r)
    assert np.abs(got - O.density_to_iso_vec(rhs)).max() < 1e-13
    assert abs(np.trace(O.iso_vec_to_density(got))) < 1e-13                 # trace preserving
    assert np.abs(G0 + G0.T).max() > 1e-2                                   # not skew-symmetric
    # without dissipation it reduces to the closed-system generator acting on both sides
    G0c, _ = O.lindblad_generators(H, [Hj], [])
    assert np.abs(G0c + G0c.T).max() < 1e-13


def test_var_G_is_the_parameter_derivative_of_the_flow():
    """d/da exp(t (G + a G_v)) x at a = 0 equals the variation block of exp(t var_G(G, [G_v])) [x; 0]."""
    from scipy.linalg import expm
    rng = np.random.default_rng(41)
    d = 4
    G = rng.normal(size=(d, d)); G = G - G.T
    Gv = rng.normal(size=(d, d)); Gv = Gv - Gv.T
    x = rng.normal(size=d)
    big = expm(0.7 * O.var_G(G, [Gv])) @ np.concatenate([x, np.zeros(d)])
    h = 1e-6
    fd = (expm(0.7 * (G + h * Gv)) @ x - expm(0.7 * (G - h * Gv)) @ x) / (2 * h)
    assert np.abs(big[:d] - expm(0.7 * G) @ x).max() < 1e-13 and np.abs(big[d:] - fd).max() < 1e-8


def test_free_phase_and_sensitivity_derivatives_by_finite_differences():
    rng = np.random.default_rng(42)
    n, nt = 3, 2
    Ug = np.linalg.qr(rng.normal(size=(n, n)) + 1j * rng.normal(size=(n, n)))[0]
    P = rng.integers(-1, 2, size=(n, nt)).astype(float)
    z = rng.normal(size=2 * n * n + nt)
    for side in (0, 1):
        F, g, H = O.free_phase_vgh(z, Ug, P, side)
        assert abs(F - O.free_phase_fidelity(z, Ug, P, side)) < 1e-14
        h = 1e-5
        gfd = np.array([(O.free_phase_fidelity(z + h * e, Ug, P, side) - O.free_phase_fidelity(z - h * e, Ug, P, side)) / (2 * h) for e in np.eye(len(z))])
        Hfd = np.array([(O.free_phase_vgh(z + h * e, Ug, P, side)[1] - O.free_phase_vgh(z - h * e, Ug, P, side)[1]) / (2 * h) for e in np.eye(len(z))])
        assert np.abs(g - gfd).max() < 1e-9 and np.abs(H - Hfd).max() < 1e-9
    N, zd = 4, 11
    Z = rng.normal(size=N * zd)
    J, g, Hs = O.sensitivity_objective(Z, N, zd, 1, 8, [2, 4], [1.5, 0.5], 1.3)
    h = 1e-6
    for i, t in enumerate([2, 4]):
        e = np.zeros_like(Z); e[(t - 1) * zd + 1 + 3] = h
        fd = (O.sensitivity_objective(Z + e, N, zd, 1, 8, [2, 4], [1.5, 0.5], 1.3)[0] - O.sensitivity_objective(Z - e, N, zd, 1, 8, [2, 4], [1.5, 0.5], 1.3)[0]) / (2 * h)
        assert abs(fd - g[i, 3]) < 1e-6 * max(1, abs(fd))
        gp = O.sensitivity_objective(Z + e, N, zd, 1, 8, [2, 4], [1.5, 0.5], 1.3)[1][i]
        gm = O.sensitivity_objective(Z - e, N, zd, 1, 8, [2, 4], [1.5, 0.5], 1.3)[1][i]
        assert np.abs((gp - gm) / (2 * h) - Hs[i][:, 3]).max() < 1e-5
    psi = rng.normal(size=n) + 1j * rng.normal(size=n); psi /= np.linalg.norm(psi)
    assert abs(O.density_pure_state_fidelity(O.density_to_iso_vec(np.outer(psi, psi.conj())), psi) - 1) < 1e-14
