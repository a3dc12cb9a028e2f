"""GPU tests of the terminal fidelity objectives / constraints against the oracle and the
reference's own known answers (src/quantum_objectives.jl:315-394)."""
import numpy as np
import pytest

import qcb200
from qcb200 import Fidelity, piccolo, workloads, isomorphisms as iso
import qc_oracle as O
from conftest import rel_err

pytestmark = pytest.mark.gpu


def test_coherent_ket_objective_reference_testitem(built_libs):
    N, ket_dim = 10, 4
    rng = np.random.default_rng(0)
    nrm = lambda a: a / np.linalg.norm(a)
    comps = {"psi1": nrm(rng.normal(size=(ket_dim, N))), "psi2": nrm(rng.normal(size=(ket_dim, N))),
             "u": rng.normal(size=(1, N)), "dt": np.full((1, N), 0.1)}
    traj = piccolo.NamedTrajectory(comps, timestep="dt")
    psi0, psi1 = np.array([1, 0], dtype=complex), np.array([0, 1], dtype=complex)
    goals = [psi1, psi0]
    obj = piccolo.CoherentKetInfidelityObjective(goals, ["psi1", "psi2"], traj, Q=100.0)
    J = piccolo.objective_value(obj, traj)
    assert isinstance(J, float) and 0.0 <= J <= 100.0
    grad = np.zeros(traj.dim * traj.N + traj.global_dim)
    piccolo.gradient_(grad, obj, traj)
    assert not np.all(grad == 0)
    assert np.all(grad[:(N - 1) * traj.dim] == 0)                 # terminal objective: knot N only
    perfect = dict(comps, psi1=np.tile(iso.ket_to_iso(psi1)[:, None], (1, N)), psi2=np.tile(iso.ket_to_iso(psi0)[:, None], (1, N)))
    assert piccolo.objective_value(obj, piccolo.NamedTrajectory(perfect, timestep="dt")) < 1e-10
    phase = dict(comps, psi1=np.tile(iso.ket_to_iso(psi1)[:, None], (1, N)), psi2=np.tile(iso.ket_to_iso(-psi0)[:, None], (1, N)))
    assert piccolo.objective_value(obj, piccolo.NamedTrajectory(phase, timestep="dt")) > 50.0
    with pytest.raises(AssertionError):
        piccolo.CoherentKetInfidelityObjective(goals, ["psi1"], traj)


@pytest.mark.parametrize("n", [2, 3, 9])
def test_unitary_fidelity_value_gradient_hessian(n, built_libs):
    rng = np.random.default_rng(n)
    Ug = workloads.haar_unitary(n, rng)
    x = rng.normal(size=2 * n * n)
    f = Fidelity(Fidelity.UNITARY, Ug)
    F, g, H = f.eval(x, want_hess=True)
    F0, g0, H0 = O.fidelity_vgh(O.KIND_UNITARY, x, Ug)
    assert abs(F - O.unitary_fidelity(x, Ug)) < 1e-12 * max(1, abs(F0))
    assert rel_err(g, g0) < 1e-12 and rel_err(H, H0) < 1e-12
    assert abs(f.eval(iso.operator_to_iso_vec(Ug), want_grad=False)[0] - 1.0) < 1e-13
    Gf, wf = f.hessian_factor()                    # the same Hessian without the dense x_dim^2 block
    assert Gf.shape == (2, 2 * n * n) and rel_err(2 * wf * Gf.T @ Gf, H0) < 1e-12


def test_ket_and_subspace_fidelity(built_libs):
    rng = np.random.default_rng(4)
    pg = rng.normal(size=3) + 1j * rng.normal(size=3); pg /= np.linalg.norm(pg)
    x = rng.normal(size=6)
    F, g, H = Fidelity(Fidelity.KET, pg).eval(x, want_hess=True)
    F0, g0, H0 = O.fidelity_vgh(O.KIND_KET, x, pg)
    assert abs(F - O.ket_fidelity(x, pg)) < 1e-13 and rel_err(g, g0) < 1e-12 and rel_err(H, H0) < 1e-12
    Ug = workloads.haar_unitary(2, rng)
    x = rng.normal(size=18)
    F, g, H = Fidelity(Fidelity.UNITARY_SUBSPACE, Ug, n=3, subspace=[0, 1]).eval(x, want_hess=True)
    F0, g0, H0 = O.fidelity_vgh(O.KIND_UNITARY_SUBSPACE, x, Ug, n_full=3, subspace=[0, 1])
    assert abs(F - O.unitary_subspace_fidelity(x, Ug, [0, 1])) < 1e-13
    assert rel_err(g, g0) < 1e-12 and rel_err(H, H0) < 1e-12


def test_final_fidelity_constraints(built_libs):
    rng = np.random.default_rng(6)
    N = 6
    U = np.stack([iso.operator_to_iso_vec(workloads.haar_unitary(2, rng)) for _ in range(N)], axis=1)
    traj = piccolo.NamedTrajectory({"U_iso": U, "dt": np.full((1, N), 0.1), "u": np.zeros((2, N))}, timestep="dt")
    con = piccolo.FinalUnitaryFidelityConstraint(workloads.H_GATE, "U_iso", 0.99, traj)
    assert con.equality is False and con.times == [N]            # src/quantum_constraints.jl:114-120
    val = con.value(traj)
    F0 = O.unitary_fidelity(U[:, -1], workloads.H_GATE)
    assert val.shape == (1,) and abs(val[0] - (0.99 - F0)) < 1e-13
    cols, vals = con.jacobian(traj)
    assert len(cols) == 8 and cols.min() == (N - 1) * traj.dim + 1   # dense 1 x x_dim row at knot N
    _, g0, H0 = O.fidelity_vgh(O.KIND_UNITARY, U[:, -1], workloads.H_GATE)
    assert rel_err(vals, -g0) < 1e-12
    assert rel_err(con.hessian_of_lagrangian(traj, 2.5), -2.5 * H0) < 1e-12
    obj = piccolo.UnitaryInfidelityObjective(workloads.H_GATE, "U_iso", traj, Q=100.0)
    assert abs(piccolo.objective_value(obj, traj) - 100 * abs(1 - F0)) < 1e-11
    kcon = piccolo.FinalKetFidelityConstraint(np.array([0, 1], dtype=complex), "u", 0.9, traj)   # 2-dim "ket": wrong size
    with pytest.raises(Exception):
        kcon.value(traj)
