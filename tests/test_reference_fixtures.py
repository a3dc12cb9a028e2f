"""Parity pinned against the reference's OWN outputs.

The reference ships solved trajectories (docs/data/*.jld2) and prints, in its documentation, the Ipopt
summary of the runs that produced them.  "Constraint violation" there is max|delta| over the dynamics
(and derivative) constraints at the saved iterate, "Objective" the template's objective on it.  Both
were computed by the real Julia evaluator (DirectTrajOpt + PiccoloQuantumObjects), so reproducing them
pins -- against the reference itself, to ~1e-11 relative --
  * the bilinear defect  x_{k+1} - exp(dt G(u)) x_k  with the knot layout, iso-vec layout and generator
    convention (3 runs; one of them to the last printed digit),
  * the (subspace) unitary infidelity objective, QuadraticRegularizer, MinimumTimeObjective and
    LeakageObjective (3 logged objective values to <= 1e-14 absolute).
The other fixtures have no log of their own run in the docs; they must still be solved trajectories
(max|delta| at solver tolerance), which any sign / layout / 2 pi mistake would break by 10 orders.

CPU tests check the oracle, GPU tests the CUDA path through the C ABI on the same data."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from conftest import rel_err  # noqa: E402
import qc_oracle as O  # noqa: E402
import reference_cases as RC  # noqa: E402

ALL = sorted(RC.CASES)
LOGGED = ["multilevel_transmon_example", "multilevel_transmon_example_leakage", "two_qubit_gates_min_time"]
# max|delta| of the fixtures without a log of their own (measured once with the oracle; solver tolerance)
SOLVED_BOUND = 1e-4
PIN = 1e-10           # relative agreement demanded with the printed 17-digit values


def oracle_layout(c):
    return O.Layout(c.d, c.c, c.m, c.N, c.S, c.dim, c.x_off, c.u_off, c.dt_off, 1)


def oracle_defect(c):
    return O.evaluate(oracle_layout(c), c.G_drift, c.G_drives, c.Zvec, None, want_jac=False, want_hess=False)[0]


def reference_objective(c, fidelity_value, leakage_value=0.0, D=0.0):
    reg = O.quadratic_regularizer(c.Zvec, c.N, c.dim, c.dt_off, c.reg_terms(1e-2), D)[0]
    return 100.0 * abs(1.0 - fidelity_value) + reg + leakage_value


def logged_objective_parts(c):
    """(fidelity kind, goal, subspace, leakage cost, D) of the three logged runs."""
    if c.name.startswith("multilevel_transmon"):
        # EmbeddedOperator(:X, sys) on levels {1,2} of 5 (multilevel_transmon.jl:54); the leakage run
        # adds LeakageObjective with leakage_cost = 1e-2 (:176-182)
        return "subspace", RC.X, [0, 1], (1e-2 if "leakage" in c.name else 0.0), 0.0
    # UnitaryMinimumTimeProblem keeps the smooth-pulse objective and adds D = 100 (two_qubit_gates.jl:203)
    return "full", c.goal_operator(), None, 0.0, 100.0


def oracle_fidelity(c, kind, goal, subspace):
    a, b = c.comp[c.states[0]]
    xN = c.data[a:b, -1]
    if kind == "subspace":
        return O.unitary_subspace_fidelity(xN, goal, subspace)      # src/quantum_objectives.jl:182-191
    return O.unitary_fidelity(xN, goal)                             # :173-180


# ------------------------------------------------------------------ CPU: the oracle ------------------
def test_fixture_file_is_complete():
    g = np.load(os.path.join(RC.GOLD, "reference_trajectories.npz"))
    assert sorted(str(n) for n in g["names"]) == ALL


@pytest.mark.parametrize("name", ALL)
def test_oracle_defect_on_reference_solutions(name):
    c = RC.Case(name)
    delta = oracle_defect(c)
    assert delta.size == c.S * (c.N - 1) * c.d * c.c
    worst = float(np.abs(delta).max())
    # derivative integrators (du, ddu rows) belong to the same constraint vector in the reference
    dd = O.derivative_integrator(c.Zvec, c.N, c.dim, c.dt_off, c.derivative_pairs())[0]
    worst = max(worst, float(np.abs(dd).max()))
    assert np.abs(O.timesteps_all_equal(c.Zvec, c.N, c.dim, c.dt_off)).max() == 0.0
    if c.log:
        assert abs(worst / c.log["constraint_violation"] - 1.0) < PIN, (worst, c.log)
    else:
        assert worst < SOLVED_BOUND, worst


@pytest.mark.parametrize("name", LOGGED)
def test_oracle_objective_on_reference_solutions(name):
    c = RC.Case(name)
    kind, goal, sub, leak_cost, D = logged_objective_parts(c)
    F = oracle_fidelity(c, kind, goal, sub)
    leak = 0.0
    if leak_cost:
        idx = O.leakage_indices(sub, c.n)
        assert len(idx) == 24
        leak = O.leakage(c.Zvec, c.N, c.dim, c.x_off[0], idx, Qs=np.full(c.N, leak_cost))[0]
    J = reference_objective(c, F, leak, D)
    assert abs(J - c.log["objective"]) <= 1e-13 * max(1.0, abs(c.log["objective"])), (J, c.log)


@pytest.mark.skipif(not os.path.isdir("/root/reference/docs/data"), reason="reference tree not present on this box")
def test_jld2_reader_matches_committed_fixtures():
    from qcb200 import jld2
    from qcb200.piccolo import NamedTrajectory
    g = np.load(os.path.join(RC.GOLD, "reference_trajectories.npz"))
    for name in ALL:
        path = f"/root/reference/docs/data/{name}_89ee72.jld2"
        t = jld2.named_trajectory(jld2.load(path)["traj"])
        assert np.array_equal(t["data"], g[f"{name}.data"])
        traj = NamedTrajectory.from_jld2(path)
        assert traj.N == t["N"] and traj.dim == t["dim"] and traj.timestep == "Δt"
        assert np.array_equal(traj.datavec, np.ascontiguousarray(t["data"].T).reshape(-1))


def test_jld2_reader_rejects_other_files(tmp_path):
    from qcb200 import jld2
    p = tmp_path / "x.jld2"
    p.write_bytes(b"not a jld2 file")
    with pytest.raises(ValueError):
        jld2.load(str(p))


# ------------------------------------------------------------------ GPU: the CUDA path ---------------
def _evaluator(c):
    from qcb200 import Evaluator
    from qcb200.layout import KnotLayout
    L = KnotLayout(d=c.d, c=c.c, m=c.m, N=c.N, S=c.S, z_dim=c.dim, x_off=c.x_off, u_off=c.u_off, dt_off=c.dt_off, jac_mode=1)
    return Evaluator(L, c.G_drift, c.G_drives), L


@pytest.mark.gpu
@pytest.mark.parametrize("name", ALL)
def test_cuda_defect_on_reference_solutions(name, built_libs):
    from qcb200 import Aux
    c = RC.Case(name)
    ev, L = _evaluator(c)
    delta = ev.eval_defect(c.Zvec)
    # on a solved trajectory delta is a difference of O(1) numbers that cancels to ~1e-6 or less, so the
    # comparison is absolute, against the size of the states (1), not of the residual
    assert np.abs(delta - oracle_defect(c)).max() < 1e-14
    aux = Aux(c.N, c.dim, c.dt_off, derivatives=c.derivative_pairs())
    dd = aux.derivative_eval(c.Zvec)[0]
    worst = max(float(np.abs(delta).max()), float(np.abs(dd).max()))
    assert np.abs(aux.timesteps_equal(c.Zvec)[0]).max() == 0.0
    if c.log:
        assert abs(worst / c.log["constraint_violation"] - 1.0) < PIN, (worst, c.log)
    else:
        assert worst < SOLVED_BOUND
    # first derivatives of the pinned function: Jacobian against the oracle on the real trajectory
    mu = np.random.default_rng(3).standard_normal(c.S * (c.N - 1) * c.d * c.c)
    d1, j1, h1 = ev.eval_all(c.Zvec, mu)
    d0, j0, h0 = O.evaluate(oracle_layout(c), c.G_drift, c.G_drives, c.Zvec, mu)
    assert rel_err(j1, j0) < 1e-10 and rel_err(h1, h0) < 1e-10
    ev.close(); aux.close()


@pytest.mark.gpu
@pytest.mark.parametrize("name", LOGGED)
def test_cuda_objective_on_reference_solutions(name, built_libs):
    from qcb200 import Aux, Fidelity
    c = RC.Case(name)
    kind, goal, sub, leak_cost, D = logged_objective_parts(c)
    a, b = c.comp["Ũ⃗"]
    xN = np.ascontiguousarray(c.data[a:b, -1])
    if kind == "subspace":
        f = Fidelity(Fidelity.UNITARY_SUBSPACE, goal, n=c.n, subspace=sub)
    else:
        f = Fidelity(Fidelity.UNITARY, goal)
    F = f.eval(xN)[0]
    assert abs(F - oracle_fidelity(c, kind, goal, sub)) < 1e-13
    leak_idx = O.leakage_indices(sub, c.n) if leak_cost else []
    aux = Aux(c.N, c.dim, c.dt_off, leak_x_off=c.x_off[0], leak_idx=leak_idx)
    reg, grad, hess = aux.regularizer_eval(c.Zvec, c.reg_terms(1e-2), D)
    leak = aux.leakage_eval(c.Zvec, np.full(c.N, leak_cost))[0] if leak_cost else 0.0
    J = 100.0 * abs(1.0 - F) + reg + leak
    assert abs(J - c.log["objective"]) <= 1e-13 * max(1.0, abs(c.log["objective"])), (J, c.log)
    # derivatives of the regulariser against the oracle restatement
    J0, g0, H0 = O.quadratic_regularizer(c.Zvec, c.N, c.dim, c.dt_off, c.reg_terms(1e-2), D)
    gc, (hr, hc) = aux.regularizer_structure(c.reg_terms(1e-2))
    gdense = np.zeros(c.N * c.dim)
    np.add.at(gdense, gc - 1, grad)
    assert abs(reg - J0) <= 1e-14 * max(1.0, abs(J0)) and rel_err(gdense, g0.reshape(-1)) < 1e-13
    dense = np.zeros((c.N * c.dim, c.N * c.dim)) if c.N * c.dim <= 6000 else None
    if dense is not None:
        np.add.at(dense, (hr - 1, hc - 1), hess)
        full = np.zeros_like(dense)
        for t in range(c.N):
            full[t * c.dim:(t + 1) * c.dim, t * c.dim:(t + 1) * c.dim] = np.triu(H0[t])
        assert rel_err(dense, full) < 1e-13 and np.all(hr <= hc)
    aux.close()
