"""Host-side mirror of the reference's interface for the hot path (Python stand-in for the Julia
host; the Julia shim with the same shape is julia/QCB200.jl).

Names, argument meaning and error behaviour follow the reference:

* ``BilinearIntegrator(qtraj, N)``            src/quantum_integrators.jl:28-134
* ``evaluate_(delta, integrator, traj)``      DirectTrajOpt.evaluate! as called at
                                              src/problem_templates/smooth_pulse_problem.jl:512-515
* ``test_integrator(integrator, traj; atol)`` src/quantum_integrators.jl:194
* ``UnitaryInfidelityObjective`` etc.         src/quantum_objectives.jl:34-42,60-68,134-166,193-201
* ``objective_value`` / ``gradient_``         src/quantum_objectives.jl:348-354
* ``Final*FidelityConstraint``                src/quantum_constraints.jl:20-35,65-100,106-121

All numerical work goes through the CUDA library (api.Evaluator / api.Fidelity).
"""
from __future__ import annotations

from collections import OrderedDict
from typing import Dict, List, Optional, Sequence

import numpy as np

from .api import Aux, Evaluator, Fidelity, FreePhaseFidelity
from .isomorphisms import G_of_H, ket_to_iso, lindblad_generators, operator_to_iso_vec, var_G
from .layout import JAC_BLOCK_DIAG, JAC_DENSE_X, KnotLayout


# --------------------------------------------------------------------------------------- #
# Minimal stand-ins for PiccoloQuantumObjects / NamedTrajectories containers
# --------------------------------------------------------------------------------------- #
class QuantumSystem:
    """``QuantumSystem(H_drift, H_drives)`` with ``.G(u) = G_drift + sum_j u_j G_j``."""

    def __init__(self, H_drift, H_drives: Sequence):
        self.H_drift = np.asarray(H_drift, dtype=complex)
        self.H_drives = [np.asarray(h, dtype=complex) for h in H_drives]
        self.levels = self.H_drift.shape[0]
        self.n_drives = len(self.H_drives)
        self.G_drift = G_of_H(self.H_drift)
        self.G_drives = np.array([G_of_H(h) for h in self.H_drives]).reshape(self.n_drives, 2 * self.levels, 2 * self.levels)

    def G(self, u, t: float = 0.0):
        return self.G_drift + np.tensordot(np.asarray(u, dtype=float), self.G_drives, axes=1)


class OpenQuantumSystem:
    """``OpenQuantumSystem(H_drift, H_drives, dissipation_operators)`` with the Lindbladian generator
    ``.𝒢(u)`` on the density iso-vec (2 n^2 entries); here ``.G_drift`` / ``.G_drives`` hold its
    constant and per-drive parts (dense, not skew-symmetric)."""

    def __init__(self, H_drift, H_drives: Sequence, dissipation_operators: Sequence = ()):
        self.H_drift = np.asarray(H_drift, dtype=complex)
        self.H_drives = [np.asarray(h, dtype=complex) for h in H_drives]
        self.dissipation_operators = [np.asarray(L, dtype=complex) for L in dissipation_operators]
        self.levels = self.H_drift.shape[0]
        self.n_drives = len(self.H_drives)
        self.G_drift, self.G_drives = lindblad_generators(self.H_drift, self.H_drives, self.dissipation_operators)

    def G(self, u, t: float = 0.0):
        return self.G_drift + np.tensordot(np.asarray(u, dtype=float), self.G_drives, axes=1)


class VariationalQuantumSystem:
    """``VariationalQuantumSystem(H_drives, H_vars)`` (docs/literate/man/unitary_problem_templates.jl:161):
    ``G(u)`` of the nominal system plus one generator per variation, each linear in the drives:
    ``G_vars[v](u) = Gv_drift[v] + sum_j u_j Gv_drives[v][j]``."""

    def __init__(self, H_drift, H_drives: Sequence, H_vars: Sequence, H_var_drives: Optional[Sequence] = None):
        self.base = QuantumSystem(H_drift, H_drives)
        self.levels, self.n_drives = self.base.levels, self.base.n_drives
        self.Gv_drift = [G_of_H(h) for h in H_vars]
        zero = np.zeros((2 * self.levels, 2 * self.levels))
        self.Gv_drives = [[zero] * self.n_drives for _ in H_vars] if H_var_drives is None else \
            [[G_of_H(h) for h in hs] for hs in H_var_drives]

    def G(self, u, t: float = 0.0):
        return self.base.G(u)

    @property
    def G_vars(self):
        return [lambda u, v=v: self.Gv_drift[v] + np.tensordot(np.asarray(u, dtype=float), np.array(self.Gv_drives[v]), axes=1)
                for v in range(len(self.Gv_drift))]


class NamedTrajectory:
    """Knot-major container: components are (dim_i, N) arrays stacked in insertion order;
    ``datavec`` is the flat ``Z`` of length ``dim * N`` (src/quantum_objectives.jl:353); global
    variables (``global_data``) follow the knots, ``global_dim`` entries (:353)."""

    def __init__(self, components: Dict[str, np.ndarray], timestep: str = "dt", controls=("u",),
                 goal: Optional[dict] = None, global_data: Optional[Dict[str, np.ndarray]] = None):
        self.names = list(components)
        comps = OrderedDict((k, np.atleast_2d(np.asarray(v, dtype=float))) for k, v in components.items())
        self.N = next(iter(comps.values())).shape[1]
        self.dims = {k: v.shape[0] for k, v in comps.items()}
        self.components = {}
        off = 0
        for k, v in comps.items():
            assert v.shape[1] == self.N, f"component {k} has {v.shape[1]} knots, expected {self.N}"
            self.components[k] = np.arange(off, off + v.shape[0])
            off += v.shape[0]
        self.dim = off
        self.data = np.vstack(list(comps.values()))          # (dim, N)
        self.timestep = timestep
        self.controls = tuple(controls)
        self.goal = goal or {}
        self.global_data = OrderedDict((k, np.atleast_1d(np.asarray(v, dtype=float))) for k, v in (global_data or {}).items())
        self.global_dims = {k: len(v) for k, v in self.global_data.items()}
        self.global_components, goff = {}, 0
        for k, v in self.global_data.items():
            self.global_components[k] = np.arange(goff, goff + len(v))
            goff += len(v)
        self.global_dim = goff

    @classmethod
    def from_jld2(cls, path: str, key: str = "traj") -> "NamedTrajectory":
        """A trajectory saved by the reference (`save(path, "traj", prob.trajectory)`; the solved
        examples under docs/data), decoded by quantumcollocation.jl_b200/jld2.py."""
        from . import jld2
        t = jld2.named_trajectory(jld2.load(path)[key])
        order = sorted(t["components"], key=lambda k: t["components"][k][0])
        traj = cls({k: t["data"][t["components"][k], :] for k in order}, timestep=t["timestep"],
                   controls=tuple(t["raw"].get("control_names", {}).values()),
                   goal={k: np.asarray(v) for k, v in t["raw"].get("goal", {}).items()})
        traj.bounds = t["raw"].get("bounds", {})
        traj.initial = {k: np.asarray(v) for k, v in t["raw"].get("initial", {}).items()}
        traj.final = {k: np.asarray(v) for k, v in t["raw"].get("final", {}).items()}
        return traj

    @property
    def datavec(self) -> np.ndarray:
        return np.ascontiguousarray(self.data.T).reshape(-1)

    def __getitem__(self, name):
        return self.data[self.components[name], :]

    def update(self, name, value):
        self.data[self.components[name], :] = value


class UnitaryTrajectory:
    def __init__(self, system: QuantumSystem, U_goal, state_name="U_iso", drive_name="u"):
        self.system, self.goal = system, np.asarray(U_goal, dtype=complex)
        self.state_name, self.drive_name = state_name, drive_name
        self.columns = system.levels


class KetTrajectory:
    def __init__(self, system: QuantumSystem, psi_init, psi_goal, state_name="psi_iso", drive_name="u"):
        self.system = system
        self.init, self.goal = np.asarray(psi_init, dtype=complex), np.asarray(psi_goal, dtype=complex)
        self.state_name, self.drive_name = state_name, drive_name
        self.columns = 1


class DensityTrajectory:
    """``DensityTrajectory(sys::OpenQuantumSystem, rho_init, rho_goal)``: state = density iso-vec."""

    def __init__(self, system: OpenQuantumSystem, rho_init, rho_goal, state_name="rho_iso", drive_name="u"):
        self.system = system
        self.init, self.goal = np.asarray(rho_init, dtype=complex), np.asarray(rho_goal, dtype=complex)
        self.state_name, self.drive_name = state_name, drive_name
        self.columns = 1


class MultiKetTrajectory:
    def __init__(self, system: QuantumSystem, psi_inits, psi_goals, drive_name="u"):
        self.system = system
        self.inits = [np.asarray(p, dtype=complex) for p in psi_inits]
        self.goals = [np.asarray(p, dtype=complex) for p in psi_goals]
        self.state_names = [f"psi_iso{i + 1}" for i in range(len(self.inits))]
        self.drive_name = drive_name
        self.columns = 1


class SamplingTrajectory:
    """One state per system, shared controls (src/problem_templates/sampling_problem.jl:125-136)."""

    def __init__(self, base_trajectory, systems: Sequence[QuantumSystem]):
        self.base_trajectory = base_trajectory
        self.systems = list(systems)
        stem = base_trajectory.state_name
        self.state_names = [f"{stem}{i + 1}" for i in range(len(self.systems))]
        self.drive_name = base_trajectory.drive_name
        self.columns = base_trajectory.columns


def _layout_from_traj(traj: NamedTrajectory, state_names: Sequence[str], drive_name: str, d: int, c: int,
                      jac_mode: Optional[int]) -> KnotLayout:
    x_off = []
    for nm in state_names:
        idx = traj.components[nm]
        if len(idx) != d * c or not np.all(np.diff(idx) == 1):
            raise ValueError(f"state component {nm} must be {d * c} contiguous entries of a knot")
        x_off.append(int(idx[0]))
    u_idx = traj.components[drive_name]
    dt_idx = traj.components[traj.timestep]
    if jac_mode is None:
        jac_mode = JAC_BLOCK_DIAG if d * c > 512 else JAC_DENSE_X
    return KnotLayout(d=d, c=c, m=len(u_idx), N=traj.N, S=len(state_names), z_dim=traj.dim, x_off=x_off,
                      u_off=int(u_idx[0]), dt_off=int(dt_idx[0]), jac_mode=jac_mode)


class Integrator:
    """The drop-in for one ``DirectTrajOpt.BilinearIntegrator`` (or, fused, for the whole vector the
    reference builds for MultiKet / Sampling trajectories).  ``dim`` = x_dim * (N - 1) [* S]."""

    def __init__(self, layout: KnotLayout, G_drift, G_drives, device=0, k_begin=0, k_end=0, force_path=0):
        self.layout = layout
        self.evaluator = Evaluator(layout, G_drift, G_drives, device=device, k_begin=k_begin, k_end=k_end,
                                   force_path=force_path)
        self.dim = self.evaluator.defect_dim

    def evaluate(self, delta: np.ndarray, traj: NamedTrajectory) -> np.ndarray:
        if delta.shape != (self.dim,):
            raise ValueError(f"delta has shape {delta.shape}, expected ({self.dim},)")
        return self.evaluator.eval_defect(traj.datavec, delta)

    def jacobian_structure(self, row_base=0):
        return self.evaluator.jacobian_structure(row_base)

    def hessian_structure(self):
        return self.evaluator.hessian_structure()

    def jacobian_values(self, traj: NamedTrajectory, out=None):
        return self.evaluator.eval_jacobian(traj.datavec, out)

    def hessian_of_lagrangian_values(self, traj: NamedTrajectory, mu: np.ndarray, out=None):
        return self.evaluator.eval_hessian(traj.datavec, mu, out)


def BilinearIntegrator(qtraj, traj: NamedTrajectory, fused: bool = False, device: int = 0,
                       jac_mode: Optional[int] = None, force_path: int = 0):
    """Dispatch on the trajectory type exactly like src/quantum_integrators.jl:28-134.

    Unitary / Ket -> one ``Integrator``; MultiKet / Sampling -> a list with one integrator per
    state (the reference's return type) or, with ``fused=True``, a single integrator that
    evaluates every state of a knot in one launch (knot x sample batching).
    """
    kw = dict(device=device, force_path=force_path)
    if isinstance(qtraj, DensityTrajectory):
        # src/quantum_integrators.jl:52-56: BilinearIntegrator(sys.𝒢, :ρ⃗̃, :u, traj); the generator acts on
        # the whole density iso-vec (c = 1, d = 2 n^2)
        sys_ = qtraj.system
        d = 2 * sys_.levels ** 2
        L = _layout_from_traj(traj, [qtraj.state_name], qtraj.drive_name, d, 1, jac_mode)
        return Integrator(L, sys_.G_drift[None], sys_.G_drives[None], **kw)
    if isinstance(qtraj, (UnitaryTrajectory, KetTrajectory)):
        sys_ = qtraj.system
        d = 2 * sys_.levels
        L = _layout_from_traj(traj, [qtraj.state_name], qtraj.drive_name, d, qtraj.columns, jac_mode)
        return Integrator(L, sys_.G_drift[None], sys_.G_drives[None], **kw)
    if isinstance(qtraj, MultiKetTrajectory):
        systems = [qtraj.system] * len(qtraj.state_names)
        names, cols = qtraj.state_names, 1
    elif isinstance(qtraj, SamplingTrajectory):
        systems, names, cols = qtraj.systems, qtraj.state_names, qtraj.columns
    else:
        raise TypeError(f"no BilinearIntegrator for {type(qtraj).__name__}")
    d = 2 * systems[0].levels
    if fused:
        L = _layout_from_traj(traj, names, qtraj.drive_name, d, cols, jac_mode)
        return Integrator(L, np.array([s.G_drift for s in systems]), np.array([s.G_drives for s in systems]), **kw)
    out = []
    for s, nm in zip(systems, names):
        L = _layout_from_traj(traj, [nm], qtraj.drive_name, d, cols, jac_mode)
        out.append(Integrator(L, s.G_drift[None], s.G_drives[None], **kw))
    return out


def _variational_integrator(sys_: VariationalQuantumSystem, traj: NamedTrajectory, names: Sequence[str], u_name: str,
                            columns: int, scales: Sequence[float], device: int, force_path: int, jac_mode):
    """Shared body of the two variational integrators: generator var_G(G(u), [G_v(u)/scale]) on the
    row-stacked panel [X; X_v1; ...]; the state is gathered from the named components through the
    state index map of qcb_desc (x_index / row_index)."""
    d0, nb = 2 * sys_.levels, len(names)
    if len(scales) != nb - 1 or nb - 1 != len(sys_.Gv_drift):
        raise ValueError("one variation component and one scale per variational generator")
    d, c = nb * d0, columns
    G0 = var_G(sys_.base.G_drift, [g / sc for g, sc in zip(sys_.Gv_drift, scales)])
    Gd = np.array([var_G(sys_.base.G_drives[j], [sys_.Gv_drives[v][j] / scales[v] for v in range(nb - 1)])
                   for j in range(sys_.n_drives)]).reshape(sys_.n_drives, d, d)
    # panel entry (block b, row r, column q) = r + b*d0 + q*d  <->  component b, entry q*d0 + r
    x_index, row_index = np.zeros(d * c, dtype=np.int64), np.zeros(d * c, dtype=np.int64)
    for b, nm in enumerate(names):
        idx = traj.components[nm]
        if len(idx) != d0 * c:
            raise ValueError(f"state component {nm} must have {d0 * c} entries")
        for q in range(c):
            for r in range(d0):
                x_index[b * d0 + r + q * d] = idx[q * d0 + r]
                row_index[b * d0 + r + q * d] = b * d0 * c + q * d0 + r
    u_idx, dt_idx = traj.components[u_name], traj.components[traj.timestep]
    if jac_mode is None:
        jac_mode = JAC_BLOCK_DIAG if d * c > 512 else JAC_DENSE_X
    L = KnotLayout(d=d, c=c, m=len(u_idx), N=traj.N, S=1, z_dim=traj.dim, x_off=[0], u_off=int(u_idx[0]),
                   dt_off=int(dt_idx[0]), jac_mode=jac_mode, x_index=[x_index.tolist()], row_index=row_index.tolist())
    return Integrator(L, G0[None], Gd[None], device=device, force_path=force_path)


def VariationalKetIntegrator(sys_: VariationalQuantumSystem, traj: NamedTrajectory, psi_name: str,
                             variation_names: Sequence[str], u_name: str, scale: float = 1.0, device: int = 0,
                             force_path: int = 0, jac_mode: Optional[int] = None):
    """src/quantum_integrators.jl:140-152: state vcat(psi, variations...), G = var_G(sys.G(u), [G_v(u)/scale])."""
    return _variational_integrator(sys_, traj, [psi_name, *variation_names], u_name, 1,
                                   [scale] * len(variation_names), device, force_path, jac_mode)


def VariationalUnitaryIntegrator(sys_: VariationalQuantumSystem, traj: NamedTrajectory, U_name: str,
                                 variation_names: Sequence[str], u_name: str, scales: Optional[Sequence[float]] = None,
                                 device: int = 0, force_path: int = 0, jac_mode: Optional[int] = None):
    """src/quantum_integrators.jl:154-169: var_G(I (x) G0, [I (x) G_v / scale]) on vcat(U_iso, variations...)
    = var_G(G0, [G_v/scale]) on the row-stacked panels."""
    scales = [1.0] * len(variation_names) if scales is None else list(scales)
    return _variational_integrator(sys_, traj, [U_name, *variation_names], u_name, sys_.levels, scales, device,
                                   force_path, jac_mode)


def rollout(integrator: Integrator, traj: NamedTrajectory, x_init=None) -> np.ndarray:
    """States [S, N, x_dim] obtained by propagating ``x_init`` (default: the states at knot 1) with the
    trajectory's controls and timesteps -- what ``sync_trajectory!`` re-computes after a solve
    (src/quantum_control_problem.jl:137-150), for controls held constant over each interval."""
    return integrator.evaluator.rollout(traj.datavec, x_init)


def evaluate_(delta: np.ndarray, integrator: Integrator, traj: NamedTrajectory) -> np.ndarray:
    """``DirectTrajOpt.evaluate!(delta, integrator, traj)``."""
    return integrator.evaluate(delta, traj)


def test_integrator(integrator: Integrator, traj: NamedTrajectory, atol: float = 1e-3, h: float = 1e-6,
                    rng=None, max_columns: int = 24) -> Dict[str, float]:
    """Derivative self-consistency, the reference's own criterion (src/quantum_integrators.jl:194):
    Jacobian values vs central differences of the defect, Hessian-of-Lagrangian values vs central
    differences of mu^T Jacobian.  Every evaluation is a call into the CUDA library.  Checks a random
    subset of ``max_columns`` variables.  Raises AssertionError beyond ``atol``."""
    rng = rng or np.random.default_rng(0)
    ev, L = integrator.evaluator, integrator.layout
    Z = traj.datavec.copy()
    rows, cols = ev.jacobian_structure()
    jac = ev.eval_jacobian(Z)
    mu = rng.normal(size=L.S * L.K * L.x_dim)
    hr, hc = ev.hessian_structure()
    hess = ev.eval_hessian(Z, mu)
    # candidate variables: everything the integrator depends on
    used = np.unique(cols) - 1
    pick = rng.choice(used, size=min(max_columns, len(used)), replace=False)
    err_j = err_h = 0.0
    mu_slab = mu.reshape(L.S, L.K, L.x_dim)[:, ev.k_begin:ev.k_end].reshape(-1)
    for v in pick:
        Zp, Zm = Z.copy(), Z.copy()
        Zp[v] += h
        Zm[v] -= h
        fd = (ev.eval_defect(Zp) - ev.eval_defect(Zm)) / (2 * h)
        col = np.zeros(integrator.dim)
        sel = cols == v + 1
        np.add.at(col, rows[sel] - 1, jac[sel])
        err_j = max(err_j, float(np.abs(fd - col).max()))
        # d/dv (mu^T J)  vs  Hessian column v
        gp = np.zeros(len(Z)); gm = np.zeros(len(Z))
        np.add.at(gp, cols - 1, ev.eval_jacobian(Zp) * mu_slab[rows - 1])
        np.add.at(gm, cols - 1, ev.eval_jacobian(Zm) * mu_slab[rows - 1])
        fdh = (gp - gm) / (2 * h)
        colh = np.zeros(len(Z))
        a = hc == v + 1
        np.add.at(colh, hr[a] - 1, hess[a])
        b = (hr == v + 1) & (hc != v + 1)
        np.add.at(colh, hc[b] - 1, hess[b])
        err_h = max(err_h, float(np.abs(fdh - colh).max()))
    assert err_j < atol, f"Jacobian inconsistent with finite differences: {err_j:.3e}"
    assert err_h < atol, f"Hessian inconsistent with finite differences: {err_h:.3e}"
    return {"jacobian": err_j, "hessian": err_h}


test_integrator.__test__ = False   # not a pytest test


# --------------------------------------------------------------------------------------- #
# Objectives (src/quantum_objectives.jl) and constraints (src/quantum_constraints.jl)
# --------------------------------------------------------------------------------------- #
class TerminalObjective:
    """``Q * |1 - F(x_N)|`` on one or several (concatenated) terminal components."""

    def __init__(self, fidelity: Fidelity, names: Sequence[str], traj: NamedTrajectory, Q: float = 100.0):
        self.fidelity, self.names, self.Q = fidelity, list(names), float(Q)
        self.idx = np.concatenate([traj.components[n] for n in self.names])
        if len(self.idx) != fidelity.x_dim:
            raise ValueError("terminal components do not match the goal's dimension")

    def _x(self, traj):
        return np.ascontiguousarray(traj.data[self.idx, traj.N - 1])

    def __call__(self, traj):
        return objective_value(self, traj)


def objective_value(obj: TerminalObjective, traj: NamedTrajectory) -> float:
    F, _, _ = obj.fidelity.eval(obj._x(traj), want_grad=False)
    return obj.Q * abs(1.0 - F)


def gradient_(grad: np.ndarray, obj: TerminalObjective, traj: NamedTrajectory) -> np.ndarray:
    """``gradient!(grad, obj, traj)`` with ``grad = zeros(traj.dim * traj.N + traj.global_dim)``."""
    if grad.shape != (traj.dim * traj.N + traj.global_dim,):
        raise ValueError("gradient buffer has the wrong length")
    F, g, _ = obj.fidelity.eval(obj._x(traj))
    grad[:] = 0.0
    grad[(traj.N - 1) * traj.dim + obj.idx] = -np.sign(1.0 - F) * obj.Q * g
    return grad


def hessian_values(obj: TerminalObjective, traj: NamedTrajectory) -> np.ndarray:
    """Dense Hessian block of the objective on its terminal components."""
    F, _, H = obj.fidelity.eval(obj._x(traj), want_grad=False, want_hess=True)
    return -np.sign(1.0 - F) * obj.Q * H


def UnitaryInfidelityObjective(U_goal, name: str, traj: NamedTrajectory, Q: float = 100.0, subspace=None, device=0):
    """src/quantum_objectives.jl:193-201; with ``subspace`` the EmbeddedOperator variant (:182-191)."""
    U_goal = np.asarray(U_goal, dtype=complex)
    if subspace is None:
        f = Fidelity(Fidelity.UNITARY, U_goal, device=device)
    else:
        n = int(round(np.sqrt(traj.dims[name] / 2)))
        f = Fidelity(Fidelity.UNITARY_SUBSPACE, U_goal, n=n, subspace=subspace, device=device)
    return TerminalObjective(f, [name], traj, Q)


def KetInfidelityObjective(psi_goal, name: str, traj: NamedTrajectory, Q: float = 100.0, device=0):
    """src/quantum_objectives.jl:60-68 (explicit goal); :34-42 reads the goal from ``traj.goal``."""
    if psi_goal is None:
        g = np.asarray(traj.goal[name])
        psi_goal = g[:len(g) // 2] + 1j * g[len(g) // 2:]
    return TerminalObjective(Fidelity(Fidelity.KET, psi_goal, device=device), [name], traj, Q)


def CoherentKetInfidelityObjective(psi_goals, names: Sequence[str], traj: NamedTrajectory, Q: float = 100.0, device=0):
    """src/quantum_objectives.jl:134-166."""
    if len(psi_goals) != len(names):
        raise AssertionError("Number of names must match number of goals")
    return TerminalObjective(Fidelity(Fidelity.COHERENT_KET, np.array(psi_goals), device=device), names, traj, Q)


def DensityMatrixPureStateInfidelityObjective(name: str, psi_goal, traj: NamedTrajectory, Q: float = 100.0, device=0):
    """src/quantum_objectives.jl:241-249: Q * |1 - real(psi' rho psi)| on the terminal density iso-vec."""
    return TerminalObjective(Fidelity(Fidelity.DENSITY_PURE, np.asarray(psi_goal, dtype=complex), device=device), [name], traj, Q)


class FreePhaseTerminalObjective:
    """``Q * |1 - F(x_N, theta)|`` on z = [x_N; theta] (TerminalObjective with global components,
    src/quantum_objectives.jl:215)."""

    def __init__(self, fidelity: FreePhaseFidelity, name: str, theta_names: Sequence[str], traj: NamedTrajectory, Q: float):
        self.fidelity, self.name, self.theta_names, self.Q = fidelity, name, list(theta_names), float(Q)
        self.idx = traj.components[name]
        self.gidx = np.concatenate([traj.global_components[n] for n in self.theta_names])
        if len(self.gidx) != fidelity.n_theta or len(self.idx) != fidelity.x_dim:
            raise ValueError("component sizes do not match the goal / phase map")

    def _z(self, traj):
        theta = np.concatenate([traj.global_data[n] for n in self.theta_names])
        return np.ascontiguousarray(traj.data[self.idx, traj.N - 1]), theta

    def __call__(self, traj):
        F, _, _ = self.fidelity.eval(*self._z(traj), want_grad=False)
        return self.Q * abs(1.0 - F)

    def gradient(self, grad: np.ndarray, traj) -> np.ndarray:
        """grad = zeros(traj.dim * traj.N + traj.global_dim)  (src/quantum_objectives.jl:353)."""
        F, g, _ = self.fidelity.eval(*self._z(traj))
        grad[:] = 0.0
        sg = -np.sign(1.0 - F) * self.Q
        grad[(traj.N - 1) * traj.dim + self.idx] = sg * g[:self.fidelity.x_dim]
        grad[traj.N * traj.dim + self.gidx] = sg * g[self.fidelity.x_dim:]
        return grad

    def hessian_values(self, traj) -> np.ndarray:
        """Dense block over [x_N; theta]."""
        F, _, H = self.fidelity.eval(*self._z(traj), want_grad=False, want_hess=True)
        return -np.sign(1.0 - F) * self.Q * H


def UnitaryFreePhaseInfidelityObjective(U_goal, phase_map, name: str, theta_names, traj: NamedTrajectory, Q: float = 100.0,
                                        side: int = 0, device=0):
    """src/quantum_objectives.jl:203-226 for the goal family U_goal(theta) = diag(e^{i P theta}) U_goal
    (side 1: U_goal diag(...)); ``theta_names`` a name or a list of global components."""
    if isinstance(theta_names, str):
        theta_names = [theta_names]
    return FreePhaseTerminalObjective(FreePhaseFidelity(U_goal, phase_map, side, device), name, theta_names, traj, Q)


class UnitarySensitivityObjective:
    """``UnitarySensitivityObjective(name, traj, times; Qs, scale)`` (src/quantum_objectives.jl:263-279):
    sum_t Q_t * scale^4 * |tr(U_t' U_t)|^2 / n^2 on a variational state component; times 1-based."""

    def __init__(self, name: str, traj: NamedTrajectory, times: Sequence[int], Qs=None, scale: float = 1.0, device=0):
        idx = traj.components[name]
        self.x_off, self.x_dim = int(idx[0]), len(idx)
        self.times = np.asarray(times, dtype=np.int64)
        self.Qs = np.ones(len(self.times)) if Qs is None else np.asarray(Qs, dtype=float)
        self.scale = float(scale)
        self.aux = Aux(traj.N, traj.dim, int(traj.components[traj.timestep][0]), device=device)
        self.cols = ((self.times - 1)[:, None] * traj.dim + idx[None, :]).reshape(-1)

    def __call__(self, traj):
        return self.aux.sensitivity_eval(traj.datavec, self.x_off, self.x_dim, self.times, self.Qs, self.scale, False)[0]

    def gradient(self, grad: np.ndarray, traj) -> np.ndarray:
        grad[:] = 0.0
        g = self.aux.sensitivity_eval(traj.datavec, self.x_off, self.x_dim, self.times, self.Qs, self.scale, False)[1]
        np.add.at(grad, self.cols, g.reshape(-1))
        return grad

    def hessian_values(self, traj) -> np.ndarray:
        """[len(times), x_dim, x_dim] dense blocks."""
        return self.aux.sensitivity_eval(traj.datavec, self.x_off, self.x_dim, self.times, self.Qs, self.scale, True)[2]


class FinalFreePhaseFidelityConstraint:
    """theta-parametrised ``FinalUnitaryFidelityConstraint`` (src/quantum_constraints.jl:123-144):
    [final_fidelity - F(x_N, theta)] <= 0 (NonlinearGlobalKnotPointConstraint, times = [N])."""

    def __init__(self, fidelity: FreePhaseFidelity, name: str, theta_names: Sequence[str], final_fidelity: float,
                 traj: NamedTrajectory):
        self._obj = FreePhaseTerminalObjective(fidelity, name, theta_names, traj, 1.0)
        self.fidelity, self.final_fidelity = fidelity, float(final_fidelity)
        self.equality, self.times = False, [traj.N]

    def value(self, traj):
        F, _, _ = self.fidelity.eval(*self._obj._z(traj), want_grad=False)
        return np.array([self.final_fidelity - F])

    def jacobian(self, traj):
        """(cols [1-based into [Z; globals]], values) of the single row."""
        _, g, _ = self.fidelity.eval(*self._obj._z(traj))
        cols = np.concatenate([(traj.N - 1) * traj.dim + self._obj.idx, traj.N * traj.dim + self._obj.gidx]) + 1
        return cols, -g

    def hessian_of_lagrangian(self, traj, mu: float):
        _, _, H = self.fidelity.eval(*self._obj._z(traj), want_grad=False, want_hess=True)
        return -mu * H


class FinalFidelityConstraint:
    """``[final_fidelity - F(x_N)] <= 0`` at knot N (equality=false, times=[traj.N])."""

    def __init__(self, fidelity: Fidelity, names: Sequence[str], final_fidelity: float, traj: NamedTrajectory):
        self.fidelity, self.names, self.final_fidelity = fidelity, list(names), float(final_fidelity)
        self.idx = np.concatenate([traj.components[n] for n in self.names])
        self.equality, self.times = False, [traj.N]

    def value(self, traj):
        F, _, _ = self.fidelity.eval(np.ascontiguousarray(traj.data[self.idx, traj.N - 1]), want_grad=False)
        return np.array([self.final_fidelity - F])

    def jacobian(self, traj):
        """(cols [1-based into Z], values) of the single row."""
        _, g, _ = self.fidelity.eval(np.ascontiguousarray(traj.data[self.idx, traj.N - 1]))
        return (traj.N - 1) * traj.dim + self.idx + 1, -g

    def hessian_of_lagrangian(self, traj, mu: float):
        _, _, H = self.fidelity.eval(np.ascontiguousarray(traj.data[self.idx, traj.N - 1]), want_grad=False, want_hess=True)
        return -mu * H


def FinalUnitaryFidelityConstraint(U_goal, name, final_fidelity, traj, subspace=None, device=0,
                                   theta_names=None, phase_map=None, side: int = 0):
    """src/quantum_constraints.jl:106-121; with ``theta_names`` + ``phase_map`` the theta-parametrised
    method (:123-144) for U_goal(theta) = diag(e^{i P theta}) U_goal."""
    U_goal = np.asarray(U_goal, dtype=complex)
    if theta_names is not None:
        if isinstance(theta_names, str):
            theta_names = [theta_names]
        return FinalFreePhaseFidelityConstraint(FreePhaseFidelity(U_goal, phase_map, side, device), name, theta_names,
                                                final_fidelity, traj)
    if subspace is None:
        f = Fidelity(Fidelity.UNITARY, U_goal, device=device)
    else:
        n = int(round(np.sqrt(traj.dims[name] / 2)))
        f = Fidelity(Fidelity.UNITARY_SUBSPACE, U_goal, n=n, subspace=subspace, device=device)
    return FinalFidelityConstraint(f, [name], final_fidelity, traj)


def FinalKetFidelityConstraint(psi_goal, name, final_fidelity, traj, device=0):
    """src/quantum_constraints.jl:20-35."""
    return FinalFidelityConstraint(Fidelity(Fidelity.KET, psi_goal, device=device), [name], final_fidelity, traj)


def FinalCoherentKetFidelityConstraint(psi_goals, names, final_fidelity, traj, device=0):
    """src/quantum_constraints.jl:65-100."""
    if len(psi_goals) != len(names):
        raise AssertionError("Number of names must match number of goals")
    return FinalFidelityConstraint(Fidelity(Fidelity.COHERENT_KET, np.array(psi_goals), device=device), names, final_fidelity, traj)


# --------------------------------------------------------------------------------------- #
# Auxiliary terms of the templates
# --------------------------------------------------------------------------------------- #
class DerivativeIntegrator:
    """``DerivativeIntegrator(x_name, dx_name, traj)``: x_{k+1} - x_k - dt_k dx_k = 0
    (src/problem_templates/smooth_pulse_problem.jl:150-152; spline_pulse_problem.jl:176)."""

    def __init__(self, x_name: str, dx_name: str, traj: NamedTrajectory, device: int = 0):
        xi, dxi = traj.components[x_name], traj.components[dx_name]
        if len(xi) != len(dxi):
            raise ValueError("x and dx must have the same dimension")
        self.aux = Aux(traj.N, traj.dim, int(traj.components[traj.timestep][0]),
                       derivatives=[(int(xi[0]), int(dxi[0]), len(xi))], device=device)
        self.dim = self.aux.derivative_dim

    def evaluate(self, delta: np.ndarray, traj: NamedTrajectory) -> np.ndarray:
        if delta.shape != (self.dim,):
            raise ValueError(f"delta has shape {delta.shape}, expected ({self.dim},)")
        delta[:] = self.aux.derivative_eval(traj.datavec)[0]
        return delta


class LeakageObjective:
    """``LeakageObjective(indices, name, traj; times, Qs)`` (src/quantum_objectives.jl:291-307):
    sum over the knots of Q_t * sum(abs2, x_t[indices]) / length(indices).  indices are 0-based here."""

    def __init__(self, indices, name: str, traj: NamedTrajectory, Qs=None, device: int = 0):
        self.aux = Aux(traj.N, traj.dim, int(traj.components[traj.timestep][0]),
                       leak_x_off=int(traj.components[name][0]), leak_idx=list(indices), device=device)
        self.Qs = np.ones(traj.N) if Qs is None else np.asarray(Qs, dtype=float)
        self.cols = (np.arange(traj.N)[:, None] * traj.dim + traj.components[name][0] + np.asarray(indices)[None, :]).reshape(-1)

    def __call__(self, traj):
        return self.aux.leakage_eval(traj.datavec, self.Qs)[0]

    def gradient(self, grad: np.ndarray, traj) -> np.ndarray:
        grad[:] = 0.0
        np.add.at(grad, self.cols, self.aux.leakage_eval(traj.datavec, self.Qs)[1])
        return grad


class LeakageConstraint:
    """``LeakageConstraint(value, indices, name, traj)`` (src/quantum_constraints.jl:156-172):
    abs2.(x_t[indices]) .- value <= 0 at every knot."""

    def __init__(self, value: float, indices, name: str, traj: NamedTrajectory, device: int = 0):
        self.value = float(value)
        self.aux = Aux(traj.N, traj.dim, int(traj.components[traj.timestep][0]),
                       leak_x_off=int(traj.components[name][0]), leak_idx=list(indices), device=device)
        self.equality = False

    def values(self, traj):
        return self.aux.leakage_eval(traj.datavec, None, self.value)[2]

    def jacobian_values(self, traj):
        return self.aux.leakage_eval(traj.datavec, None, self.value)[3]


def get_iso_vec_leakage_indices(subspace, levels: int) -> np.ndarray:
    """0-based iso-vec operator indices coupling `subspace` to its complement (both directions, real
    and imaginary parts) -- the set PiccoloQuantumObjects' helper of this name yields: with it the
    leakage example's logged objective is reproduced to 1e-18 (tests/test_reference_fixtures.py)."""
    n = int(levels)
    sub = set(int(s) for s in subspace)
    return np.asarray([j * 2 * n + part + i for j in range(n) for part in (0, n) for i in range(n)
                       if (i in sub) != (j in sub)], dtype=np.int64)


class QuadraticRegularizer:
    """``QuadraticRegularizer(name, traj, R)`` (src/problem_templates/smooth_pulse_problem.jl:139-152):
    J = 1/2 sum_t (dt_t v_t)' diag(R) (dt_t v_t) over all N knots; R scalar or one weight per entry."""

    def __init__(self, name: str, traj: NamedTrajectory, R, device: int = 0):
        idx = traj.components[name]
        self.term = (int(idx[0]), len(idx), R)
        self.aux = Aux(traj.N, traj.dim, int(traj.components[traj.timestep][0]), device=device)

    def __call__(self, traj):
        return self.aux.regularizer_eval(traj.datavec, [self.term], want_hessian=False)[0]

    def gradient(self, grad: np.ndarray, traj) -> np.ndarray:
        grad[:] = self.aux.regularizer_gradient(traj.datavec, [self.term])
        return grad

    def hessian_structure(self):
        return self.aux.regularizer_structure([self.term])[1]

    def hessian_values(self, traj):
        return self.aux.regularizer_eval(traj.datavec, [self.term])[2]


class MinimumTimeObjective:
    """``MinimumTimeObjective(traj; D)`` (src/problem_templates/minimum_time_problem.jl:84-131):
    J = D * sum of the first N-1 timesteps."""

    def __init__(self, traj: NamedTrajectory, D: float = 100.0, device: int = 0):
        self.D = float(D)
        self.aux = Aux(traj.N, traj.dim, int(traj.components[traj.timestep][0]), device=device)

    def __call__(self, traj):
        return self.aux.regularizer_eval(traj.datavec, [], self.D, want_hessian=False)[0]

    def gradient(self, grad: np.ndarray, traj) -> np.ndarray:
        grad[:] = self.aux.regularizer_gradient(traj.datavec, [], self.D)
        return grad


class TimeStepsAllEqualConstraint:
    """dt_t - dt_N = 0 for t < N ("timesteps all equal constraint" in the reference's solver logs)."""

    def __init__(self, traj: NamedTrajectory, device: int = 0):
        self.aux = Aux(traj.N, traj.dim, int(traj.components[traj.timestep][0]), device=device)
        self.dim = traj.N - 1
        self.equality = True

    def values(self, traj):
        return self.aux.timesteps_equal(traj.datavec)[0]

    def jacobian(self, traj, row_base: int = 0):
        _, rc, vals = self.aux.timesteps_equal(traj.datavec, row_base)
        return rc, vals


# aliases for the old (docs / BASELINE.json) names, SURVEY.md Appendix A
UnitaryIntegrator = KetIntegrator = BilinearIntegrator
