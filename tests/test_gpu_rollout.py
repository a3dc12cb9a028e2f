"""GPU tests of qcb_rollout: x_{k+1} = exp(dt_k G(u_k)) x_k with all propagators computed in parallel."""
import numpy as np
import pytest
from scipy.linalg import expm

import qcb200
from qcb200 import Evaluator, workloads, _lib
from conftest import oracle_layout, rel_err

pytestmark = pytest.mark.gpu


def oracle_rollout(w, x_init=None):
    L = w.layout
    Zm = w.Z.reshape(L.N, L.z_dim)
    out = np.zeros((L.S, L.N, L.x_dim))
    for s in range(L.S):
        x = Zm[0, L.x_off[s]:L.x_off[s] + L.x_dim] if x_init is None else x_init[s]
        X = x.reshape(L.c, L.d).T
        out[s, 0] = x
        for k in range(L.K):
            G = w.G_drift[s] + np.tensordot(Zm[k, L.u_off:L.u_off + L.m], w.G_drives[s], axes=1)
            X = expm(Zm[k, L.dt_off] * G) @ X
            out[s, k + 1] = X.T.reshape(-1)
    return out


@pytest.mark.parametrize("name,N,S,path", [("cfg1_drift", 12, None, 0), ("cfg2", 10, None, 0), ("cfg3_ket", 8, None, 0),
                                           ("cfg4", 7, 3, 0), ("cfg2", 6, None, 2), ("cfg5", 6, None, 0)])
def test_rollout_matches_scipy_chain(name, N, S, path, built_libs):
    w = workloads.make(name, N=N, S=S)
    ev = Evaluator(w.layout, w.G_drift, w.G_drives, force_path=path)
    states = ev.rollout(w.Z)
    assert rel_err(states, oracle_rollout(w)) < 1e-12
    # closed systems: the rollout of a unitary / normalised ket keeps its norm
    n0 = np.linalg.norm(states[0, 0])
    assert abs(np.linalg.norm(states[0, -1]) - n0) < 1e-12 * n0
    ev.close()


def test_large_d_rollout_by_parallel_prefix(built_libs, monkeypatch):
    """d = 128: the propagators are combined by a Hillis-Steele scan over the knots (batched DMMA
    products), not by a sequential chain; same states as scipy's chain and as the chain kernel."""
    w = workloads.make("cfg5", N=23)
    ev = Evaluator(w.layout, w.G_drift, w.G_drives)
    l0 = ev.kernel_launches
    states = ev.rollout(w.Z)
    n_scan = ev.kernel_launches - l0
    assert rel_err(states, oracle_rollout(w)) < 1e-12
    monkeypatch.setenv("QCB_ROLLOUT_CHAIN", "1")
    chain = ev.rollout(w.Z)
    assert rel_err(states, chain) < 1e-13
    # 22 intervals: 5 scan rounds
    x0 = np.random.default_rng(1).normal(size=(1, w.layout.x_dim))
    monkeypatch.delenv("QCB_ROLLOUT_CHAIN")
    a = ev.rollout(w.Z, x_init=x0)
    monkeypatch.setenv("QCB_ROLLOUT_CHAIN", "1")
    b = ev.rollout(w.Z, x_init=x0)
    assert rel_err(a, b) < 1e-13 and np.array_equal(a[0, 0], x0[0])
    assert n_scan > 0
    ev.close()


def test_rollout_closes_the_dynamics_constraint(built_libs):
    """Feeding the rolled-out states back as the trajectory gives a zero defect (the reference's
    post-solve criterion ||delta||_inf < 1e-3, smooth_pulse_problem.jl:511-515 -- here to rounding)."""
    w = workloads.make("cfg2", N=40)
    L = w.layout
    ev = Evaluator(L, w.G_drift, w.G_drives)
    rng = np.random.default_rng(0)
    x0 = qcb200.isomorphisms.operator_to_iso_vec(np.eye(3))
    states = ev.rollout(w.Z, x_init=x0[None])
    Z = w.Z.reshape(L.N, L.z_dim).copy()
    Z[:, :L.x_dim] = states[0]
    assert np.abs(ev.eval_defect(Z.reshape(-1))).max() < 1e-13
    U = qcb200.isomorphisms.iso_vec_to_operator(states[0, -1])
    assert np.abs(U.conj().T @ U - np.eye(3)).max() < 1e-12
    ev.close()


def test_rollout_needs_the_whole_trajectory(built_libs):
    w = workloads.make("cfg2", N=8)
    ev = Evaluator(w.layout, w.G_drift, w.G_drives, k_begin=2, k_end=5)
    with pytest.raises(_lib.QcbError) as ei:
        ev.rollout(w.Z)
    assert ei.value.status == 1
    ev.close()
