"""GPU tests at BASELINE.json's full sizes: direct parity where the CPU port finishes in seconds,
size-independent properties elsewhere."""
import numpy as np
import pytest

import qcb200
from qcb200 import Evaluator, workloads
import qc_port as P
from conftest import oracle_layout, rel_err

pytestmark = pytest.mark.gpu
TOL = 1e-10


@pytest.mark.parametrize("name", ["cfg1", "cfg2", "cfg3_ket"])
def test_full_size_small_configs_against_port(name, built_libs):
    w = workloads.make(name)
    ev = Evaluator(w.layout, w.G_drift, w.G_drives)
    d1, j1, h1 = ev.eval_all(w.Z, w.mu)
    d0, j0, h0 = P.evaluate(oracle_layout(w.layout), w.G_drift, w.G_drives, w.Z, w.mu)
    assert rel_err(d1, d0) < TOL and rel_err(j1, j0) < TOL and rel_err(h1, h0) < TOL
    ev.close()


def test_full_size_cfg3_unitary_against_port(built_libs):
    w = workloads.make("cfg3")                      # 499 units, x_dim 162, dense-x Jacobian (13.6 M nnz)
    ev = Evaluator(w.layout, w.G_drift, w.G_drives)
    d1, j1, h1 = ev.eval_all(w.Z, w.mu)
    d0, ja, ha = P.evaluate_arenas(oracle_layout(w.layout), w.G_drift, w.G_drives, w.Z, w.mu)
    ji, js, hi = P.arena_codes(oracle_layout(w.layout))
    assert rel_err(d1, d0) < TOL
    assert rel_err(j1.reshape(499, -1), ja[:, ji] * js) < TOL
    assert rel_err(h1.reshape(499, -1), ha[:, hi]) < TOL
    ev.close()


def test_full_size_cfg4_against_port(built_libs):
    w = workloads.make("cfg4")                      # 102 144 units (knot x sample)
    L = w.layout
    ev = Evaluator(L, w.G_drift, w.G_drives)
    d1, j1, h1 = ev.eval_all(w.Z, w.mu)
    OL = oracle_layout(L)
    d0, ja, ha = P.evaluate_arenas(OL, w.G_drift, w.G_drives, w.Z, w.mu)
    assert rel_err(d1, d0) < TOL
    j1 = j1.reshape(L.S, L.K, -1); h1 = h1.reshape(L.S, L.K, -1)
    ja = ja.reshape(L.S, L.K, -1); ha = ha.reshape(L.S, L.K, -1)
    for s in (0, 1, 100, 255):                      # x_off differs per system: per-system codes
        ji, js, hi = P.arena_codes(OL, s)
        assert rel_err(j1[s], ja[s][:, ji] * js) < TOL and rel_err(h1[s], ha[s][:, hi]) < TOL
    ji, js, hi = P.arena_codes(OL, 0)               # same pattern for all s (states precede t, dt, u)
    assert rel_err(j1, ja[:, :, ji] * js) < TOL and rel_err(h1, ha[:, :, hi]) < TOL
    ev.close()


def test_cfg5_full_size_properties(built_libs):
    """d = 128, T = 2000 (1999 units, 2.3e9 Jacobian values): properties that need no oracle, plus a
    16-unit sample against the CPU port."""
    import torch
    w = workloads.make("cfg5")
    L = w.layout
    ev = Evaluator(L, w.G_drift, w.G_drives)
    dZ, dmu = torch.from_numpy(w.Z).cuda(), torch.from_numpy(w.mu).cuda()
    dd = torch.empty(ev.defect_dim, dtype=torch.float64, device="cuda")
    dj = torch.empty(ev.jacobian_nnz, dtype=torch.float64, device="cuda")
    dh = torch.empty(ev.hessian_nnz, dtype=torch.float64, device="cuda")
    ev.eval_all_device(dZ, dmu, dd, dj, dh)
    J = dj.view(L.K, L.jac_nnz_unit)
    d, c, m = L.d, L.c, L.m
    # (1) block-diagonal x-block: c identical copies of -E per unit, and E is orthogonal (closed system)
    xb = J[:, :c * d * d].view(L.K, c, d, d)       # [unit][block][col][row]
    assert float((xb - xb[:, :1]).abs().max()) == 0.0
    E = -xb[:, 0].transpose(1, 2)                   # [unit][row][col]
    eye = torch.eye(d, dtype=torch.float64, device="cuda")
    assert float((E.transpose(1, 2) @ E - eye).abs().max()) < 1e-12
    # (2) defect consistent with the Jacobian's own E:  delta = x_{k+1} - E x_k
    Zm = dZ.view(L.N, L.z_dim)
    X = Zm[:-1, :L.x_dim].reshape(L.K, c, d).transpose(1, 2)
    Xn = Zm[1:, :L.x_dim].reshape(L.K, c, d).transpose(1, 2)
    delta = (Xn - E @ X).transpose(1, 2).reshape(-1)
    assert float((dd - delta).abs().max()) < 1e-12
    # (3) the x_{k+1} block is the identity: last x_dim entries of a unit are +1
    assert float((J[:, -L.x_dim:] - 1).abs().max()) == 0.0
    # (4) Hessian of the Lagrangian is linear in mu
    dh2 = torch.empty_like(dh)
    ev.eval_all_device(dZ, -2.0 * dmu, dd, None, dh2)
    assert float((dh2 + 2.0 * dh).abs().max()) < 1e-10 * float(dh.abs().max())
    # (5) a sample of units against the CPU port
    OL = oracle_layout(L)
    ks = [0, 1, 2, 3, 500, 501, 999, 1000, 1500, 1501, 1995, 1996, 1997, 1998, 777, 1234]
    ji, js, hi = P.arena_codes(OL)
    Jh, Hh, Dh = J.cpu().numpy(), dh.view(L.K, -1).cpu().numpy(), dd.view(L.K, -1).cpu().numpy()
    for k in ks[::4]:
        d0, ja, ha = P.evaluate_arenas(OL, w.G_drift, w.G_drives, w.Z, w.mu, k_range=(k, k + 1))
        assert rel_err(Dh[k], d0) < TOL and rel_err(Jh[k], ja[0][ji] * js) < TOL and rel_err(Hh[k], ha[0][hi]) < TOL
    ev.close()
