"""GPU parity tests: the CUDA library, called through the C ABI, against the oracle on the same
seeded inputs.  Tolerance: 1e-10 relative, block-normwise (BASELINE.json north_star; SURVEY
Appendix B), FP64.  Structures must be bit-exact."""
import os

import numpy as np
import pytest

import qcb200
from qcb200 import Evaluator, KnotLayout, workloads, piccolo, _lib
import qc_oracle as O
import qc_port as P
from conftest import ROOT, oracle_layout, rel_err

pytestmark = pytest.mark.gpu
TOL = 1e-10
GOLD = os.path.join(ROOT, "tests", "golden")


def check_all(ev, OL, w, Z=None, mu=None, ref=None, tol=TOL):
    Z = w.Z if Z is None else Z
    mu = w.mu if mu is None else mu
    d0, j0, h0 = ref or O.evaluate(OL, w.G_drift, w.G_drives, Z, mu)
    d1, j1, h1 = ev.eval_all(Z, mu)
    assert rel_err(d1, d0) < tol, ("defect", rel_err(d1, d0))
    assert rel_err(j1, j0) < tol, ("jacobian", rel_err(j1, j0))
    assert rel_err(h1, h0) < tol, ("hessian", rel_err(h1, h0))
    return d1, j1, h1


# ---- every BASELINE config at oracle-sized N, both kernels paths where they apply ----------------
CASES = [("cfg1", 8, None, 0), ("cfg1_drift", 12, None, 0), ("cfg2", 9, None, 0), ("cfg3", 5, None, 0),
         ("cfg3_ket", 6, None, 0), ("cfg4", 6, 5, 0),
         ("cfg1_drift", 9, None, 2), ("cfg2", 5, None, 2), ("cfg3", 4, None, 2), ("cfg3_ket", 5, None, 2), ("cfg4", 5, 3, 2)]


@pytest.mark.parametrize("name,N,S,path", CASES)
def test_cuda_matches_scipy_oracle(name, N, S, path, built_libs):
    w = workloads.make(name, N=N, S=S)
    OL = oracle_layout(w.layout)
    ev = Evaluator(w.layout, w.G_drift, w.G_drives, force_path=path)
    check_all(ev, OL, w)
    # structures bit-exact against the oracle's generator
    r, c = ev.jacobian_structure(row_base=7)
    ro, co = OL.jacobian_structure(row_base=7)
    assert np.array_equal(r, ro) and np.array_equal(c, co)
    r, c = ev.hessian_structure()
    ro, co = OL.hessian_structure()
    assert np.array_equal(r, ro) and np.array_equal(c, co)
    assert ev.kernel_launches > 0
    ev.close()


@pytest.mark.parametrize("name,N,S", [("cfg1_drift", 5, None), ("cfg2", 4, None), ("cfg3_ket", 4, None),
                                      ("cfg3", 3, None), ("cfg4", 4, 3)])
def test_cuda_matches_committed_golden_vectors(name, N, S, built_libs):
    g = np.load(os.path.join(GOLD, "oracle_small.npz"))
    w = workloads.make(name, N=N, S=S, seed=7)
    ev = Evaluator(w.layout, w.G_drift, w.G_drives)
    d1, j1, h1 = ev.eval_all(g[f"{name}.Z"], g[f"{name}.mu"])
    assert rel_err(d1, g[f"{name}.delta"]) < TOL and rel_err(j1, g[f"{name}.jac"]) < TOL
    assert rel_err(h1, g[f"{name}.hess"]) < TOL
    r, c = ev.jacobian_structure()
    assert np.array_equal(r, g[f"{name}.jac_rows"]) and np.array_equal(c, g[f"{name}.jac_cols"])


def test_cfg5_d128_against_port_and_scipy_spot(built_libs):
    """cfg5 shape (d = 128, c = 64, m = 12, DMMA path): all values of 6 units against the CPU port
    (itself pinned to scipy by tests/test_oracle.py), plus a direct scipy check of E."""
    from scipy.linalg import expm
    w = workloads.make("cfg5", N=7)
    L = w.layout
    OL = oracle_layout(L)
    ev = Evaluator(L, w.G_drift, w.G_drives)
    ref = P.evaluate(OL, w.G_drift, w.G_drives, w.Z, w.mu)
    d1, j1, h1 = check_all(ev, OL, w, ref=ref)
    X, Xn, M, u, dt = OL.unit_inputs(w.Z, w.mu, 0, 3)
    E = expm(dt * (w.G_drift[0] + np.tensordot(u, w.G_drives[0], axes=1)))
    assert rel_err(d1[3 * L.x_dim:4 * L.x_dim], (Xn - E @ X).T.reshape(-1)) < TOL
    # block-diagonal structure: the first d*d Jacobian entries of a unit are -E, column-major
    assert rel_err(j1[3 * L.jac_nnz_unit:3 * L.jac_nnz_unit + L.d ** 2], -E.T.reshape(-1)) < TOL
    ev.close()


def _block_errs(L, got, ref):
    """Block-normwise errors of one unit's Jacobian / Hessian COO vectors, block by block, from the
    per-entry source tags of the oracle's structure generator."""
    OL = oracle_layout(L)
    out = {}
    for nm, pat, a, b in (("jac", OL.jac_unit_pattern(0), got[1], ref[1]), ("hess", OL.hess_unit_pattern(0), got[2], ref[2])):
        tags = np.array([e[2][0] + (str(e[2][1]) if e[2][0] in ("U", "XU") else "") for e in pat])
        for t in np.unique(tags):
            sel = tags == t
            out[f"{nm}.{t}"] = rel_err(a[sel], b[sel]) if np.abs(b[sel]).max() > 0 else float(np.abs(a[sel]).max())
    out["delta"] = rel_err(got[0], ref[0])
    return out


@pytest.mark.parametrize("storage", ["iso", "full"])
def test_cfg5_d128_against_committed_scipy_goldens(storage, built_libs, monkeypatch):
    """Headline shape, independent check: every block (-E, all 12 L_j X panels, G E X, all of (x,u),
    (x,dt), (u,u), (u,dt), (dt,dt)) of two full cfg5 units -- one without squarings, one with three --
    against tests/golden/oracle_cfg5.npz (scipy Pade expm / expm_frechet / block-triangular second
    order, generated offline by tests/golden/make_golden_cfg5.py), 1e-10 block-normwise, on the iso
    half-storage kernels and on the full-storage kernels."""
    from test_oracle import cfg5_golden
    if storage == "full":
        monkeypatch.setenv("QCB_NO_ISO", "1")
    w, OL, g, units = cfg5_golden()
    L = w.layout
    ev = Evaluator(L, w.G_drift, w.G_drives)
    d1, j1, h1 = ev.eval_all(g["Z"], g["mu"])
    nj, nh, xd = L.jac_nnz_unit, L.hess_nnz_unit, L.x_dim
    for k, ref in enumerate(units):
        got = (d1[k * xd:(k + 1) * xd], j1[k * nj:(k + 1) * nj], h1[k * nh:(k + 1) * nh])
        errs = _block_errs(L, got, ref)
        assert len(errs) == 1 + 3 + L.m + L.m + 1 + 3, sorted(errs)      # delta; E, T, I; U0..; XU0..; XT; UU, UT, TT
        bad = {b: e for b, e in errs.items() if e > TOL}
        assert not bad, (storage, k, bad)
    # device-resident route (sorted squaring classes, two lanes) gives the same numbers
    import torch
    dZ, dmu = torch.from_numpy(g["Z"]).cuda(), torch.from_numpy(g["mu"]).cuda()
    dd = torch.zeros(ev.defect_dim, dtype=torch.float64, device="cuda")
    dj = torch.zeros(ev.jacobian_nnz, dtype=torch.float64, device="cuda")
    dh = torch.zeros(ev.hessian_nnz, dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    ev.eval_all_device(dZ, dmu, dd, dj, dh)
    assert rel_err(dj.cpu().numpy(), j1) < 1e-13 and rel_err(dh.cpu().numpy(), h1) < 1e-13
    ev.close()


@pytest.mark.parametrize("mix,iso", [(1, False), (2, True), (3, True), (3, False), (12, True)])
def test_d128_drive_sparsity_classes(mix, iso, built_libs):
    """Large-d path with drive generators of 2, 3 (ELL width 2 / 4: gather products) and 12 nonzeros
    per row (beyond the ELL limit of 8: dense DMMA products) against the CPU port; iso=False breaks
    the real-iso block structure of the drift so that the full-storage kernels run."""
    w = workloads.make("cfg5", N=4)
    L = w.layout
    OL = oracle_layout(L)
    Gd = np.array(w.G_drives)
    m = Gd.shape[1]
    mixed = np.stack([sum((0.5 + 0.1 * t) * Gd[0, (j + t) % m] for t in range(mix)) for j in range(m)])[None]
    assert max(int((g != 0).sum(axis=0).max()) for g in mixed[0]) == mix
    G0 = np.array(w.G_drift)
    if not iso:
        G0[0, 0, 1] += 0.05
    ev = Evaluator(L, G0, mixed)
    ref = P.evaluate(OL, G0, mixed, w.Z, w.mu)
    d1, j1, h1 = ev.eval_all(w.Z, w.mu)
    assert rel_err(d1, ref[0]) < TOL and rel_err(j1, ref[1]) < TOL and rel_err(h1, ref[2]) < TOL
    ev.close()


@pytest.mark.parametrize("iso,dense", [(False, False), (True, True), (False, True)])
def test_d128_device_route_variants(iso, dense, built_libs):
    """Device-resident calls on the large-d path with timesteps spread over several squaring classes
    (sorted units, concurrent lanes) for the full-storage kernels (iso=False) and for dense drive
    generators (no ELL products), against the port."""
    import torch
    w = workloads.make("cfg5", N=11)
    L = w.layout
    OL = oracle_layout(L)
    G0, Gd = np.array(w.G_drift), np.array(w.G_drives)
    if not iso:
        G0[0, 0, 1] += 0.05
    if dense:
        m = Gd.shape[1]
        Gd = np.stack([sum((0.4 + 0.05 * t) * Gd[0, (j + t) % m] for t in range(m)) for j in range(m)])[None]
    Z = w.Z.reshape(L.N, L.z_dim).copy()
    Z[:, L.dt_off] *= np.random.default_rng(9).permutation(np.linspace(0.4, 5.0, L.N))
    if dense:
        Z[:, L.u_off:L.u_off + L.m] *= 0.3
    Z = Z.reshape(-1)
    ev = Evaluator(L, G0, Gd)
    ref = P.evaluate(OL, G0, Gd, Z, w.mu)
    dZ, dmu = torch.from_numpy(Z).cuda(), torch.from_numpy(w.mu).cuda()
    dd = torch.zeros(ev.defect_dim, dtype=torch.float64, device="cuda")
    dj = torch.zeros(ev.jacobian_nnz, dtype=torch.float64, device="cuda")
    dh = torch.zeros(ev.hessian_nnz, dtype=torch.float64, device="cuda")
    ev.eval_all_device(dZ, dmu, dd, dj, dh)
    dd, dj, dh = dd.cpu().numpy(), dj.cpu().numpy(), dh.cpu().numpy()
    nj, nh = L.jac_nnz_unit, L.hess_nnz_unit
    assert rel_err(dd, ref[0]) < TOL
    for k in range(L.N - 1):
        assert rel_err(dj[k * nj:(k + 1) * nj], ref[1][k * nj:(k + 1) * nj]) < TOL, k
        assert rel_err(dh[k * nh:(k + 1) * nh], ref[2][k * nh:(k + 1) * nh]) < TOL, k
    ev.close()


def test_d128_two_systems_with_different_sparse_drives(built_libs):
    """Sampling batch on the large-d path (S = 2, d = 128): every system has its own ELL tables for
    the sparse drive products (second system: drives in reverse order and rescaled, detuned drift)."""
    from qcb200.layout import KnotLayout, JAC_BLOCK_DIAG
    w = workloads.make("cfg5", N=3)
    d, c, m, N = w.layout.d, w.layout.c, w.layout.m, 3
    L = KnotLayout.smooth_pulse(d, c, m, N, 2, JAC_BLOCK_DIAG, with_derivatives=False)
    G0 = np.stack([w.G_drift[0], 1.1 * w.G_drift[0]])
    Gd = np.stack([w.G_drives[0], 0.9 * w.G_drives[0][::-1]])
    rng = np.random.default_rng(8)
    Z = np.zeros((N, L.z_dim))
    Zs = w.Z.reshape(N, w.layout.z_dim)
    for s in range(2):
        Z[:, L.x_off[s]:L.x_off[s] + L.x_dim] = np.roll(Zs[:, :L.x_dim], s, axis=0)
    Z[:, L.u_off:L.u_off + m] = 0.1 * rng.normal(size=(N, m))
    Z[:, L.dt_off] = 0.1 * (1 + 0.1 * rng.uniform(-1, 1, N))
    Z = Z.reshape(-1)
    mu = rng.normal(size=2 * (N - 1) * L.x_dim)
    OL = oracle_layout(L)
    ev = Evaluator(L, G0, Gd)
    ref = P.evaluate(OL, G0, Gd, Z, mu)
    d1, j1, h1 = ev.eval_all(Z, mu)
    assert rel_err(d1, ref[0]) < TOL and rel_err(j1, ref[1]) < TOL and rel_err(h1, ref[2]) < TOL
    ev.close()


def test_d128_chunks_take_their_own_squaring_count(built_libs, monkeypatch):
    """Large-d path: with a scratch budget of a few units per chunk and timesteps that grow along the
    trajectory, successive chunks run 0, 1, 2, ... squaring levels; every unit must still match the port."""
    monkeypatch.setenv("QCB_SCRATCH_MB", "64")      # the minimum: about 7 units of d = 128 per chunk
    w = workloads.make("cfg5", N=17)
    L = w.layout
    OL = oracle_layout(L)
    Z = w.Z.reshape(L.N, L.z_dim).copy()
    Z[:, L.dt_off] *= np.linspace(0.3, 6.0, L.N)
    Z = Z.reshape(-1)
    ev = Evaluator(L, w.G_drift, w.G_drives)
    ref = P.evaluate(OL, w.G_drift, w.G_drives, Z, w.mu)
    d1, j1, h1 = ev.eval_all(Z, w.mu)
    assert rel_err(d1, ref[0]) < TOL and rel_err(j1, ref[1]) < TOL and rel_err(h1, ref[2]) < TOL
    # per unit, not only in the block norm: late units have much larger entries than early ones
    nj, nh = L.jac_nnz_unit, L.hess_nnz_unit
    for k in range(L.N - 1):
        assert rel_err(j1[k * nj:(k + 1) * nj], ref[1][k * nj:(k + 1) * nj]) < TOL
        assert rel_err(h1[k * nh:(k + 1) * nh], ref[2][k * nh:(k + 1) * nh]) < TOL
    # device-resident call: the units are visited sorted by squaring count (shuffle the timesteps so
    # that the sorted order differs from the natural one) and written back to their own rows
    import torch
    Zs = Z.reshape(L.N, L.z_dim).copy()
    Zs[:-1, L.dt_off] = Zs[np.random.default_rng(5).permutation(L.N - 1), L.dt_off]
    Zs = Zs.reshape(-1)
    ref = P.evaluate(OL, w.G_drift, w.G_drives, Zs, w.mu)
    dZ, dmu = torch.from_numpy(Zs).cuda(), torch.from_numpy(w.mu).cuda()
    dd = torch.empty(ev.defect_dim, dtype=torch.float64, device="cuda")
    dj = torch.empty(ev.jacobian_nnz, dtype=torch.float64, device="cuda")
    dh = torch.empty(ev.hessian_nnz, dtype=torch.float64, device="cuda")
    ev.eval_all_device(dZ, dmu, dd, dj, dh)
    dd, dj, dh = dd.cpu().numpy(), dj.cpu().numpy(), dh.cpu().numpy()
    assert rel_err(dd, ref[0]) < TOL
    for k in range(L.N - 1):
        assert rel_err(dj[k * nj:(k + 1) * nj], ref[1][k * nj:(k + 1) * nj]) < TOL, k
        assert rel_err(dh[k * nh:(k + 1) * nh], ref[2][k * nh:(k + 1) * nh]) < TOL, k
    ev.close()
    # default scratch budget: the whole slab fits, so the squaring classes run as concurrent lanes on
    # two streams over disjoint scratch slices; twice in a row (the second call reuses everything)
    monkeypatch.delenv("QCB_SCRATCH_MB")
    ev = Evaluator(L, w.G_drift, w.G_drives)
    for _ in range(2):
        dd = torch.zeros(ev.defect_dim, dtype=torch.float64, device="cuda")
        dj = torch.zeros(ev.jacobian_nnz, dtype=torch.float64, device="cuda")
        dh = torch.zeros(ev.hessian_nnz, dtype=torch.float64, device="cuda")
        ev.eval_all_device(dZ, dmu, dd, dj, dh)
        dd, dj, dh = dd.cpu().numpy(), dj.cpu().numpy(), dh.cpu().numpy()
        assert rel_err(dd, ref[0]) < TOL
        for k in range(L.N - 1):
            assert rel_err(dj[k * nj:(k + 1) * nj], ref[1][k * nj:(k + 1) * nj]) < TOL, k
            assert rel_err(dh[k * nh:(k + 1) * nh], ref[2][k * nh:(k + 1) * nh]) < TOL, k
    # lanes and the serial order give the same numbers (same chunks, same kernels)
    monkeypatch.setenv("QCB_NO_LANES", "1")
    d2 = torch.zeros(ev.defect_dim, dtype=torch.float64, device="cuda")
    j2 = torch.zeros(ev.jacobian_nnz, dtype=torch.float64, device="cuda")
    h2 = torch.zeros(ev.hessian_nnz, dtype=torch.float64, device="cuda")
    ev.eval_all_device(dZ, dmu, d2, j2, h2)
    assert np.array_equal(d2.cpu().numpy(), dd) and np.array_equal(j2.cpu().numpy(), dj) and np.array_equal(h2.cpu().numpy(), dh)
    ev.close()


# ---- call variants of the ABI ----------------------------------------------------------------------
def test_separate_callbacks_equal_fused_call(built_libs):
    w = workloads.make("cfg2", N=7)
    ev = Evaluator(w.layout, w.G_drift, w.G_drives)
    d, j, h = ev.eval_all(w.Z, w.mu)
    assert np.array_equal(ev.eval_defect(w.Z), d)
    assert np.array_equal(ev.eval_jacobian(w.Z), j)
    assert np.array_equal(ev.eval_hessian(w.Z, w.mu), h)
    with pytest.raises(_lib.QcbError):
        ev.eval_all_raw(w.Z, None, None, None, np.empty(ev.hessian_nnz))   # Hessian needs mu
    ev.close()


@pytest.mark.parametrize("name,N,path", [("cfg2", 6, 0), ("cfg3", 4, 2)])
def test_device_buffer_entry_point(name, N, path, built_libs):
    import torch
    w = workloads.make(name, N=N)
    ev = Evaluator(w.layout, w.G_drift, w.G_drives, force_path=path)
    d, j, h = ev.eval_all(w.Z, w.mu)
    dZ, dmu = torch.from_numpy(w.Z).cuda(), torch.from_numpy(w.mu).cuda()
    dd = torch.empty(ev.defect_dim, dtype=torch.float64, device="cuda")
    dj = torch.empty(ev.jacobian_nnz, dtype=torch.float64, device="cuda")
    dh = torch.empty(ev.hessian_nnz, dtype=torch.float64, device="cuda")
    ev.eval_all_device(dZ, dmu, dd, dj, dh)
    assert np.array_equal(dd.cpu().numpy(), d) and np.array_equal(dj.cpu().numpy(), j)
    assert np.array_equal(dh.cpu().numpy(), h)
    ev.close()


@pytest.mark.parametrize("name,N,S,path", [("cfg2", 11, None, 0), ("cfg4", 9, 3, 0), ("cfg3", 7, None, 2)])
def test_time_slab_handles_tile_the_trajectory(name, N, S, path, built_libs):
    w = workloads.make(name, N=N, S=S)
    L = w.layout
    full = Evaluator(L, w.G_drift, w.G_drives, force_path=path)
    d, j, h = full.eval_all(w.Z, w.mu)
    d = d.reshape(L.S, L.K, -1); j = j.reshape(L.S, L.K, -1); h = h.reshape(L.S, L.K, -1)
    for k0, k1 in [(0, 3), (3, L.K - 1), (L.K - 1, L.K)]:
        ev = Evaluator(L, w.G_drift, w.G_drives, k_begin=k0, k_end=k1, force_path=path)
        ds, js, hs = ev.eval_all(w.Z, w.mu)
        # the fused small-d kernel picks the squaring count per unit: slabs are bit-identical to the
        # full pass; the large-d path uses one count per slab (its maximum), so only to rounding
        same = np.array_equal if path != 2 else (lambda a, b: rel_err(a, b) < 1e-13)
        assert same(ds.reshape(L.S, k1 - k0, -1), d[:, k0:k1])
        assert same(js.reshape(L.S, k1 - k0, -1), j[:, k0:k1])
        assert same(hs.reshape(L.S, k1 - k0, -1), h[:, k0:k1])
        r, c = ev.jacobian_structure()
        assert c.min() >= k0 * L.z_dim + 1 and c.max() <= (k1 + 1) * L.z_dim
        ev.close()
    full.close()


# ---- caller-supplied (reference-side) sparsity order ----------------------------------------------
@pytest.mark.parametrize("path", [0, 2])
def test_custom_structure_permutation_and_structural_zeros(path, built_libs):
    w = workloads.make("cfg2", N=6)
    L = w.layout
    ev = Evaluator(L, w.G_drift, w.G_drives, force_path=path)
    d, j, h = ev.eval_all(w.Z, w.mu)
    jr, jc = ev.jacobian_structure()
    hr, hc = ev.hessian_structure()
    rng = np.random.default_rng(0)
    nj, nh = L.jac_nnz_unit, L.hess_nnz_unit
    pj, ph = rng.permutation(nj), rng.permutation(nh)
    # same permutation inside every unit, plus one extra structurally-zero entry per unit
    # (a (row, t-column) Jacobian entry; a (t, t) Hessian entry), lower-triangular Hessian order
    t_col = L.dt_off - 1
    jr2 = np.concatenate([np.append(jr[k * nj:(k + 1) * nj][pj], k * L.x_dim + 1) for k in range(L.K)])
    jc2 = np.concatenate([np.append(jc[k * nj:(k + 1) * nj][pj], k * L.z_dim + t_col + 1) for k in range(L.K)])
    hr2 = np.concatenate([np.append(hc[k * nh:(k + 1) * nh][ph], k * L.z_dim + t_col + 1) for k in range(L.K)])
    hc2 = np.concatenate([np.append(hr[k * nh:(k + 1) * nh][ph], k * L.z_dim + t_col + 1) for k in range(L.K)])
    ev.set_structure((jr2, jc2), (hr2, hc2))
    assert ev.jacobian_nnz == L.K * (nj + 1) and ev.hessian_nnz == L.K * (nh + 1)
    d2, j2, h2 = ev.eval_all(w.Z, w.mu)
    assert np.array_equal(d2, d)
    j2 = j2.reshape(L.K, nj + 1); h2 = h2.reshape(L.K, nh + 1)
    assert np.array_equal(j2[:, :nj], j.reshape(L.K, nj)[:, pj]) and np.all(j2[:, nj] == 0)
    assert np.array_equal(h2[:, :nh], h.reshape(L.K, nh)[:, ph]) and np.all(h2[:, nh] == 0)
    # a structure that is not periodic / not grouped by unit is refused
    with pytest.raises(_lib.QcbError) as ei:
        ev.set_structure((jr2[::-1].copy(), jc2[::-1].copy()), None)
    assert ei.value.status == 4
    ev.close()


def test_custom_structure_with_both_triangles_counts_each_pair_once(built_libs):
    """A reference-side Hessian structure that lists (a, b) and (b, a): MOI sums the two, so only the
    first occurrence may carry the value.  Also: adopting a structure after a host-buffer evaluation
    (staging buffers sized with the old nnz) must not corrupt anything."""
    w = workloads.make("cfg2", N=5)
    L = w.layout
    ev = Evaluator(L, w.G_drift, w.G_drives, force_path=2)
    d, j, h = ev.eval_all(w.Z, w.mu)                    # allocates the large-path staging first
    hr, hc = ev.hessian_structure()
    nh = L.hess_nnz_unit
    off = hr != hc
    rows2, cols2 = [], []
    for k in range(L.K):
        sl = slice(k * nh, (k + 1) * nh)
        o = off[sl]
        rows2.append(np.concatenate([hr[sl], hc[sl][o]]))     # upper triangle, then its mirror
        cols2.append(np.concatenate([hc[sl], hr[sl][o]]))
    n_off = int(off[:nh].sum())
    ev.set_structure(None, (np.concatenate(rows2), np.concatenate(cols2)))
    assert ev.hessian_nnz == L.K * (nh + n_off)
    d2, j2, h2 = ev.eval_all(w.Z, w.mu)
    h2 = h2.reshape(L.K, nh + n_off)
    assert np.array_equal(h2[:, :nh], h.reshape(L.K, nh)) and np.all(h2[:, nh:] == 0)
    assert np.array_equal(j2, j) and np.array_equal(d2, d)
    ev.close()


# ---- edge cases -------------------------------------------------------------------------------------
def test_minimal_trajectory_two_knots_and_zero_controls(built_libs):
    w = workloads.make("cfg1_drift", N=2, u_scale=0.0)
    ev = Evaluator(w.layout, w.G_drift, w.G_drives)
    check_all(ev, oracle_layout(w.layout), w)
    ev.close()


def test_no_drives(built_libs):
    """m = 0: only the x, dt blocks exist."""
    rng = np.random.default_rng(2)
    n, N = 3, 5
    H = rng.normal(size=(n, n)) + 1j * rng.normal(size=(n, n)); H = H + H.conj().T
    L = KnotLayout(d=2 * n, c=1, m=0, N=N, S=1, z_dim=2 * n + 1, x_off=[0], u_off=2 * n, dt_off=2 * n)
    G0 = O.G_of_H(H)[None]
    Gd = np.zeros((1, 0, 2 * n, 2 * n))
    Z = rng.normal(size=(N, L.z_dim)); Z[:, L.dt_off] = 0.3
    mu = rng.normal(size=L.K * L.x_dim)
    ev = Evaluator(L, G0, Gd)
    d0, j0, h0 = O.evaluate(oracle_layout(L), G0, Gd, Z.reshape(-1), mu)
    d1, j1, h1 = ev.eval_all(Z.reshape(-1), mu)
    assert rel_err(d1, d0) < TOL and rel_err(j1, j0) < TOL and rel_err(h1, h0) < TOL
    ev.close()


@pytest.mark.parametrize("path", [0, 2])
def test_large_norm_many_squarings(path, built_libs):
    w = workloads.make("cfg2", N=5)
    L = w.layout
    Z = w.Z.reshape(L.N, L.z_dim).copy()
    Z[:, L.dt_off] = np.array([9.0, 0.01, 3.0, 0.5, 1.0])      # per-unit squarings differ: 6, 0, 4, 2, ...
    ev = Evaluator(L, w.G_drift, w.G_drives, force_path=path)
    check_all(ev, oracle_layout(L), w, Z=Z.reshape(-1))
    ev.close()


def test_non_skew_generator(built_libs):
    """Lindblad-like / variational generators are not skew-symmetric: exp is not orthogonal."""
    rng = np.random.default_rng(5)
    d, m, N = 6, 2, 5
    L = KnotLayout.smooth_pulse(d, 1, m, N)
    G0 = 0.5 * rng.normal(size=(1, d, d))
    Gd = 0.5 * rng.normal(size=(1, m, d, d))
    Z = rng.normal(size=(N, L.z_dim)); Z[:, L.dt_off] = 0.4; Z[:, L.u_off:L.u_off + m] *= 0.3
    mu = rng.normal(size=L.K * L.x_dim)
    for path in (0, 2):
        ev = Evaluator(L, G0, Gd, force_path=path)
        d0, j0, h0 = O.evaluate(oracle_layout(L), G0, Gd, Z.reshape(-1), mu)
        d1, j1, h1 = ev.eval_all(Z.reshape(-1), mu)
        assert rel_err(d1, d0) < TOL and rel_err(j1, j0) < TOL and rel_err(h1, h0) < TOL
        ev.close()


@pytest.mark.parametrize("path", [0, 2])
def test_non_finite_and_out_of_range_inputs_fail_loudly(path, built_libs):
    w = workloads.make("cfg2", N=5)
    L = w.layout
    ev = Evaluator(L, w.G_drift, w.G_drives, force_path=path)
    Z = w.Z.reshape(L.N, L.z_dim).copy()
    Z[2, L.u_off] = np.nan
    with pytest.raises(_lib.QcbError) as ei:
        ev.eval_all(Z.reshape(-1), w.mu)
    assert ei.value.status == 3
    Z = w.Z.reshape(L.N, L.z_dim).copy()
    Z[1, L.dt_off] = 1e6                      # ||dt G|| far beyond theta * 2^10
    with pytest.raises(_lib.QcbError) as ei:
        ev.eval_all(Z.reshape(-1), w.mu)
    assert ei.value.status == 3
    ev.eval_all(w.Z, w.mu)                    # the handle stays usable afterwards
    ev.close()


def test_knot_offsets_are_runtime_data(built_libs):
    """Component order inside a knot is the caller's: put u first, dt in the middle, x last."""
    rng = np.random.default_rng(9)
    w = workloads.make("cfg2", N=6)
    L0 = w.layout
    L = KnotLayout(d=L0.d, c=L0.c, m=L0.m, N=L0.N, S=1, z_dim=L0.x_dim + 7, x_off=[7], u_off=1, dt_off=4)
    Z0 = w.Z.reshape(L0.N, L0.z_dim)
    Z = rng.normal(size=(L.N, L.z_dim))
    Z[:, 7:] = Z0[:, :L0.x_dim]; Z[:, 1:3] = Z0[:, L0.u_off:L0.u_off + 2]; Z[:, 4] = Z0[:, L0.dt_off]
    ev = Evaluator(L, w.G_drift, w.G_drives)
    d0, j0, h0 = O.evaluate(oracle_layout(L), w.G_drift, w.G_drives, Z.reshape(-1), w.mu)
    d1, j1, h1 = ev.eval_all(Z.reshape(-1), w.mu)
    assert rel_err(d1, d0) < TOL and rel_err(j1, j0) < TOL and rel_err(h1, h0) < TOL
    r, c = ev.hessian_structure()
    ro, co = oracle_layout(L).hessian_structure()
    assert np.array_equal(r, ro) and np.array_equal(c, co)
    ev.close()


# ---- the reference's own operator-level tests, restated (src/quantum_integrators.jl:175-329) ------
def _test_system():
    # QuantumSystem(GATES[:Z], [GATES[:X], GATES[:Y]], [1.0, 1.0]), N = 11, zero controls (:181-191)
    return piccolo.QuantumSystem(workloads.Z2, [workloads.X2, workloads.Y2])


def _traj(state: dict, N=11, m=2, rng=None):
    rng = rng or np.random.default_rng(0)
    comps = dict(state)
    comps.update({"t": np.arange(N)[None] * 0.1, "dt": np.full((1, N), 0.1), "u": np.zeros((m, N))})
    return piccolo.NamedTrajectory(comps, timestep="dt")


def test_reference_test_integrator_unitary_and_ket(built_libs):
    sys_ = _test_system()
    rng = np.random.default_rng(1)
    N = 11
    U = np.stack([qcb200.isomorphisms.operator_to_iso_vec(workloads.haar_unitary(2, rng)) for _ in range(N)], axis=1)
    traj = _traj({"U_iso": U})
    integ = piccolo.BilinearIntegrator(piccolo.UnitaryTrajectory(sys_, workloads.H_GATE), traj)
    assert integ.dim == 8 * (N - 1)
    errs = piccolo.test_integrator(integ, traj, atol=1e-3)       # the reference's tolerance
    assert errs["jacobian"] < 1e-6 and errs["hessian"] < 1e-5     # ours is far tighter
    psi = rng.normal(size=(4, N))
    trajk = _traj({"psi_iso": psi})
    integ = piccolo.BilinearIntegrator(piccolo.KetTrajectory(sys_, [1, 0], [0, 1]), trajk)
    piccolo.test_integrator(integ, trajk, atol=1e-3)
    delta = np.zeros(integ.dim)
    piccolo.evaluate_(delta, integ, trajk)
    assert np.abs(delta).max() > 0
    with pytest.raises(ValueError):
        piccolo.evaluate_(np.zeros(3), integ, trajk)


def test_reference_multiket_and_sampling_dispatch(built_libs):
    sys_ = _test_system()
    rng = np.random.default_rng(2)
    N = 11
    traj = _traj({"psi_iso1": rng.normal(size=(4, N)), "psi_iso2": rng.normal(size=(4, N))})
    mk = piccolo.MultiKetTrajectory(sys_, [[1, 0], [0, 1]], [[0, 1], [1, 0]])
    ints = piccolo.BilinearIntegrator(mk, traj)
    assert isinstance(ints, list) and len(ints) == 2            # one integrator per ket (:63-74)
    for it in ints:
        piccolo.test_integrator(it, traj, atol=1e-3)
    fused = piccolo.BilinearIntegrator(mk, traj, fused=True)
    d_f = fused.evaluate(np.zeros(fused.dim), traj)
    d_s = np.concatenate([it.evaluate(np.zeros(it.dim), traj) for it in ints])
    assert np.array_equal(d_f, d_s)
    # SamplingTrajectory: one integrator per system, shared controls (:91-101)
    systems = [piccolo.QuantumSystem((1 + e) * workloads.Z2, [workloads.X2, workloads.Y2]) for e in (-0.1, 0.0, 0.1)]
    base = piccolo.UnitaryTrajectory(systems[1], workloads.H_GATE)
    st = piccolo.SamplingTrajectory(base, systems)
    U = lambda: np.stack([qcb200.isomorphisms.operator_to_iso_vec(workloads.haar_unitary(2, rng)) for _ in range(N)], axis=1)
    trajs = _traj({nm: U() for nm in st.state_names})
    ints = piccolo.BilinearIntegrator(st, trajs)
    assert len(ints) == 3
    fused = piccolo.BilinearIntegrator(st, trajs, fused=True)
    piccolo.test_integrator(fused, trajs, atol=1e-3)
    d_f = fused.evaluate(np.zeros(fused.dim), trajs)
    d_s = np.concatenate([it.evaluate(np.zeros(it.dim), trajs) for it in ints])
    assert np.array_equal(d_f, d_s)


@pytest.mark.parametrize("name,N,S", [("cfg3", 6, None), ("cfg4", 5, 3), ("cfg5", 4, None)])
def test_large_path_compact_host_transfer_equals_direct(name, N, S, built_libs):
    """qcb_eval_all on the large-d path ships the compact value arenas and expands them on the host
    cores; QCB_E2E_DIRECT=1 selects device-side assembly + full-size D2H.  Same numbers either way."""
    w = workloads.make(name, N=N, S=S)
    ev = Evaluator(w.layout, w.G_drift, w.G_drives, force_path=2)
    d1, j1, h1 = ev.eval_all(w.Z, w.mu)
    os.environ["QCB_E2E_DIRECT"] = "1"
    try:
        d2, j2, h2 = ev.eval_all(w.Z, w.mu)
    finally:
        del os.environ["QCB_E2E_DIRECT"]
    assert np.array_equal(d1, d2) and np.array_equal(j1, j2) and np.array_equal(h1, h2)
    # Jacobian-only and Hessian-only callbacks through the compact route
    assert np.array_equal(ev.eval_jacobian(w.Z), j1) and np.array_equal(ev.eval_hessian(w.Z, w.mu), h1)
    ev.close()


def _random_system(n, m, rng, scale=0.05):
    rh = lambda: (lambda A: (A + A.conj().T) / 2)(rng.normal(size=(n, n)) + 1j * rng.normal(size=(n, n)))
    return scale * O.G_of_H(rh())[None], scale * np.array([O.G_of_H(rh()) for _ in range(m)])[None]


def test_large_path_d256_two_row_tiles(built_libs):
    """d = 256 (7 qubits): the iso half-storage GEMM runs two 128-row tiles per matrix, so the
    row-rotation / sign logic is exercised with m0 > 0.  Ket panel, 2 drives, against the CPU port."""
    rng = np.random.default_rng(3)
    n, m, N = 128, 2, 3
    G0, Gd = _random_system(n, m, rng, scale=0.02)
    L = KnotLayout.smooth_pulse(2 * n, 1, m, N, jac_mode=1)
    Z = rng.normal(size=(N, L.z_dim)); Z[:, L.dt_off] = 0.3; Z[:, L.u_off:L.u_off + m] *= 0.2
    mu = rng.normal(size=L.K * L.x_dim)
    ev = Evaluator(L, G0, Gd)
    d1, j1, h1 = ev.eval_all(Z.reshape(-1), mu)
    d0, j0, h0 = P.evaluate(oracle_layout(L), G0, Gd, Z.reshape(-1), mu)
    assert rel_err(d1, d0) < TOL and rel_err(j1, j0) < TOL and rel_err(h1, h0) < TOL
    ev.close()


def test_large_path_full_storage_d128_without_iso_structure(built_libs):
    """A d = 128 generator WITHOUT the real-iso block form (e.g. a variational / block-triangular
    generator): the full-storage 128 x 128-tile DMMA kernel is used instead of the half-storage one."""
    rng = np.random.default_rng(4)
    d, m, N = 128, 2, 3
    G0 = 0.01 * rng.normal(size=(1, d, d))
    Gd = 0.01 * rng.normal(size=(1, m, d, d))
    L = KnotLayout.smooth_pulse(d, 2, m, N, jac_mode=1)
    Z = rng.normal(size=(N, L.z_dim)); Z[:, L.dt_off] = 0.25; Z[:, L.u_off:L.u_off + m] *= 0.3
    mu = rng.normal(size=L.K * L.x_dim)
    ev = Evaluator(L, G0, Gd)
    d1, j1, h1 = ev.eval_all(Z.reshape(-1), mu)
    d0, j0, h0 = P.evaluate(oracle_layout(L), G0, Gd, Z.reshape(-1), mu)
    assert rel_err(d1, d0) < TOL and rel_err(j1, j0) < TOL and rel_err(h1, h0) < TOL
    ev.close()


@pytest.mark.parametrize("d,c,m,S", [(2, 1, 1, 1), (5, 2, 3, 2), (7, 1, 2, 1), (9, 3, 1, 2), (14, 2, 2, 1), (20, 1, 5, 1),
                                     (22, 1, 2, 1), (23, 2, 1, 1), (40, 3, 2, 1)])
def test_odd_shapes_and_general_generators(d, c, m, S, built_libs):
    """Dimensions without a specialised instance (run-time-d interpreter), odd d, several systems,
    non-skew generators, and d = 40 (large-d path padded to 64)."""
    rng = np.random.default_rng(d * 100 + c * 10 + m)
    N = 4
    G0 = 0.3 * rng.normal(size=(S, d, d)) / np.sqrt(d)
    Gd = 0.3 * rng.normal(size=(S, m, d, d)) / np.sqrt(d)
    L = KnotLayout.smooth_pulse(d, c, m, N, S=S, with_derivatives=False)
    Z = rng.normal(size=(N, L.z_dim)); Z[:, L.dt_off] = rng.uniform(0.1, 1.5, N); Z[:, L.u_off:L.u_off + m] *= 0.5
    mu = rng.normal(size=S * L.K * L.x_dim)
    ev = Evaluator(L, G0, Gd)
    d1, j1, h1 = ev.eval_all(Z.reshape(-1), mu)
    ref = O.evaluate if d <= 14 else P.evaluate
    d0, j0, h0 = ref(oracle_layout(L), G0, Gd, Z.reshape(-1), mu)
    assert rel_err(d1, d0) < TOL and rel_err(j1, j0) < TOL and rel_err(h1, h0) < TOL
    ev.close()


@pytest.mark.parametrize("d,m", [(6, 17), (40, 17), (128, 18)])
def test_many_drives(d, m, built_libs):
    """More than 16 drives: the small-d kernel loops its contraction chunks, the large-d path falls
    back from the 16 x 16 DMMA contraction to the generic one."""
    rng = np.random.default_rng(d + m)
    N = 3
    if d == 128:
        G0, Gd = _random_system(64, m, rng, scale=0.02)      # iso structure -> half-storage path
    else:
        G0 = 0.2 * rng.normal(size=(1, d, d)) / np.sqrt(d)
        Gd = 0.2 * rng.normal(size=(1, m, d, d)) / np.sqrt(d)
    L = KnotLayout.smooth_pulse(d, 1, m, N, with_derivatives=False)
    Z = rng.normal(size=(N, L.z_dim)); Z[:, L.dt_off] = 0.4; Z[:, L.u_off:L.u_off + m] *= 0.2
    mu = rng.normal(size=L.K * L.x_dim)
    ev = Evaluator(L, G0, Gd)
    d1, j1, h1 = ev.eval_all(Z.reshape(-1), mu)
    d0, j0, h0 = P.evaluate(oracle_layout(L), G0, Gd, Z.reshape(-1), mu)
    assert rel_err(d1, d0) < TOL and rel_err(j1, j0) < TOL and rel_err(h1, h0) < TOL
    ev.close()


# ---- Ipopt's new_x flag: eval_g -> eval_jac_g -> eval_h on one iterate reuse the Z-only intermediates ----
@pytest.mark.parametrize("storage", ["iso", "full"])
def test_new_x_reuse_is_bit_identical_and_a_new_point_recomputes(storage, built_libs, monkeypatch):
    import torch
    if storage == "full":
        monkeypatch.setenv("QCB_NO_ISO", "1")
    w = workloads.make("cfg5", N=9)
    L = w.layout
    Zk = w.Z.reshape(L.N, L.z_dim).copy()
    Zk[2, L.dt_off] *= 3.0        # two squaring classes -> two lanes, sorted units
    Zk[5, L.dt_off] *= 6.0
    Z1 = Zk.reshape(-1)
    ev = Evaluator(L, w.G_drift, w.G_drives)
    cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    nanv = lambda n: torch.full((n,), float("nan"), dtype=torch.float64, device="cuda")
    dZ, dmu = cu(Z1), cu(w.mu)
    ref_d, ref_j, ref_h = nanv(ev.defect_dim), nanv(ev.jacobian_nnz), nanv(ev.hessian_nnz)
    ev.eval_all_device(dZ, dmu, ref_d, ref_j, ref_h)                       # plain call: the reference
    launches_full = ev.kernel_launches
    # Ipopt's order on a new iterate: eval_g(new_x = 1), eval_jac_g(0), eval_h(0, new multipliers)
    d, j, h = nanv(ev.defect_dim), nanv(ev.jacobian_nnz), nanv(ev.hessian_nnz)
    ev.eval_all_device_x(dZ, None, True, d, None, None)
    l0 = ev.kernel_launches
    ev.eval_all_device_x(dZ, None, False, d, j, None)
    l1 = ev.kernel_launches
    ev.eval_all_device_x(dZ, dmu, False, d, None, h)
    l2 = ev.kernel_launches
    assert torch.equal(d, ref_d) and torch.equal(j, ref_j) and torch.equal(h, ref_h)
    assert (l1 - l0) + (l2 - l1) < launches_full + 8, "the cached calls must not redo the whole pass"
    # new multipliers on the same point
    mu2 = cu(2.0 * w.mu)
    h2, h2ref = nanv(ev.hessian_nnz), nanv(ev.hessian_nnz)
    ev.eval_all_device_x(dZ, mu2, False, d, None, h2)
    ev.eval_all_device(dZ, mu2, d, None, h2ref)
    assert torch.equal(h2, h2ref) and not torch.equal(h2, ref_h)
    # a Jacobian request served from the arena the Hessian call left behind
    j3 = nanv(ev.jacobian_nnz)
    ev.eval_all_device_x(dZ, None, False, d, j3, None)
    assert torch.equal(j3, ref_j)
    # a new point (new_x = 1) is recomputed: compare with a fresh handle
    Z2 = Z1.copy(); Z2.reshape(L.N, L.z_dim)[:, L.u_off:L.u_off + L.m] *= 1.5
    dZ2 = cu(Z2)
    ev.eval_all_device_x(dZ2, dmu, True, d, j, h)
    fresh = Evaluator(L, w.G_drift, w.G_drives)
    fd, fj, fh = nanv(ev.defect_dim), nanv(ev.jacobian_nnz), nanv(ev.hessian_nnz)
    fresh.eval_all_device(dZ2, dmu, fd, fj, fh)
    assert torch.equal(d, fd) and torch.equal(j, fj) and torch.equal(h, fh) and not torch.equal(j, ref_j)
    # and against the independent port
    OL = oracle_layout(L)
    pd_, pj, ph = P.evaluate(OL, w.G_drift, w.G_drives, Z2, w.mu)
    assert rel_err(j.cpu().numpy(), pj) < TOL and rel_err(h.cpu().numpy(), ph) < TOL
    # small-d path: the flag is accepted and the call recomputes
    ws = workloads.make("cfg2", N=6)
    es = Evaluator(ws.layout, ws.G_drift, ws.G_drives)
    sd, sj, sh = nanv(es.defect_dim), nanv(es.jacobian_nnz), nanv(es.hessian_nnz)
    es.eval_all_device_x(cu(ws.Z), cu(ws.mu), False, sd, sj, sh)
    r = es.eval_all(ws.Z, ws.mu)
    assert np.array_equal(sj.cpu().numpy(), r[1]) and np.array_equal(sh.cpu().numpy(), r[2])
    ev.close(); fresh.close(); es.close()


def test_host_route_with_pinned_buffers_takes_direct_panels_and_matches(built_libs, monkeypatch):
    """Host-buffer calls on the large-d path: with pinned result buffers the long plain runs of the COO vectors
    (u / dt blocks) are DMA'd straight into place, with pageable ones everything goes through the staging
    buffers and the host cores; both give the same bytes, and the direct form moves fewer of them twice."""
    import torch
    w = workloads.make("cfg5", N=12)
    L = w.layout
    ev = Evaluator(L, w.G_drift, w.G_drives)
    d0, j0, h0 = ev.eval_all(w.Z, w.mu)                              # numpy: pageable
    staged_bytes = ev.transfer_bytes()[1]
    pin = lambda n: torch.full((n,), float("nan"), dtype=torch.float64).pin_memory()
    pd_, pj, ph = pin(ev.defect_dim), pin(ev.jacobian_nnz), pin(ev.hessian_nnz)
    hZ, hmu = torch.from_numpy(w.Z).pin_memory(), torch.from_numpy(w.mu).pin_memory()
    ev.eval_all_raw(hZ.data_ptr(), hmu.data_ptr(), pd_.data_ptr(), pj.data_ptr(), ph.data_ptr())
    assert np.array_equal(pd_.numpy(), d0) and np.array_equal(pj.numpy(), j0) and np.array_equal(ph.numpy(), h0)
    assert staged_bytes - 4096 <= ev.transfer_bytes()[1] <= staged_bytes    # the same bytes over PCIe (unused arena entries aside), other destinations
    # Jacobian only / Hessian only, and the switch that disables the direct form
    pj.fill_(float("nan")); ph.fill_(float("nan"))
    ev.eval_all_raw(hZ.data_ptr(), None, None, pj.data_ptr(), None)
    ev.eval_all_raw(hZ.data_ptr(), hmu.data_ptr(), None, None, ph.data_ptr())
    assert np.array_equal(pj.numpy(), j0) and np.array_equal(ph.numpy(), h0)
    # an ordinary numpy array, page-locked through the library, takes the direct form as well
    jn = np.full(ev.jacobian_nnz, np.nan)
    lib = _lib.load()
    assert lib.qcb_host_register(jn.ctypes.data, jn.nbytes) == 0 and lib.qcb_host_register(jn.ctypes.data, jn.nbytes) == 0
    ev.eval_all_raw(hZ.data_ptr(), None, None, jn.ctypes.data, None)
    lib.qcb_host_unregister(jn.ctypes.data)
    assert np.array_equal(jn, j0)
    monkeypatch.setenv("QCB_E2E_NO_DIRECT_PANELS", "1")
    ev2 = Evaluator(L, w.G_drift, w.G_drives)
    pj.fill_(float("nan"))
    ev2.eval_all_raw(hZ.data_ptr(), hmu.data_ptr(), pd_.data_ptr(), pj.data_ptr(), ph.data_ptr())
    assert np.array_equal(pj.numpy(), j0)
    ev.close(); ev2.close()


def test_new_x_reuse_through_the_host_buffer_calls(built_libs):
    """qcb_eval_defect_x / jacobian_x / hessian_x: Ipopt's callback order on one iterate through HOST buffers
    (what the Julia shim calls) equals the recomputing calls bit for bit, skips the upload of Z and most of
    the kernels, and a new iterate is recomputed."""
    w = workloads.make("cfg5", N=9)
    L = w.layout
    ev = Evaluator(L, w.G_drift, w.G_drives)
    d0, j0, h0 = ev.eval_all(w.Z, w.mu)
    l_full = ev.kernel_launches
    d, j, h = np.full_like(d0, np.nan), np.full_like(j0, np.nan), np.full_like(h0, np.nan)
    ev.eval_all_x(w.Z, None, True, delta=d)
    a = ev.kernel_launches
    ev.eval_all_x(w.Z, None, False, jac=j)
    assert ev.transfer_bytes()[0] == 0                                # Z stayed on the device
    b = ev.kernel_launches
    ev.eval_all_x(w.Z, w.mu, False, hess=h)
    c = ev.kernel_launches
    assert np.array_equal(d, d0) and np.array_equal(j, j0) and np.array_equal(h, h0)
    assert (c - a) < (l_full - 0) and (b - a) > 0 and (c - b) > 0
    # the three callbacks together launch about what one fused pass does (not three passes)
    assert (c - l_full) < 1.35 * l_full
    # new multipliers, same point
    h2 = np.full_like(h0, np.nan)
    ev.eval_all_x(w.Z, 3.0 * w.mu, False, hess=h2)
    assert rel_err(h2, 3.0 * h0) < 1e-13
    # a new point: new_x = 1 recomputes; a stale new_x = 0 after a device-buffer call falls back to recomputing
    Z2 = w.Z.copy(); Z2.reshape(L.N, L.z_dim)[:, L.u_off:L.u_off + L.m] *= 0.5
    ev.eval_all_x(Z2, w.mu, True, delta=d, jac=j, hess=h)
    fresh = Evaluator(L, w.G_drift, w.G_drives)
    fd, fj, fh = fresh.eval_all(Z2, w.mu)
    assert np.array_equal(d, fd) and np.array_equal(j, fj) and np.array_equal(h, fh)
    import torch
    dZ = torch.from_numpy(w.Z).cuda(); dd = torch.empty(ev.defect_dim, dtype=torch.float64, device="cuda")
    ev.eval_all_device(dZ, None, dd, None, None)                     # the scratch now belongs to another Z
    ev.eval_all_x(Z2, None, False, jac=j)                             # promise cannot be honoured: recompute
    assert np.array_equal(j, fj)
    ev.close(); fresh.close()


def test_host_thread_choice_settles_and_does_not_change_results(built_libs):
    """Large-d host-buffer calls time themselves over their first nine calls and keep the fastest number of
    expansion threads (qcb_host_threads_in_use): every call of the exploration returns the same bits."""
    w = workloads.make("cfg5", N=4)
    ev = Evaluator(w.layout, w.G_drift, w.G_drives)
    lib = qcb200.load_library()
    pool = lib.qcb_host_threads()
    assert lib.qcb_host_threads_in_use(ev.h, 1) == pool
    ref = ev.eval_all(w.Z, w.mu)
    for _ in range(9):
        out = ev.eval_all(w.Z, w.mu)
        assert all(np.array_equal(a, b) for a, b in zip(out, ref))
    allowed = {pool} if pool < 8 else {pool, max(2, pool * 6 // 8), max(2, pool * 4 // 8), max(2, pool * 3 // 8)}
    assert lib.qcb_host_threads_in_use(ev.h, 1) in allowed
    assert lib.qcb_host_threads_in_use(ev.h, 0) == pool          # no Jacobian-only call has been made
    ev.close()
