"""CPU tests of the host-side logic: layouts, workloads, trajectory containers, slab partition and
the multi-rank broadcast/gather driver (gloo, world_size 2, CPU stand-in for the GPU evaluator)."""
import os
import socket

import numpy as np
import pytest

import qcb200
from qcb200 import KnotLayout, workloads, piccolo, distributed, isomorphisms as iso
import qc_oracle as O
import qc_port as P
from conftest import oracle_layout, rel_err


def test_smooth_pulse_layout_offsets():
    L = KnotLayout.smooth_pulse(6, 3, 2, 200)
    assert (L.x_dim, L.K, L.x_off, L.dt_off, L.u_off, L.z_dim) == (18, 199, [0], 19, 20, 26)
    Ls = KnotLayout.smooth_pulse(4, 2, 2, 400, S=256, with_derivatives=False)
    assert Ls.x_off[255] == 255 * 8 and Ls.z_dim == 256 * 8 + 2 + 2
    assert L.jac_nnz_unit == 18 * 18 + 18 * 4 and L.hess_nnz_unit == 18 * 3 + 3 + 2 + 1


@pytest.mark.parametrize("name,units,d,c,m", [("cfg1", 49, 4, 2, 2), ("cfg2", 199, 6, 3, 2), ("cfg3", 499, 18, 9, 4),
                                              ("cfg3_ket", 499, 18, 1, 4)])
def test_workload_shapes(name, units, d, c, m):
    w = workloads.make(name)
    L = w.layout
    assert (w.units, L.d, L.c, L.m) == (units, d, c, m)
    assert w.Z.shape == (L.N * L.z_dim,) and w.mu.shape == (L.S * L.K * L.x_dim,)
    assert np.array_equal(workloads.make(name).Z, w.Z)          # deterministic
    G = w.G_drift[0]
    assert np.abs(G + G.T).max() < 1e-14                         # closed system: skew-symmetric
    # states are unitary / normalised
    x0 = w.Z[:L.x_dim]
    if c > 1:
        U = iso.iso_vec_to_operator(x0)
        assert np.abs(U.conj().T @ U - np.eye(d // 2)).max() < 1e-12
    else:
        assert abs(np.linalg.norm(x0) - 1) < 1e-12


def test_cfg4_and_cfg5_descriptors():
    w4 = workloads.make("cfg4", N=3)
    assert w4.layout.S == 256 and w4.units == 512
    w5 = workloads.make("cfg5", N=3)
    assert (w5.layout.d, w5.layout.c, w5.layout.m, w5.layout.jac_mode) == (128, 64, 12, 1)
    assert w5.layout.jac_nnz_unit == 1163264 and w5.layout.hess_nnz_unit == 106587


def test_isomorphism_roundtrip():
    rng = np.random.default_rng(0)
    U = workloads.haar_unitary(5, rng)
    assert np.abs(iso.iso_vec_to_operator(iso.operator_to_iso_vec(U)) - U).max() < 1e-15
    psi = rng.normal(size=4) + 1j * rng.normal(size=4)
    assert np.abs(iso.iso_to_ket(iso.ket_to_iso(psi)) - psi).max() == 0
    H = rng.normal(size=(3, 3)) + 1j * rng.normal(size=(3, 3)); H = H + H.conj().T
    # G(H) acts on iso kets like -iH on kets
    assert np.abs(iso.G_of_H(H) @ iso.ket_to_iso(psi[:3]) - iso.ket_to_iso(-1j * H @ psi[:3])).max() < 1e-13


def test_named_trajectory_layout_and_integrator_descriptor():
    sys_ = piccolo.QuantumSystem(0.1 * workloads.Z2, [workloads.X2, workloads.Y2])
    N = 11
    traj = piccolo.NamedTrajectory({"U_iso": np.zeros((8, N)), "t": np.zeros((1, N)), "dt": np.full((1, N), 0.2),
                                    "u": np.zeros((2, N)), "du": np.zeros((2, N))}, timestep="dt")
    assert traj.dim == 14 and traj.datavec.shape == (14 * N,)
    assert list(traj.components["u"]) == [10, 11]
    traj.update("U_iso", np.arange(8)[:, None] * np.ones((1, N)))
    assert np.array_equal(traj.datavec[:8], np.arange(8))           # knot-major
    L = piccolo._layout_from_traj(traj, ["U_iso"], "u", 4, 2, None)
    assert (L.x_off, L.u_off, L.dt_off, L.z_dim, L.m) == ([0], 10, 9, 14, 2)
    with pytest.raises(ValueError):
        piccolo._layout_from_traj(traj, ["u"], "u", 4, 2, None)


def test_slab_bounds_partition():
    for K, world in [(1999, 8), (49, 4), (7, 7), (499, 3)]:
        b = [distributed.slab_bounds(K, r, world) for r in range(world)]
        assert b[0][0] == 0 and b[-1][1] == K
        assert all(b[i][1] == b[i + 1][0] for i in range(world - 1))
        sizes = [y - x for x, y in b]
        assert min(sizes) >= 1 and max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        distributed.slab_bounds(3, 0, 4)


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _gloo_worker(rank, world, port, name, N, S, ret):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    w = workloads.make(name, N=N, S=S)
    L = w.layout
    OL = oracle_layout(L)

    def compute(k0, k1, Z, mu):     # CPU stand-in for the GPU evaluator (test infrastructure)
        d, j, h = P.evaluate(OL, w.G_drift, w.G_drives, Z.numpy(), None if mu is None else mu.numpy(), k_range=(k0, k1))
        return torch.from_numpy(d), torch.from_numpy(j), (None if h is None else torch.from_numpy(h))

    co = distributed.SlabCoordinator(L, compute, torch.device("cpu"))
    if rank == 0:
        out = co.evaluate(torch.from_numpy(w.Z.copy()), torch.from_numpy(w.mu.copy()))
        d0, j0, h0 = P.evaluate(OL, w.G_drift, w.G_drives, w.Z, w.mu)
        ret["err"] = max(rel_err(out[0].numpy(), d0), rel_err(out[1].numpy(), j0), rel_err(out[2].numpy(), h0))
    else:
        assert co.evaluate() == (None, None, None)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("name,N,S", [("cfg1_drift", 8, None), ("cfg4", 6, 3)])
def test_slab_coordinator_two_ranks_gloo(name, N, S, built_libs):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, name, N, S, ret)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert ret["err"] == 0.0        # slabs are evaluated independently: bit-identical to one pass


# ---- the pipelined compact delivery (ArenaDelivery) as a protocol, on CPU: gloo, two ranks ----------------
class _ShmBuffers:
    """Stand-in for the root GPU's buffers: one shared-memory block, `base` = addresses inside it."""

    def __init__(self, sizes, device_index, root, group):
        import torch.distributed as dist
        from multiprocessing import shared_memory
        self.rank, self.root, self.sizes = dist.get_rank(group), root, list(sizes)
        name = [f"qcb200_test_{os.getpid()}" if self.rank == root else None]
        total = 8 * (sum(self.sizes) + 16)
        if self.rank == root:
            self.shm = shared_memory.SharedMemory(create=True, size=total, name=name[0])
        dist.broadcast_object_list(name, src=root, group=group)
        if self.rank != root:
            self.shm = shared_memory.SharedMemory(name=name[0])
            try:
                from multiprocessing import resource_tracker
                resource_tracker.unregister(self.shm._name, "shared_memory")
            except Exception:
                pass
        self.arr = np.ndarray((total // 8,), dtype=np.float64, buffer=self.shm.buf)
        if self.rank == root:
            self.arr[:] = np.nan
        dist.barrier(group=group)
        a0 = self.arr.ctypes.data
        self.base = [a0 + 8 * int(sum(self.sizes[:i])) for i in range(len(self.sizes))]

    def view(self, addr, n):
        off = (addr - self.arr.ctypes.data) // 8
        return self.arr[off:off + n]

    def fetch(self, i):
        return self.view(self.base[i], self.sizes[i]).copy()

    def close(self):
        del self.arr
        self.shm.close()
        if self.rank == self.root:
            self.shm.unlink()


class _StubEvaluator:
    """Speaks api.Evaluator's delivery interface with the C port behind it (test infrastructure): arenas and
    COO vectors are written at raw addresses, the call order is logged."""

    def __init__(self, w, k0, k1, bufs_getter):
        self.w, self.L, self.OL = w, w.layout, oracle_layout(w.layout)
        self.k_begin, self.k_end = k0, k1
        self.parts, self.log, self.bufs = 3, [], bufs_getter
        self.ji, self.js, self.hi = P.arena_codes(self.OL, 0)
        self.JA = int(P.lib().qc_port_jac_arena(self.L.d, self.L.c, self.L.m))
        self.HA = int(P.lib().qc_port_hess_arena(self.L.d, self.L.c, self.L.m))
        self.e_part = self.L.d * self.L.d

    def arena_doubles(self):
        return self.JA, self.HA

    def _arenas(self, Z, mu):
        return P.evaluate_arenas(self.OL, self.w.G_drift, self.w.G_drives, Z.numpy(), None if mu is None else mu.numpy(),
                                 k_range=(self.k_begin, self.k_end))

    def _put(self, addr, values):
        self.bufs().view(addr, len(values))[:] = values

    def eval_all_device(self, Z, mu, pd, pj, ph):
        d, ja, ha = self._arenas(Z, mu)
        self._put(pd, d)
        if pj is not None:
            self._put(pj, (ja[:, self.ji] * self.js).reshape(-1))
        if ph is not None and ha is not None:
            self._put(ph, ha[:, self.hi].reshape(-1))

    def eval_all_device_x(self, Z, mu, new_x, pd, pj, ph):
        self.log.append(("all", bool(new_x), pj is not None, ph is not None))
        self.eval_all_device(Z, mu, pd, pj, ph)

    def eval_arenas_device(self, Z, mu, pd, aj, ah):
        d, ja, ha = self._arenas(Z, mu)
        self._put(pd, d)
        nk = self.k_end - self.k_begin
        if aj is not None:
            dst = self.bufs().view(aj, nk * self.JA).reshape(nk, self.JA)
            if self.parts & 1:
                dst[:, :self.e_part] = ja[:, :self.e_part]
            if self.parts & 2:
                dst[:, self.e_part:] = ja[:, self.e_part:]
        if ah is not None and ha is not None:
            self._put(ah, ha.reshape(-1))

    def eval_arenas_device_x(self, Z, mu, new_x, pd, aj, ah):
        self.log.append(("arenas", bool(new_x), self.parts, aj is not None, ah is not None))
        self.eval_arenas_device(Z, mu, pd, aj, ah)

    def assemble_device(self, aj, ah, n, nk_src, pj, ph, parts=3):
        self.log.append(("assemble", aj is not None, ah is not None, parts))
        if aj is not None:
            ja = self.bufs().view(aj, n * self.JA).reshape(n, self.JA)
            out = self.bufs().view(pj, n * self.L.jac_nnz_unit).reshape(n, -1)
            head = self.ji < self.e_part
            sel = head if parts == 1 else (~head if parts == 2 else np.ones_like(head))
            out[:, sel] = (ja[:, self.ji] * self.js)[:, sel]
        if ah is not None:
            ha = self.bufs().view(ah, n * self.HA).reshape(n, self.HA)
            self.bufs().view(ph, n * self.L.hess_nnz_unit).reshape(n, -1)[:] = ha[:, self.hi]

    def set_async(self, on):
        self.log.append(("async", bool(on)))

    def set_arena_parts(self, parts):
        self.parts = parts

    def wait(self, what=0):
        self.log.append(("wait", what))


def _arena_worker(rank, world, port, ret):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    w = workloads.make("cfg2", N=12, jac_mode=qcb200.JAC_BLOCK_DIAG)
    L = w.layout
    k0, k1 = distributed.slab_bounds(L.K, rank, world)
    holder = {}
    ev = _StubEvaluator(w, k0, k1, lambda: holder["ad"].buf)
    ad = holder["ad"] = distributed.ArenaDelivery(ev, L, 0, buffers=_ShmBuffers)
    Z = torch.from_numpy(w.Z.copy()) if rank == 0 else torch.zeros(len(w.Z), dtype=torch.float64)
    mu = torch.from_numpy(w.mu.copy()) if rank == 0 else torch.zeros(len(w.mu), dtype=torch.float64)
    ad.broadcast(Z, mu)
    errs = []
    for pipelined in (True, False):
        ev.log.clear()
        ad.evaluate(Z, mu, pipelined=pipelined)
        dist.barrier()
        if rank == 0:
            ref = P.evaluate(oracle_layout(L), w.G_drift, w.G_drives, w.Z, w.mu)
            errs.append(max(float(np.abs(a - b).max()) for a, b in zip(ad.fetch(), ref)))
            ad.buf.arr[:] = np.nan
        if pipelined:
            ret[f"log{rank}"] = list(ev.log)
        dist.barrier()
    if rank == 0:
        ret["errs"] = errs
    ad.close()
    dist.destroy_process_group()


def test_arena_delivery_protocol_two_ranks_gloo(built_libs):
    """Host logic of the pipelined compact delivery on CPU: the phases are queued in order, the root
    assembles foreign slabs part by part after the matching barriers, and the vectors that end up in the
    root's buffers equal a single-pass evaluation bit for bit (C port behind a stub evaluator)."""
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=_arena_worker, args=(r, 2, port, ret)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
        assert p.exitcode == 0
    assert ret["errs"] == [0.0, 0.0]
    root, other = ret["log0"], ret["log1"]
    # other rank: async on, propagators (E part) with new_x = 1, then the rest of the Jacobian arena and the
    # Hessian side with new_x = 0, waits in delivery order, async off
    assert other[0] == ("async", True) and other[-1] == ("async", False)
    calls = [c for c in other if c[0] == "arenas"]
    assert calls == [("arenas", True, 1, True, False), ("arenas", False, 2, True, False), ("arenas", False, 3, False, True)]
    assert [c[1] for c in other if c[0] == "wait"] == [1, 2, 0, 0]
    # root: own slab straight into COO (Jacobian side first), then assembly of the foreign slab in three parts
    assert [c for c in root if c[0] == "all"] == [("all", True, True, False), ("all", False, False, True)]
    assert [c for c in root if c[0] == "assemble"] == [("assemble", True, False, 1), ("assemble", True, False, 2), ("assemble", False, True, 3)]
