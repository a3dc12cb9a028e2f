"""Multi-GPU time-slab driver: one process per GPU, ``torch.distributed`` plumbing.

Units (interval k, system s) are independent given Z and mu, so the trajectory is cut into
contiguous time slabs, one per rank.  The only exchange steps are the ones the path really has:
broadcast of the updated Z (and mu) from the rank that hosts Ipopt, and the gather of the
defect / Jacobian / Hessian value slabs back to it.  With the "nccl" backend both run over
NVLink on device tensors; the "gloo" backend (CPU tensors) exists for the host-logic tests.
"""
from __future__ import annotations

from typing import Callable, Optional, Tuple

import numpy as np


def slab_bounds(K: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous, balanced partition of the K intervals; never empty when world <= K."""
    if world > K:
        raise ValueError(f"cannot cut {K} intervals into {world} slabs")
    return (K * rank) // world, (K * (rank + 1)) // world


class SlabCoordinator:
    """Broadcast inputs from ``root``, evaluate this rank's slab, gather value slabs on ``root``.

    ``compute(k0, k1, Z, mu) -> (delta, jac, hess)`` evaluates intervals [k0, k1) and returns flat
    arrays ordered [system][interval][...]; in the product it is backed by ``api.Evaluator`` on the
    rank's GPU (see ``gpu_compute``), in the CPU tests by an injected stand-in.
    Arrays are torch tensors on ``device`` (cuda for nccl, cpu for gloo).
    """

    def __init__(self, layout, compute: Callable, device, root: int = 0, group=None):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.L = layout
        self.compute = compute
        self.device = device
        self.root = root
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.k0, self.k1 = slab_bounds(layout.K, self.rank, self.world)
        self.bounds = [slab_bounds(layout.K, r, self.world) for r in range(self.world)]
        self.max_nk = max(b - a for a, b in self.bounds)

    def _gather(self, slab, per_unit: int):
        """slab: [S * nk * per_unit] on this rank -> full [S * K * per_unit] on root (None elsewhere)."""
        torch, dist, L = self.torch, self.dist, self.L
        nk = self.k1 - self.k0
        padded = torch.zeros(L.S, self.max_nk, per_unit, dtype=torch.float64, device=self.device)
        padded[:, :nk] = slab.view(L.S, nk, per_unit)
        if self.rank == self.root:
            bufs = [torch.empty_like(padded) for _ in range(self.world)]
            dist.gather(padded, bufs, dst=self.root, group=self.group)
            full = torch.empty(L.S, L.K, per_unit, dtype=torch.float64, device=self.device)
            for (a, b), buf in zip(self.bounds, bufs):
                full[:, a:b] = buf[:, :b - a]
            return full.view(-1)
        dist.gather(padded, None, dst=self.root, group=self.group)
        return None

    def evaluate(self, Z=None, mu=None, want_jac=True, want_hess=True):
        """On root pass the full Z (N*z_dim) and mu (S*K*x_dim); other ranks pass None.
        Returns (delta, jac_vals, hess_vals) on root in the single-handle order, None elsewhere."""
        torch, dist, L = self.torch, self.dist, self.L
        if self.rank != self.root:
            Z = torch.empty(L.N * L.z_dim, dtype=torch.float64, device=self.device)
            mu = torch.empty(L.S * L.K * L.x_dim, dtype=torch.float64, device=self.device) if want_hess else None
        dist.broadcast(Z, src=self.root, group=self.group)
        if want_hess:
            dist.broadcast(mu, src=self.root, group=self.group)
        delta, jac, hess = self.compute(self.k0, self.k1, Z, mu if want_hess else None)
        out_d = self._gather(delta, L.x_dim)
        out_j = self._gather(jac, L.jac_nnz_unit) if want_jac else None
        out_h = self._gather(hess, L.hess_nnz_unit) if want_hess else None
        return out_d, out_j, out_h


def gpu_compute(evaluator):
    """compute() backed by the CUDA library on this rank's device (device tensors in and out)."""
    import torch

    def compute(k0, k1, Z, mu):
        assert (k0, k1) == (evaluator.k_begin, evaluator.k_end)
        dev = Z.device
        # the library launches on its own stream: the NCCL broadcast that produced Z / mu is only
        # ordered on torch's stream, so finish it before handing the buffers over
        torch.cuda.current_stream(dev).synchronize()
        d = torch.empty(evaluator.defect_dim, dtype=torch.float64, device=dev)
        j = torch.empty(evaluator.jacobian_nnz, dtype=torch.float64, device=dev)
        h = torch.empty(evaluator.hessian_nnz, dtype=torch.float64, device=dev) if mu is not None else None
        evaluator.eval_all_device(Z, mu, d, j, h)
        return d, j, h

    return compute


class _RootBuffers:
    """Device buffers on the root GPU, mapped into every other rank's process with CUDA IPC."""

    def __init__(self, sizes, device_index, root, group):
        import ctypes
        import torch
        import torch.distributed as dist
        from . import _lib
        self.lib, self.dev, self.root = _lib.load(), device_index, root
        self.rank = dist.get_rank(group)
        self.sizes = list(sizes)
        handles = torch.zeros(len(self.sizes), 64, dtype=torch.uint8)
        self.base = []
        if self.rank == root:
            for i, n in enumerate(self.sizes):
                p = ctypes.c_void_p()
                if self.lib.qcb_device_alloc(device_index, ctypes.byref(p), max(n, 1) * 8) != 0:
                    raise MemoryError("qcb_device_alloc failed")
                buf = ctypes.create_string_buffer(64)
                assert self.lib.qcb_ipc_export(device_index, p, buf) == 0
                handles[i] = torch.frombuffer(bytearray(buf.raw), dtype=torch.uint8)
                self.base.append(p.value)
        obj = [handles.numpy().tobytes() if self.rank == root else None]
        dist.broadcast_object_list(obj, src=root, group=group)
        if self.rank != root:
            raw = obj[0]
            for i in range(len(self.sizes)):
                p = ctypes.c_void_p()
                if self.lib.qcb_ipc_open(device_index, raw[64 * i:64 * (i + 1)], ctypes.byref(p)) != 0:
                    raise RuntimeError("qcb_ipc_open failed (no peer access between the GPUs?)")
                self.base.append(p.value)

    def fetch(self, i):
        """Root only: buffer i as a host numpy array (verification)."""
        assert self.rank == self.root
        a = np.empty(self.sizes[i])
        assert self.lib.qcb_device_to_host(self.dev, a.ctypes.data, self.base[i], self.sizes[i] * 8) == 0
        return a

    def close(self):
        for p in self.base:
            if self.rank == self.root:
                self.lib.qcb_device_free(self.dev, p)
            else:
                self.lib.qcb_ipc_close(self.dev, p)
        self.base = []


class ArenaDelivery:
    """The data plane of SURVEY.md 8e with the gather fused into the producing kernels AND kept compact:
    ``broadcast`` sends Z and mu from the Ipopt rank (NCCL); every other rank's kernels store the value
    arenas of their slab (every distinct number once: ~5x fewer bytes than the COO vectors on the unitary
    path, whose Jacobian repeats -E per state column) into the Ipopt rank's HBM over NVLink; the Ipopt
    rank evaluates its own slab straight into the COO vectors and, after the barrier, assembles the
    other slabs from the delivered arenas.  S == 1 layouts on the large-d path."""

    def __init__(self, evaluator, layout, device_index: int, root: int = 0, group=None, buffers=None):
        """buffers: factory `(sizes, device_index, root, group) -> object with .base, .fetch(i), .close()`
        for the five root-side buffers (default: device memory on the root, mapped with CUDA IPC; the CPU
        tests of the protocol inject shared host memory)."""
        import torch.distributed as dist
        if layout.S != 1:
            raise ValueError("arena delivery needs S == 1")
        self.dist, self.group = dist, group
        self.ev, self.L, self.dev, self.root = evaluator, layout, device_index, root
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        L = layout
        self.JA, self.HA = evaluator.arena_doubles()
        self.bounds = [slab_bounds(L.K, r, self.world) for r in range(self.world)]
        # [0] delta, [1] Jacobian values, [2] Hessian values (the results, on the root);
        # [3] / [4] arenas of every unit the root does not evaluate itself
        self.buf = (buffers or _RootBuffers)((L.K * L.x_dim, L.K * L.jac_nnz_unit, L.K * L.hess_nnz_unit, L.K * self.JA, L.K * self.HA),
                                             device_index, root, group)
        self.k0, self.k1 = evaluator.k_begin, evaluator.k_end
        self.trace = None       # set to {} to collect host timestamps of the pipelined step's stages

    def broadcast(self, Z, mu):
        """NCCL broadcast of Z and mu (device tensors; allocated by the caller on every rank)."""
        import torch
        self.dist.broadcast(Z, src=self.root, group=self.group)
        if mu is not None:
            self.dist.broadcast(mu, src=self.root, group=self.group)
        if Z.is_cuda:
            torch.cuda.current_stream(Z.device).synchronize()     # the library runs on its own stream

    def evaluate(self, dZ, dmu, pipelined: bool = True):
        """One evaluation with every value vector complete in the root's buffers on return.

        pipelined: propagators, tangents and the Hessian side of the iterate are queued back to back as three
        calls (each reuses what the previous one computed, new_x = 0).  Other ranks: the E part of the
        Jacobian arena leaves as soon as the squarings are done, the rest of it while the Hessian side is
        computed, all on a side stream.  Root: after the first barrier it assembles the x-blocks of the foreign
        slabs (c copies of -E per unit: 90 % of the Jacobian entries on the unitary path) on its side stream
        while everybody, itself included, is still computing tangents; after the second barrier the remaining
        Jacobian entries, after the third the Hessian values."""
        L, b, k0, ev = self.L, self.buf.base, self.k0, self.ev
        root = self.rank == self.root
        pd, pj, ph = b[0] + 8 * k0 * L.x_dim, b[1] + 8 * k0 * L.jac_nnz_unit, b[2] + 8 * k0 * L.hess_nnz_unit
        aj, ah = b[3] + 8 * k0 * self.JA, b[4] + 8 * k0 * self.HA
        runs = self._foreign_runs() if root else []
        if not pipelined or dmu is None:
            if root:
                ev.eval_all_device(dZ, dmu, pd, pj, ph)
            else:
                ev.eval_arenas_device(dZ, dmu, pd, aj, ah)
            self.dist.barrier(group=self.group)
            for a, e in runs:
                ev.assemble_device(b[3] + 8 * a * self.JA, b[4] + 8 * a * self.HA if dmu is not None else None, e - a, e - a,
                                   b[1] + 8 * a * L.jac_nnz_unit, b[2] + 8 * a * L.hess_nnz_unit)
            return
        import time
        tr = self.trace
        t0 = time.perf_counter()
        mark = (lambda name: tr.setdefault(name, []).append(time.perf_counter() - t0)) if tr is not None else (lambda name: None)
        ev.set_async(True)
        try:
            if root:
                ev.eval_all_device_x(dZ, None, True, pd, pj, None)        # defect + Jacobian values of the own slab
                ev.eval_all_device_x(dZ, dmu, False, pd, None, ph)        # Hessian values, queued right behind
            else:
                ev.set_arena_parts(1)
                ev.eval_arenas_device_x(dZ, None, True, pd, aj, None)     # propagators: the E part goes ahead
                ev.set_arena_parts(2)
                ev.eval_arenas_device_x(dZ, None, False, pd, aj, None)    # tangents; the rest of the Jacobian arena
                ev.set_arena_parts(3)
                ev.eval_arenas_device_x(dZ, dmu, False, pd, None, ah)     # Hessian side overlaps that delivery
                mark("queued")
                ev.wait(1)
            mark("E shipped")
            self.dist.barrier(group=self.group)                           # every E part has landed
            mark("barrier 1")
            for a, e in runs:                                             # x-blocks: c copies of -E per unit
                ev.assemble_device(b[3] + 8 * a * self.JA, None, e - a, e - a, b[1] + 8 * a * L.jac_nnz_unit, None, parts=1)
            if not root:
                ev.wait(2)
            mark("J shipped")
            self.dist.barrier(group=self.group)                           # every Jacobian arena is complete
            mark("barrier 2")
            for a, e in runs:
                ev.assemble_device(b[3] + 8 * a * self.JA, None, e - a, e - a, b[1] + 8 * a * L.jac_nnz_unit, None, parts=2)
            ev.wait(0)
            mark("own work done")
            self.dist.barrier(group=self.group)                           # every Hessian arena has landed
            mark("barrier 3")
            for a, e in runs:
                ev.assemble_device(None, b[4] + 8 * a * self.HA, e - a, e - a, None, b[2] + 8 * a * L.hess_nnz_unit)
            ev.wait(0)
            mark("end")
        finally:
            ev.set_arena_parts(3)
            ev.set_async(False)

    def _foreign_runs(self):
        runs, K = [], self.L.K
        if self.k0 > 0:
            runs.append((0, self.k0))
        if self.k1 < K:
            runs.append((self.k1, K))
        return runs

    def fetch(self):
        return [self.buf.fetch(i) for i in range(3)]

    def close(self):
        self.buf.close()


class PeerDelivery:
    """Fused gather: every rank's kernels store their slab of the value vectors directly into buffers
    that live in the root GPU's HBM, mapped into each process with CUDA IPC (NVLink peer stores);
    no collective follows the evaluation -- a barrier is all that is left.

    Only layouts with S == 1 (a slab is one contiguous range of every output) are supported.
    """

    def __init__(self, evaluator, layout, device_index: int, root: int = 0, group=None):
        import ctypes
        import torch
        import torch.distributed as dist
        from . import _lib
        if layout.S != 1:
            raise ValueError("peer delivery needs S == 1")
        self.lib, self.dist, self.group = _lib.load(), dist, group
        self.ev, self.L, self.dev, self.root = evaluator, layout, device_index, root
        self.rank = dist.get_rank(group)
        L = layout
        self.sizes = (L.K * L.x_dim, L.K * L.jac_nnz_unit, L.K * L.hess_nnz_unit)
        handles = torch.zeros(3, 64, dtype=torch.uint8)
        self.base = []
        if self.rank == root:
            for i, n in enumerate(self.sizes):
                p = ctypes.c_void_p()
                st = self.lib.qcb_device_alloc(device_index, ctypes.byref(p), n * 8)
                if st != 0:
                    raise MemoryError("qcb_device_alloc failed")
                buf = ctypes.create_string_buffer(64)
                assert self.lib.qcb_ipc_export(device_index, p, buf) == 0
                handles[i] = torch.frombuffer(bytearray(buf.raw), dtype=torch.uint8)
                self.base.append(p.value)
        obj = [handles.numpy().tobytes() if self.rank == root else None]
        dist.broadcast_object_list(obj, src=root, group=group)
        if self.rank != root:
            raw = obj[0]
            for i in range(3):
                p = ctypes.c_void_p()
                st = self.lib.qcb_ipc_open(device_index, raw[64 * i:64 * (i + 1)], ctypes.byref(p))
                if st != 0:
                    raise RuntimeError("qcb_ipc_open failed (no peer access between the GPUs?)")
                self.base.append(p.value)
        k0 = evaluator.k_begin
        self.out = (self.base[0] + 8 * k0 * L.x_dim, self.base[1] + 8 * k0 * L.jac_nnz_unit,
                    self.base[2] + 8 * k0 * L.hess_nnz_unit)

    def evaluate(self, dZ, dmu):
        """Evaluate this rank's slab with the outputs landing in the root's buffers; returns after
        every rank is done (barrier)."""
        self.ev.eval_all_device(dZ, dmu, self.out[0], self.out[1], self.out[2])
        self.dist.barrier(group=self.group)

    def fetch(self):
        """Root only: copy the gathered vectors to host numpy arrays (verification)."""
        import numpy as np
        assert self.rank == self.root
        outs = []
        for p, n in zip(self.base, self.sizes):
            a = np.empty(n)
            assert self.lib.qcb_device_to_host(self.dev, a.ctypes.data, p, n * 8) == 0
            outs.append(a)
        return outs

    def close(self):
        for p in self.base:
            if self.rank == self.root:
                self.lib.qcb_device_free(self.dev, p)
            else:
                self.lib.qcb_ipc_close(self.dev, p)
        self.base = []
