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
