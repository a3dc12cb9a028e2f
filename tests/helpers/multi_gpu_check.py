"""Run under torchrun (one rank per GPU) by tests/test_gpu_multi.py: every multi-GPU driver of
qcb200.distributed against a single-handle evaluation of the same trajectory, bit for bit."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import qcb200  # noqa: E402
from qcb200 import Evaluator, distributed, workloads  # noqa: E402

rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)


def reference(w):
    full = Evaluator(w.layout, w.G_drift, w.G_drives, device=local)
    out = full.eval_all(w.Z, w.mu)
    full.close()
    return out


def same(got, ref, what):
    for nm, a, b in zip(("delta", "jac", "hess"), got, ref):
        a = a.cpu().numpy() if hasattr(a, "cpu") else a
        err = float(np.abs(a - b).max() / np.abs(b).max())
        assert err < 1e-13, (what, nm, err)
    print(f"ok {what} world={world}", flush=True)


# --- what north_star prescribes: NCCL broadcast of Z, mu + gather of the value slabs -------------------
for name, N, S in [("cfg2", 41, None), ("cfg4", 13, 5), ("cfg5", 21, None)]:
    w = workloads.make(name, N=N, S=S)
    L = w.layout
    k0, k1 = distributed.slab_bounds(L.K, rank, world)
    ev = Evaluator(L, w.G_drift, w.G_drives, device=local, k_begin=k0, k_end=k1)
    co = distributed.SlabCoordinator(L, distributed.gpu_compute(ev), dev)
    if rank == 0:
        out = co.evaluate(torch.from_numpy(w.Z).to(dev), torch.from_numpy(w.mu).to(dev))
        same(out, reference(w), f"SlabCoordinator {name}")
    else:
        co.evaluate()
    dist.barrier()
    ev.close()

# --- gather fused into the producing kernels: COO slabs (PeerDelivery), compact arenas (ArenaDelivery) ---
for cls, cases in ((distributed.PeerDelivery, [("cfg2", 41), ("cfg5", 21)]), (distributed.ArenaDelivery, [("cfg5", 21), ("cfg5", 34)])):
    for name, N in cases:
        w = workloads.make(name, N=N, seed=N)
        L = w.layout
        k0, k1 = distributed.slab_bounds(L.K, rank, world)
        ev = Evaluator(L, w.G_drift, w.G_drives, device=local, k_begin=k0, k_end=k1)
        dl = cls(ev, L, local)
        if cls is distributed.ArenaDelivery:
            # inputs exist on the root only; the broadcast is part of the driver
            dZ = torch.from_numpy(w.Z).to(dev) if rank == 0 else torch.empty(len(w.Z), dtype=torch.float64, device=dev)
            dmu = torch.from_numpy(w.mu).to(dev) if rank == 0 else torch.empty(len(w.mu), dtype=torch.float64, device=dev)
            dl.broadcast(dZ, dmu)
        else:
            dZ, dmu = torch.from_numpy(w.Z).to(dev), torch.from_numpy(w.mu).to(dev)
            torch.cuda.synchronize()
        ref = reference(w) if rank == 0 else None
        modes = (True, False) if cls is distributed.ArenaDelivery else (None,)
        for mode in modes:
            for rep in range(2):        # twice: the buffers are reused
                if mode is None:
                    dl.evaluate(dZ, dmu)
                else:
                    dl.evaluate(dZ, dmu, pipelined=mode)
            if rank == 0:
                same(dl.fetch(), ref, f"{cls.__name__} {name} N={N}" + ("" if mode is None else (" pipelined" if mode else " plain")))
            dist.barrier()
        dist.barrier()
        dl.close()
        ev.close()
dist.destroy_process_group()
