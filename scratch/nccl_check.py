"""torchrun --nproc-per-node N scratch/nccl_check.py : SlabCoordinator (NCCL broadcast + gather) vs one handle."""
import os, sys, numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qcb200
from qcb200 import workloads, Evaluator, distributed
rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
for name, N, S in [("cfg2", 41, None), ("cfg4", 13, 5), ("cfg5", 33, None)]:
    w = workloads.make(name, N=N, S=S); L = w.layout
    k0, k1 = distributed.slab_bounds(L.K, rank, world)
    ev = Evaluator(L, w.G_drift, w.G_drives, device=local, k_begin=k0, k_end=k1)
    co = distributed.SlabCoordinator(L, distributed.gpu_compute(ev), dev)
    if rank == 0:
        out = co.evaluate(torch.from_numpy(w.Z).to(dev), torch.from_numpy(w.mu).to(dev))
        full = Evaluator(L, w.G_drift, w.G_drives, device=local)
        d, j, h = full.eval_all(w.Z, w.mu)
        rel = lambda a, b: float(np.abs(a.cpu().numpy() - b).max() / np.abs(b).max())
        print(name, "world", world, "delta", rel(out[0], d), "jac", rel(out[1], j), "hess", rel(out[2], h), flush=True)
        assert max(rel(out[0], d), rel(out[1], j), rel(out[2], h)) < 1e-12
    else:
        co.evaluate()
    dist.barrier()

# ---- peer delivery (CUDA IPC): slabs stored straight into rank 0's HBM by the producing kernels ----
for name, N in [("cfg2", 41), ("cfg5", 33)]:
    w = workloads.make(name, N=N); L = w.layout
    k0, k1 = distributed.slab_bounds(L.K, rank, world)
    ev = Evaluator(L, w.G_drift, w.G_drives, device=local, k_begin=k0, k_end=k1)
    pd = distributed.PeerDelivery(ev, L, local)
    dZ, dmu = torch.from_numpy(w.Z).to(dev), torch.from_numpy(w.mu).to(dev)
    pd.evaluate(dZ, dmu)
    if rank == 0:
        got = pd.fetch()
        full = Evaluator(L, w.G_drift, w.G_drives, device=local)
        ref = full.eval_all(w.Z, w.mu)
        print(name, "peer delivery world", world, [float(np.abs(a - b).max() / np.abs(b).max()) for a, b in zip(got, ref)], flush=True)
        assert all(np.abs(a - b).max() <= 1e-12 * np.abs(b).max() for a, b in zip(got, ref))
    dist.barrier()
    pd.close()
dist.destroy_process_group()
