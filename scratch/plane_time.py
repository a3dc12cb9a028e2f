"""torchrun --nproc-per-node N scratch/plane_time.py : ArenaDelivery step time for several root shares."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
import qcb200
from qcb200 import workloads, Evaluator, distributed as qd
rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
w = workloads.make("cfg5"); L = w.layout; K = L.K
dZ = torch.from_numpy(w.Z).to(dev) if rank == 0 else torch.empty(len(w.Z), dtype=torch.float64, device=dev)
dmu = torch.from_numpy(w.mu).to(dev) if rank == 0 else torch.empty(len(w.mu), dtype=torch.float64, device=dev)
def weighted_bounds(K, rank, world, a):
    """Smaller slab on the root (it also assembles): n0 + a (K - n0) = (K - n0) / (world - 1).  Measured slower."""
    w_ = world - 1
    n0 = max(1, min(int(round(K * (1.0 / w_ - a) / (1.0 - a + 1.0 / w_))), K - w_))
    if rank == 0:
        return 0, n0
    rest = K - n0
    return n0 + (rest * (rank - 1)) // w_, n0 + (rest * rank) // w_
def mx(x):
    t = torch.tensor([x], dtype=torch.float64, device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX); return float(t.item())
for cost in [float(c) for c in os.environ.get("QCB_COSTS", "0.0,0.03,0.06,0.095").split(",")]:
    k0, k1 = weighted_bounds(K, rank, world, cost) if cost > 0 else qd.slab_bounds(K, rank, world)
    ev = Evaluator(L, w.G_drift, w.G_drives, device=local, k_begin=k0, k_end=k1)
    ad = qd.ArenaDelivery(ev, L, local)
    if os.environ.get("QCB_PLANE_LOCAL") and rank != 0:       # diagnostic: arenas stay on this GPU (no NVLink traffic)
        ad._loc = [torch.empty(K * ad.JA, dtype=torch.float64, device=dev), torch.empty(K * ad.HA, dtype=torch.float64, device=dev)]
        ad.buf.base[3], ad.buf.base[4] = ad._loc[0].data_ptr(), ad._loc[1].data_ptr()
    def step():
        ad.broadcast(dZ, dmu); ad.evaluate(dZ, dmu)
    for _ in range(3): step()
    dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(5): step()
    torch.cuda.synchronize()
    dt = mx(time.perf_counter() - t0) / 5
    if os.environ.get("QCB_PLANE_TRACE"):
        ad.trace = {}
        for _ in range(5): step()
        if rank in (0, 1, world - 1):
            print(f"rank {rank} stages (ms after evaluate() starts): " + ", ".join(f"{k} {1e3 * np.mean(v):.2f}" for k, v in ad.trace.items()), flush=True)
        ad.trace = None
    if rank == 0:
        print(f"world {world} assemble_cost {cost}: root units {k1 - k0}  {1e3 * dt:.3f} ms/step  {K / dt:.0f} knot-evals/s", flush=True)
    dist.barrier(); ad.close(); ev.close()
dist.destroy_process_group()
