"""Device-resident timing of the small-d configs (cfg1-4): python scratch/small_time.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import qcb200
from qcb200 import workloads, Evaluator
dev = torch.device("cuda", 0)
flush = torch.empty(64 * 1024 * 1024, dtype=torch.float64, device=dev)
names = os.environ.get("QCB_SMALL_CFGS", "cfg1,cfg2,cfg3,cfg3_ket,cfg4").split(",")
reps = int(os.environ.get("QCB_SMALL_REPS", "10"))
for nm in names:
    w = workloads.make(nm); L = w.layout
    ev = Evaluator(L, w.G_drift, w.G_drives)
    Z, mu = torch.from_numpy(w.Z).to(dev), torch.from_numpy(w.mu).to(dev)
    d = torch.empty(ev.defect_dim, dtype=torch.float64, device=dev)
    j = torch.empty(ev.jacobian_nnz, dtype=torch.float64, device=dev)
    h = torch.empty(ev.hessian_nnz, dtype=torch.float64, device=dev)
    st = torch.cuda.ExternalStream(ev.stream, device=dev)
    for _ in range(3): ev.eval_all_device(Z, mu, d, j, h)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ts = []
    for _ in range(reps):
        flush.add_(1.0); torch.cuda.synchronize()
        a.record(st); ev.eval_all_device(Z, mu, d, j, h); b.record(st); b.synchronize()
        ts.append(a.elapsed_time(b))
    line = f"{nm:9s} d={L.d:3d} units={L.S*L.K:7d}  all {np.median(ts):.4f} ms (min {min(ts):.4f})"
    for label, args in (("defect", (Z, None, d, None, None)), ("jac", (Z, None, d, j, None)), ("hess", (Z, mu, d, None, h))):
        ts = []
        for _ in range(max(3, reps // 2)):
            flush.add_(1.0); torch.cuda.synchronize()
            a.record(st); ev.eval_all_device(*args); b.record(st); b.synchronize()
            ts.append(a.elapsed_time(b))
        line += f"  {label} {np.median(ts):.4f}"
    print(line, flush=True)
    ev.close()
