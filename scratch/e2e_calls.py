"""Per-call wall times of the host-buffer fused evaluation on cfg5 (diagnostic; run under gpurun).
usage: python scratch/e2e_calls.py [tag]   (QCB200_LIB selects the library)"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qcb200
from qcb200 import Evaluator, workloads

tag = sys.argv[1] if len(sys.argv) > 1 else ""
w = workloads.make("cfg5")
L = w.layout
ev = Evaluator(L, w.G_drift, w.G_drives, device=0)
dev = torch.device("cuda", 0)
hZ = torch.from_numpy(w.Z).pin_memory()
hmu = torch.from_numpy(w.mu).pin_memory()
h_delta = torch.empty(ev.defect_dim, dtype=torch.float64, pin_memory=True)
h_jac = torch.empty(ev.jacobian_nnz, dtype=torch.float64, pin_memory=True)
h_hess = torch.empty(ev.hessian_nnz, dtype=torch.float64, pin_memory=True)
pZ, pmu, pd, pj, ph = (t.data_ptr() for t in (hZ, hmu, h_delta, h_jac, h_hess))
lib = qcb200.load_library()


def calls(n, what):
    out = []
    for _ in range(n):
        t0 = time.perf_counter()
        ev.eval_all_raw(pZ, pmu, pd, pj, ph)
        out.append(1e3 * (time.perf_counter() - t0))
    print(tag, what, " ".join(f"{t:.1f}" for t in out), flush=True)


calls(16, "fresh:")
print(tag, "threads in use (fused, jac only):", lib.qcb_host_threads_in_use(ev.h, 1), lib.qcb_host_threads_in_use(ev.h, 0), "of", lib.qcb_host_threads(), flush=True)
if os.environ.get("E2E_QUICK"):
    print(tag, "timings", ev.timings(), flush=True)
    sys.exit(0)
dZ, dmu = hZ.to(dev), hmu.to(dev)
d_delta = torch.empty(ev.defect_dim, dtype=torch.float64, device=dev)
d_jac = torch.empty(ev.jacobian_nnz, dtype=torch.float64, device=dev)
d_hess = torch.empty(ev.hessian_nnz, dtype=torch.float64, device=dev)
for _ in range(3):
    ev.eval_all_device(dZ, dmu, d_delta, d_jac, d_hess)
torch.cuda.synchronize()
calls(5, "after device fused:")
for _ in range(2):
    ev.eval_all_device_x(dZ, None, True, d_delta, None, None)
    ev.eval_all_device_x(dZ, None, False, d_delta, d_jac, None)
    ev.eval_all_device_x(dZ, dmu, False, d_delta, None, d_hess)
torch.cuda.synchronize()
calls(5, "after device new_x protocol:")
acc = []
for _ in range(3):
    t0 = time.perf_counter(); lib.qcb_eval_defect_x(ev.h, pZ, 1, pd)
    t1 = time.perf_counter(); lib.qcb_eval_jacobian_x(ev.h, pZ, 0, pj)
    t2 = time.perf_counter(); lib.qcb_eval_hessian_x(ev.h, pZ, 0, pmu, ph)
    t3 = time.perf_counter()
    acc.append((1e3 * (t1 - t0), 1e3 * (t2 - t1), 1e3 * (t3 - t2)))
print(tag, "host mix g/jac/h:", " | ".join("%.1f %.1f %.1f" % a for a in acc), flush=True)
calls(5, "after host mix:")
