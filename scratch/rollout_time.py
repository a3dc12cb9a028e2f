"""Wall time of qcb_rollout on cfg5 (N = 2000): parallel prefix over the knots vs the sequential chain kernel."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import qcb200
from qcb200 import workloads, Evaluator
w = workloads.make("cfg5"); ev = Evaluator(w.layout, w.G_drift, w.G_drives)
for mode in ("prefix", "chain"):
    if mode == "chain": os.environ["QCB_ROLLOUT_CHAIN"] = "1"
    ev.rollout(w.Z)
    t = time.perf_counter(); s = ev.rollout(w.Z); dt = time.perf_counter() - t
    print(f"rollout cfg5 N=2000 {mode}: {1e3 * dt:.1f} ms (host to host, 131 MB of states returned)", flush=True)
    if mode == "prefix": ref = s
print("max rel diff prefix vs chain", float(np.abs(ref - s).max() / np.abs(s).max()))
