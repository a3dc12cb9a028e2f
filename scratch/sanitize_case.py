"""Small mixed case for compute-sanitizer: both paths, host + device entry points, fidelity."""
import sys, numpy as np
sys.path.insert(0, '/root/repo')
import qcb200
from qcb200 import workloads, Evaluator, Fidelity
for name, N, S, fp in [("cfg1_drift", 5, None, 0), ("cfg2", 4, None, 0), ("cfg3_ket", 3, None, 0), ("cfg4", 4, 3, 0),
                       ("cfg2", 3, None, 2), ("cfg5", 3, None, 0)]:
    w = workloads.make(name, N=N, S=S)
    ev = Evaluator(w.layout, w.G_drift, w.G_drives, force_path=fp)
    d, j, h = ev.eval_all(w.Z, w.mu)
    assert np.isfinite(d).all() and np.isfinite(j).all() and np.isfinite(h).all()
    ev.eval_defect(w.Z); ev.eval_jacobian(w.Z)
    ev.close()
f = Fidelity(Fidelity.UNITARY, workloads.H_GATE)
print(f.eval(np.arange(8.0), want_hess=True)[0])
print("sanitize case ok")
