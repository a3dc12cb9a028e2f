"""Generates tests/golden/oracle_small.npz from the scipy oracle (oracle/qc_oracle.py).

No Julia exists in this image, so these are NOT outputs of the reference evaluator; they freeze
the independent scipy oracle (Pade expm + expm_frechet + block-triangular second-order terms) so
that the C port and the CUDA library are checked against fixed numbers across rounds.
Run from the repo root:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import qcb200  # noqa: E402
from qcb200 import workloads  # noqa: E402
import qc_oracle as O  # noqa: E402

CASES = [("cfg1_drift", 5, None), ("cfg2", 4, None), ("cfg3_ket", 4, None), ("cfg3", 3, None), ("cfg4", 4, 3)]

out = {}
for name, N, S in CASES:
    w = workloads.make(name, N=N, S=S, seed=7)
    L = w.layout
    OL = O.Layout(L.d, L.c, L.m, L.N, L.S, L.z_dim, L.x_off, L.u_off, L.dt_off, L.jac_mode)
    d, j, h = O.evaluate(OL, w.G_drift, w.G_drives, w.Z, w.mu)
    out[f"{name}.Z"], out[f"{name}.mu"] = w.Z, w.mu
    out[f"{name}.delta"], out[f"{name}.jac"], out[f"{name}.hess"] = d, j, h
    r, c = OL.jacobian_structure()
    out[f"{name}.jac_rows"], out[f"{name}.jac_cols"] = r.astype(np.int32), c.astype(np.int32)
    r, c = OL.hessian_structure()
    out[f"{name}.hess_rows"], out[f"{name}.hess_cols"] = r.astype(np.int32), c.astype(np.int32)
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "oracle_small.npz"), **out)
print("wrote", len(out), "arrays")
