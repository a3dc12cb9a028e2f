"""Generates tests/golden/oracle_cfg5.npz: every block of TWO full cfg5 units (d = 128, c = 64,
m = 12) from the independent scipy oracle (oracle/qc_oracle.py: Pade expm, expm_frechet,
block-triangular second-order terms) -- not from the C port, which shares the op tables with the
CUDA kernels.  Unit 0 has ||dt G(u)||_1 < 0.5 (no squaring), unit 1 has dt scaled so that three
squaring levels run.  Blocks stored: E, delta, J_u (12 panels), J_dt, H_xu (12 panels), H_xdt,
H_uu, H_udt, H_dtdt.  No Julia exists in this image or on the GPU box (profiles/r02_gpu_box_probe.txt),
so these are NOT outputs of the reference evaluator.
Run from the repo root (a few minutes on the CPU):  python tests/golden/make_golden_cfg5.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import qcb200  # noqa: E402,F401
from qcb200 import workloads  # noqa: E402
import qc_oracle as O  # noqa: E402

DT_SCALE = (0.8, 6.0)      # unit 0: s = 0; unit 1: s = 3


def inputs():
    w = workloads.make("cfg5", N=3, seed=11)
    L = w.layout
    Z = w.Z.reshape(L.N, L.z_dim).copy()
    for k, f in enumerate(DT_SCALE):
        Z[k, L.dt_off] *= f
    return w, Z.reshape(-1)


if __name__ == "__main__":
    w, Z = inputs()
    L = w.layout
    OL = O.Layout(L.d, L.c, L.m, L.N, L.S, L.z_dim, L.x_off, L.u_off, L.dt_off, L.jac_mode)
    out = {"Z": Z, "mu": w.mu}
    for k in range(L.K):
        X, Xn, M, u, dt = OL.unit_inputs(Z, w.mu, 0, k)
        G = w.G_drift[0] + np.tensordot(u, w.G_drives[0], axes=1)
        out[f"u{k}.norm1"] = np.array(np.abs(dt * G).sum(axis=0).max())
        ev = O.unit_eval(w.G_drift[0], w.G_drives[0], X, Xn, M, u, dt, want_hessian=True)
        for key, val in ev.items():
            out[f"u{k}.{key}"] = np.asarray(val)
        print("unit", k, "norm1", float(out[f"u{k}.norm1"]), flush=True)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "oracle_cfg5.npz"), **out)
    print("wrote", len(out), "arrays")
