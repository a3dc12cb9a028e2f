"""Accuracy of the degree-12 Taylor + squaring program as a function of theta (scratch study)."""
import sys, os, ctypes, numpy as np
sys.path.insert(0, '.'); sys.path.insert(0, 'oracle'); sys.path.insert(0, 'tests')
import qc_port as P, qc_oracle as O
from qcb200 import workloads as wl
from conftest import oracle_layout, rel_err

def run(th, OL, w, u_scale=1.0):
    P._libs.clear()
    L = ctypes.CDLL(f'/tmp/theta/libqc_port_{th}.so')
    # reuse the binding code by faking the cache
    real = P.lib.__globals__['ctypes'].CDLL
    P.lib.__globals__['ctypes'].CDLL = lambda path: real(f'/tmp/theta/libqc_port_{th}.so')
    try:
        return P.evaluate(OL, w.G_drift, w.G_drives, w.Z, w.mu)
    finally:
        P.lib.__globals__['ctypes'].CDLL = real

for name, N, S, us in [('cfg3', 6, None, 1.0), ('cfg3', 6, None, 3.0), ('cfg5', 4, None, 1.0), ('cfg5', 4, None, 2.0), ('cfg2', 8, None, 10.0)]:
    w = wl.make(name, N=N, S=S, u_scale=us)
    OL = oracle_layout(w.layout)
    ref = O.evaluate(OL, w.G_drift, w.G_drives, w.Z, w.mu)
    L = w.layout; Z = w.Z.reshape(L.N, L.z_dim)
    nm = max(np.abs(Z[k, L.dt_off] * (w.G_drift[0] + np.tensordot(Z[k, L.u_off:L.u_off + L.m], w.G_drives[0], axes=1))).sum(axis=0).max() for k in range(L.N - 1))
    print(name, 'u_scale', us, 'max norm %.3f' % nm)
    for th in ['0.25', '0.4', '0.5', '0.7']:
        r = run(th, OL, w)
        print('   theta', th, ' '.join('%.2e' % rel_err(a, b) for a, b in zip(r, ref)))
