"""ctypes binding of the CPU port (oracle/qc_port.c).  TEST / BASELINE INFRASTRUCTURE ONLY."""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_dp = ctypes.POINTER(ctypes.c_double)
_ip = ctypes.POINTER(ctypes.c_int)


def build(native: bool = False) -> str:
    target = "libqc_port_native.so" if native else "libqc_port.so"
    subprocess.run(["make", "-s", "-C", _HERE, target], check=True)
    return os.path.join(_HERE, target)


_libs = {}


def lib(native: bool = False):
    if native not in _libs:
        path = os.path.join(_HERE, "libqc_port_native.so" if native else "libqc_port.so")
        src_m = max(os.path.getmtime(os.path.join(_HERE, "qc_port.c")),
                    os.path.getmtime(os.path.join(_HERE, "..", "quantumcollocation.jl_b200", "csrc", "qcb_program.h")))
        if not os.path.exists(path) or os.path.getmtime(path) < src_m:
            build(native)
        L = ctypes.CDLL(path)
        L.qc_port_jac_arena.restype = ctypes.c_size_t
        L.qc_port_hess_arena.restype = ctypes.c_size_t
        L.qc_port_eval.restype = ctypes.c_int
        L.qc_port_eval.argtypes = [ctypes.c_int] * 6 + [_ip, ctypes.c_int, ctypes.c_int,
                                   _dp, _dp, _dp, _dp, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                   _dp, _dp, _dp, ctypes.c_int]
        L.qc_port_eval_coo.restype = ctypes.c_int
        _up = ctypes.POINTER(ctypes.c_uint32)
        L.qc_port_eval_coo.argtypes = [ctypes.c_int] * 6 + [_ip, ctypes.c_int, ctypes.c_int, _dp, _dp, _dp, _dp,
                                       ctypes.c_int, ctypes.c_int, _up, ctypes.c_longlong, _up, ctypes.c_longlong,
                                       _dp, _dp, _dp, ctypes.c_int]
        _libs[native] = L
    return _libs[native]


def _p(a):
    return None if a is None else a.ctypes.data_as(_dp)


_blas = {}


def use_numpy_openblas(native: bool = False, enable: bool = True) -> str:
    """Route the port's matrix products (dimension >= 32) through the OpenBLAS that numpy bundles
    (numpy.libs/libscipy_openblas64_*.so: scipy_cblas_dgemm64_, ILP64), single-threaded per call --
    the port itself threads over units with OpenMP.  Returns a description, or "" if unavailable."""
    lb = lib(native)
    lb.qc_port_set_dgemm.argtypes = [ctypes.c_void_p]
    if not enable:
        lb.qc_port_set_dgemm(None)
        return ""
    if "lib" not in _blas:
        import glob
        cands = glob.glob(os.path.join(os.path.dirname(os.path.dirname(np.__file__)), "numpy.libs", "libscipy_openblas64_*.so"))
        _blas["lib"] = None
        for c in cands:
            try:
                B = ctypes.CDLL(c)
                fn = ctypes.cast(B.scipy_cblas_dgemm64_, ctypes.c_void_p).value
                B.scipy_openblas_set_num_threads64_.argtypes = [ctypes.c_int]
                B.scipy_openblas_set_num_threads64_(1)
                ver = ""
                try:
                    B.scipy_openblas_get_config64_.restype = ctypes.c_char_p
                    ver = B.scipy_openblas_get_config64_().decode()
                except Exception:
                    pass
                _blas.update(lib=B, fn=fn, desc=f"numpy's bundled OpenBLAS ({os.path.basename(c)}; {ver})".strip())
                break
            except Exception:
                continue
    if not _blas.get("lib"):
        return ""
    lb.qc_port_set_dgemm(_blas["fn"])
    return _blas["desc"]


def evaluate_arenas(layout, G0, Gd, Z, mu=None, want_jac=True, want_hess=True, k_range=None,
                    nthreads=0, native=False):
    """Returns (delta [S*nk*x_dim], jac_arena [S*nk, JA] or None, hess_arena [S*nk, HA] or None)."""
    L = layout
    k0, k1 = (0, L.K) if k_range is None else k_range
    nk = k1 - k0
    lb = lib(native)
    JA = lb.qc_port_jac_arena(L.d, L.c, L.m)
    HA = lb.qc_port_hess_arena(L.d, L.c, L.m)
    G0 = np.ascontiguousarray(np.transpose(G0, (0, 2, 1)), dtype=np.float64)       # col-major
    Gd = np.ascontiguousarray(np.transpose(Gd, (0, 1, 3, 2)), dtype=np.float64)
    Z = np.ascontiguousarray(Z, dtype=np.float64)
    mu = None if mu is None else np.ascontiguousarray(mu, dtype=np.float64)
    want = (1 if want_jac else 0) | (2 if (want_hess and mu is not None) else 0)
    delta = np.zeros(L.S * nk * L.x_dim)
    ja = np.zeros((L.S * nk, JA)) if want & 1 else None
    ha = np.zeros((L.S * nk, HA)) if want & 2 else None
    xo = np.asarray(L.x_off, dtype=np.int32)
    fails = lb.qc_port_eval(L.d, L.c, L.m, L.N, L.S, L.z_dim, xo.ctypes.data_as(_ip), L.u_off, L.dt_off,
                            _p(G0), _p(Gd), _p(Z), _p(mu), k0, k1, want, _p(delta), _p(ja), _p(ha), nthreads)
    if fails:
        raise FloatingPointError(f"{fails} units had ||dt*G(u)||_1 out of range or non-finite")
    return delta, ja, ha


def arena_codes(layout, s=0):
    """(jac_idx, jac_sign, hess_idx) mapping the native per-unit COO order to arena offsets (dp=d, cp=c)."""
    L = layout
    d, c, m = L.d, L.c, L.m
    dd, dc = d * d, d * c
    ji, js = [], []
    for (_, _, src) in L.jac_unit_pattern(s):
        t = src[0]
        if t == "E":
            ji.append(src[1] + src[2] * d); js.append(-1.0)
        elif t == "Z":
            ji.append(dd + (m + 1) * dc); js.append(1.0)
        elif t == "U":
            ji.append(dd + src[1] * dc + src[2]); js.append(1.0)
        elif t == "T":
            ji.append(dd + m * dc + src[1]); js.append(1.0)
        else:
            ji.append(dd + (m + 1) * dc + 1); js.append(1.0)
    hi = []
    for (_, _, src) in L.hess_unit_pattern(s):
        t = src[0]
        if t == "XU":
            hi.append(src[1] * dc + src[2])
        elif t == "XT":
            hi.append(m * dc + src[1])
        elif t == "UU":
            hi.append((m + 1) * dc + src[1] * m + src[2])
        elif t == "UT":
            hi.append((m + 1) * dc + m * m + src[1])
        else:
            hi.append((m + 1) * dc + m * m + m)
    return np.array(ji), np.array(js), np.array(hi)


def evaluate(layout, G0, Gd, Z, mu=None, want_jac=True, want_hess=True, k_range=None, nthreads=0, native=False):
    """Same contract as qc_oracle.evaluate: (delta, jac_vals, hess_vals) in native COO order."""
    L = layout
    delta, ja, ha = evaluate_arenas(L, G0, Gd, Z, mu, want_jac, want_hess, k_range, nthreads, native)
    nk = (L.K if k_range is None else k_range[1] - k_range[0])
    jac = hess = None
    if ja is not None:
        jac = np.empty((L.S, nk, L.jac_nnz_unit))
        for s in range(L.S):
            ji, js, _ = arena_codes(L, s)
            jac[s] = ja[s * nk:(s + 1) * nk][:, ji] * js
        jac = jac.reshape(-1)
    if ha is not None:
        hess = np.empty((L.S, nk, L.hess_nnz_unit))
        for s in range(L.S):
            _, _, hi = arena_codes(L, s)
            hess[s] = ha[s * nk:(s + 1) * nk][:, hi]
        hess = hess.reshape(-1)
    return delta, jac, hess


_code_cache = {}


def slot_codes(layout):
    """uint32 slot codes [S, nnz] (bit 31 = negate) of the native structure, cached per layout."""
    L = layout
    key = (L.d, L.c, L.m, L.S, L.z_dim, tuple(L.x_off), L.u_off, L.dt_off, L.jac_mode)
    if key not in _code_cache:
        jc, hc = [], []
        for s in range(L.S):
            ji, js, hi = arena_codes(L, s)
            jc.append(ji.astype(np.uint32) | np.where(js < 0, np.uint32(0x80000000), np.uint32(0)).astype(np.uint32))
            hc.append(hi.astype(np.uint32))
        _code_cache[key] = (np.ascontiguousarray(np.stack(jc)), np.ascontiguousarray(np.stack(hc)))
    return _code_cache[key]


def evaluate_coo(layout, G0, Gd, Z, mu, k_range=None, nthreads=0, native=False, out=None):
    """CPU evaluator delivering the COO value vectors (what the reference hands to Ipopt), assembled
    inside the C port.  Returns (delta, jac_vals, hess_vals)."""
    L = layout
    k0, k1 = (0, L.K) if k_range is None else k_range
    nk = k1 - k0
    lb = lib(native)
    jc, hc = slot_codes(L)
    G0 = np.ascontiguousarray(np.transpose(G0, (0, 2, 1)), dtype=np.float64)
    Gd = np.ascontiguousarray(np.transpose(Gd, (0, 1, 3, 2)), dtype=np.float64)
    Z = np.ascontiguousarray(Z, dtype=np.float64)
    mu = np.ascontiguousarray(mu, dtype=np.float64)
    if out is None:
        out = (np.empty(L.S * nk * L.x_dim), np.empty(L.S * nk * L.jac_nnz_unit), np.empty(L.S * nk * L.hess_nnz_unit))
    delta, jac, hess = out
    xo = np.asarray(L.x_off, dtype=np.int32)
    _up = ctypes.POINTER(ctypes.c_uint32)
    fails = lb.qc_port_eval_coo(L.d, L.c, L.m, L.N, L.S, L.z_dim, xo.ctypes.data_as(_ip), L.u_off, L.dt_off,
                                _p(G0), _p(Gd), _p(Z), _p(mu), k0, k1, jc.ctypes.data_as(_up), L.jac_nnz_unit,
                                hc.ctypes.data_as(_up), L.hess_nnz_unit, _p(delta), _p(jac), _p(hess), nthreads)
    if fails:
        raise FloatingPointError(f"{fails} units had ||dt*G(u)||_1 out of range or non-finite")
    return delta, jac, hess
