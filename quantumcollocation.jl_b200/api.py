"""Thin object layer over the C ABI: one ``Evaluator`` per integrator set (handle)."""
from __future__ import annotations

import ctypes
from typing import Optional

import numpy as np

from . import _lib
from .layout import KnotLayout


def _ptr(a):
    """Raw address of a numpy array / torch tensor / int / None."""
    if a is None:
        return None
    if isinstance(a, int):
        return a
    if isinstance(a, np.ndarray):
        return a.ctypes.data
    return a.data_ptr()   # torch tensor


def _sync_producers(*buffers):
    """Synchronise torch's current stream on the device of any torch tensor among `buffers`."""
    for b in buffers:
        if b is not None and not isinstance(b, (int, np.ndarray)) and getattr(b, "is_cuda", False):
            import torch
            torch.cuda.current_stream(b.device).synchronize()
            return


def _colmajor(M):
    """[..., d, d] row-major numpy (M[r, c]) -> contiguous buffer holding column-major matrices."""
    return np.ascontiguousarray(np.swapaxes(np.asarray(M, dtype=np.float64), -1, -2))


class Evaluator:
    """Owns a ``qcb_handle``: defect / Jacobian / Hessian-of-Lagrangian values of a bilinear
    integrator set on one GPU, for the interval slab ``[k_begin, k_end)``."""

    def __init__(self, layout: KnotLayout, G_drift, G_drives, device: int = 0,
                 k_begin: int = 0, k_end: int = 0, force_path: int = 0):
        self.lib = _lib.load()
        self.layout = L = layout
        G_drift = np.asarray(G_drift, dtype=np.float64).reshape(L.S, L.d, L.d)
        G_drives = np.asarray(G_drives, dtype=np.float64).reshape(L.S, L.m, L.d, L.d)
        self._g0 = _colmajor(G_drift)
        self._gd = _colmajor(G_drives) if L.m else np.zeros(1)
        self._xo = np.asarray(L.x_off, dtype=np.int32)
        self._xi = None if L.x_index is None else np.ascontiguousarray(L.x_index, dtype=np.int32).reshape(L.S, L.x_dim)
        self._ri = None if L.row_index is None else np.ascontiguousarray(L.row_index, dtype=np.int32).reshape(L.x_dim)
        desc = _lib.QcbDesc(
            struct_bytes=ctypes.sizeof(_lib.QcbDesc), d=L.d, c=L.c, m=L.m, N=L.N, S=L.S, z_dim=L.z_dim,
            u_off=L.u_off, dt_off=L.dt_off, jac_mode=L.jac_mode, device=device, k_begin=k_begin, k_end=k_end,
            force_path=force_path,
            x_off=self._xo.ctypes.data_as(_lib.c_ip), G_drift=self._g0.ctypes.data_as(_lib.c_dp),
            G_drives=self._gd.ctypes.data_as(_lib.c_dp),
            x_index=None if self._xi is None else self._xi.ctypes.data_as(_lib.c_ip),
            row_index=None if self._ri is None else self._ri.ctypes.data_as(_lib.c_ip))
        h = ctypes.c_void_p()
        st = self.lib.qcb_create(ctypes.byref(h), ctypes.byref(desc))
        if st != 0:
            raise _lib.QcbError(st, self.lib.qcb_last_error(None).decode())
        self.h = h
        self.k_begin = k_begin
        self.k_end = k_end or L.K

    # -- sizes --
    @property
    def defect_dim(self) -> int:
        return self.lib.qcb_defect_dim(self.h)

    @property
    def jacobian_nnz(self) -> int:
        return self.lib.qcb_jacobian_nnz(self.h)

    @property
    def hessian_nnz(self) -> int:
        return self.lib.qcb_hessian_nnz(self.h)

    @property
    def kernel_launches(self) -> int:
        return self.lib.qcb_kernel_launches(self.h)

    @property
    def stream(self) -> int:
        return self.lib.qcb_stream(self.h) or 0

    def _check(self, st):
        if st != 0:
            raise _lib.QcbError(st, self.lib.qcb_last_error(self.h).decode())

    # -- structure --
    def jacobian_structure(self, row_base: int = 0):
        n = self.jacobian_nnz
        rows, cols = np.empty(n, dtype=np.int64), np.empty(n, dtype=np.int64)
        self._check(self.lib.qcb_jacobian_structure(self.h, row_base, rows.ctypes.data_as(_lib.c_lp), cols.ctypes.data_as(_lib.c_lp)))
        return rows, cols

    def hessian_structure(self):
        n = self.hessian_nnz
        rows, cols = np.empty(n, dtype=np.int64), np.empty(n, dtype=np.int64)
        self._check(self.lib.qcb_hessian_structure(self.h, rows.ctypes.data_as(_lib.c_lp), cols.ctypes.data_as(_lib.c_lp)))
        return rows, cols

    def set_structure(self, jac=None, hess=None, row_base: int = 0):
        def pr(t):
            if t is None:
                return None, None, 0
            r = np.ascontiguousarray(t[0], dtype=np.int64)
            c = np.ascontiguousarray(t[1], dtype=np.int64)
            return r, c, len(r)
        jr, jc, jn = pr(jac)
        hr, hc, hn = pr(hess)
        p = lambda a: None if a is None else a.ctypes.data_as(_lib.c_lp)
        self._check(self.lib.qcb_set_structure(self.h, row_base, p(jr), p(jc), jn, p(hr), p(hc), hn))

    # -- evaluation with host buffers --
    def eval_all(self, Z, mu=None, delta=None, jac=None, hess=None, want_jac=True, want_hess=True):
        """Returns (delta, jac_vals, hess_vals) numpy arrays (caller may pass preallocated ones)."""
        Z = np.ascontiguousarray(Z, dtype=np.float64)
        if mu is not None:
            mu = np.ascontiguousarray(mu, dtype=np.float64)
        if delta is None:
            delta = np.empty(self.defect_dim)
        if jac is None and want_jac:
            jac = np.empty(self.jacobian_nnz)
        if hess is None and want_hess and mu is not None:
            hess = np.empty(self.hessian_nnz)
        self._check(self.lib.qcb_eval_all(self.h, _ptr(Z), _ptr(mu), _ptr(delta), _ptr(jac), _ptr(hess)))
        return delta, jac, hess

    def eval_all_raw(self, Z, mu, delta, jac, hess):
        """Pointer-level call (numpy arrays, pinned torch tensors or raw addresses)."""
        self._check(self.lib.qcb_eval_all(self.h, _ptr(Z), _ptr(mu), _ptr(delta), _ptr(jac), _ptr(hess)))

    def eval_defect(self, Z, delta=None):
        Z = np.ascontiguousarray(Z, dtype=np.float64)
        delta = np.empty(self.defect_dim) if delta is None else delta
        self._check(self.lib.qcb_eval_defect(self.h, _ptr(Z), _ptr(delta)))
        return delta

    def eval_jacobian(self, Z, jac=None):
        Z = np.ascontiguousarray(Z, dtype=np.float64)
        jac = np.empty(self.jacobian_nnz) if jac is None else jac
        self._check(self.lib.qcb_eval_jacobian(self.h, _ptr(Z), _ptr(jac)))
        return jac

    def eval_hessian(self, Z, mu, hess=None):
        Z = np.ascontiguousarray(Z, dtype=np.float64)
        mu = np.ascontiguousarray(mu, dtype=np.float64)
        hess = np.empty(self.hessian_nnz) if hess is None else hess
        self._check(self.lib.qcb_eval_hessian(self.h, _ptr(Z), _ptr(mu), _ptr(hess)))
        return hess

    def eval_all_x(self, Z, mu, new_x, delta=None, jac=None, hess=None):
        """Host-buffer evaluation with Ipopt's new_x flag (include/qcb200.h, qcb_eval_all_x): outputs are
        written where given (numpy arrays or raw addresses); None skips that output."""
        Z = np.ascontiguousarray(Z, dtype=np.float64)
        mu = None if mu is None else np.ascontiguousarray(mu, dtype=np.float64)
        self._check(self.lib.qcb_eval_all_x(self.h, _ptr(Z), _ptr(mu), 1 if new_x else 0, _ptr(delta), _ptr(jac), _ptr(hess)))

    # -- evaluation with device buffers (torch CUDA tensors or raw device addresses) --
    def eval_all_device(self, dZ, dmu, d_delta, d_jac, d_hess):
        """Device-resident evaluation.  The library runs on its own stream, so torch tensors are made
        idle first (include/qcb200.h: every buffer must be idle when the call is made)."""
        _sync_producers(dZ, dmu, d_delta, d_jac, d_hess)
        self._check(self.lib.qcb_eval_all_device(self.h, _ptr(dZ), _ptr(dmu), _ptr(d_delta), _ptr(d_jac), _ptr(d_hess)))

    def eval_all_device_x(self, dZ, dmu, new_x, d_delta, d_jac, d_hess):
        """eval_all_device with Ipopt's new_x flag: new_x = False reuses what the previous call on the
        same Z left in the handle's scratch (include/qcb200.h, qcb_eval_all_device_x)."""
        _sync_producers(dZ, dmu, d_delta, d_jac, d_hess)
        self._check(self.lib.qcb_eval_all_device_x(self.h, _ptr(dZ), _ptr(dmu), 1 if new_x else 0, _ptr(d_delta), _ptr(d_jac), _ptr(d_hess)))

    def arena_doubles(self):
        """(Jacobian, Hessian) value-arena sizes per unit, in doubles."""
        return self.lib.qcb_arena_doubles(self.h, 0), self.lib.qcb_arena_doubles(self.h, 1)

    def eval_arenas_device(self, dZ, dmu, d_delta, d_jac_arenas, d_hess_arenas):
        """Like eval_all_device, but the per-unit value arenas are delivered instead of COO vectors."""
        _sync_producers(dZ, dmu, d_delta, d_jac_arenas, d_hess_arenas)
        self._check(self.lib.qcb_eval_arenas_device(self.h, _ptr(dZ), _ptr(dmu), _ptr(d_delta), _ptr(d_jac_arenas), _ptr(d_hess_arenas)))

    def eval_arenas_device_x(self, dZ, dmu, new_x, d_delta, d_jac_arenas, d_hess_arenas):
        _sync_producers(dZ, dmu, d_delta, d_jac_arenas, d_hess_arenas)
        self._check(self.lib.qcb_eval_arenas_device_x(self.h, _ptr(dZ), _ptr(dmu), 1 if new_x else 0, _ptr(d_delta),
                                                      _ptr(d_jac_arenas), _ptr(d_hess_arenas)))

    def set_async(self, on: bool):
        """Device-pointer calls return once queued; collect completion with wait()."""
        self._check(self.lib.qcb_set_async(self.h, 1 if on else 0))

    def wait(self, what: int = 0):
        """0: everything queued on this handle; 1 / 2 / 3: the last delivery of the E part / of the rest of
        the Jacobian arena / of the Hessian arena."""
        self._check(self.lib.qcb_wait(self.h, what))

    def assemble_device(self, d_jac_arenas, d_hess_arenas, n_units, nk_src, d_jac, d_hess, parts: int = 3):
        """COO value vectors of `n_units` units from their arenas (runs on this handle's GPU).
        parts: 1 = Jacobian entries that copy the E part of the arena, 2 = all others, 3 = all."""
        _sync_producers(d_jac_arenas, d_hess_arenas, d_jac, d_hess)
        self._check(self.lib.qcb_assemble_device(self.h, _ptr(d_jac_arenas), _ptr(d_hess_arenas), n_units, nk_src,
                                                 _ptr(d_jac), _ptr(d_hess), parts))

    def set_arena_parts(self, parts: int):
        """1: deliver the propagators (E part) only; 2: the rest of the Jacobian arena; 3: everything."""
        self._check(self.lib.qcb_set_arena_parts(self.h, parts))

    def rollout(self, Z, x_init=None):
        """States [S, N, x_dim] of x_{k+1} = exp(dt_k G(u_k)) x_k (x_init [S, x_dim] or knot 1 of Z)."""
        L = self.layout
        Z = np.ascontiguousarray(Z, dtype=np.float64)
        if x_init is not None:
            x_init = np.ascontiguousarray(x_init, dtype=np.float64).reshape(L.S, L.x_dim)
        states = np.empty((L.S, L.N, L.x_dim))
        self._check(self.lib.qcb_rollout(self.h, _ptr(Z), _ptr(x_init), _ptr(states)))
        return states

    def timings(self):
        ms = (ctypes.c_double * 4)()
        self.lib.qcb_get_timings(self.h, ms)
        return {"h2d_ms": ms[0], "kernels_ms": ms[1], "d2h_ms": ms[2], "total_ms": ms[3]}

    def transfer_bytes(self):
        b = (ctypes.c_int64 * 2)()
        self.lib.qcb_get_transfer_bytes(self.h, b)
        return int(b[0]), int(b[1])

    def profile(self, on: bool):
        self.lib.qcb_profile_enable(self.h, 1 if on else 0)

    def profile_read(self):
        out = (ctypes.c_double * 4)()
        self.lib.qcb_profile_read(self.h, out)
        return {"launches": out[0], "ms": out[1], "flops": out[2]}

    def close(self):
        if getattr(self, "h", None):
            self.lib.qcb_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Fidelity:
    """Terminal fidelity F(x_N) and its derivatives on the GPU (include/qcb200.h, qcb_fidelity_*)."""
    UNITARY, KET, COHERENT_KET, UNITARY_SUBSPACE, DENSITY_PURE = 0, 1, 2, 3, 4

    def __init__(self, kind: int, goal, n: Optional[int] = None, subspace=None, device: int = 0):
        self.lib = _lib.load()
        goal = np.asarray(goal, dtype=complex)
        if kind == self.UNITARY:
            n = goal.shape[0]; count = 1
            flat = goal.T.reshape(-1)
        elif kind in (self.KET, self.DENSITY_PURE):
            n = goal.shape[0]; count = 1
            flat = goal.reshape(-1)
        elif kind == self.COHERENT_KET:
            count, n = goal.shape
            flat = goal.reshape(-1)
        else:
            count = 1
            flat = goal.T.reshape(-1)
        gi = np.ascontiguousarray(np.stack([flat.real, flat.imag], axis=-1).reshape(-1))
        sub = None if subspace is None else np.asarray(subspace, dtype=np.int32)
        h = ctypes.c_void_p()
        st = self.lib.qcb_fidelity_create(ctypes.byref(h), kind, n, count, gi.ctypes.data_as(_lib.c_dp),
                                          None if sub is None else sub.ctypes.data_as(_lib.c_ip),
                                          0 if sub is None else len(sub), device)
        if st != 0:
            raise _lib.QcbError(st, self.lib.qcb_last_error(None).decode())
        self.h = h
        self.x_dim = self.lib.qcb_fidelity_x_dim(h)

    def eval(self, x, want_grad=True, want_hess=False):
        x = np.ascontiguousarray(x, dtype=np.float64)
        assert x.shape == (self.x_dim,)
        F = np.zeros(1)
        g = np.empty(self.x_dim) if want_grad else None
        H = np.empty((self.x_dim, self.x_dim)) if want_hess else None
        st = self.lib.qcb_fidelity_eval(self.h, _ptr(x), _ptr(F), _ptr(g), _ptr(H))
        if st != 0:
            raise _lib.QcbError(st, "fidelity evaluation failed")
        return float(F[0]), g, H

    def hessian_factor(self):
        """(G [rank, x_dim], w): Hess F = 2 w G'G, constant in x (rank 0 for the linear density kind)."""
        R = self.lib.qcb_fidelity_rank(self.h)
        G = np.empty((max(R, 1), self.x_dim))
        w = ctypes.c_double(0.0)
        if self.lib.qcb_fidelity_factor(self.h, _ptr(G), ctypes.byref(w)) != 0:
            raise _lib.QcbError(2, "qcb_fidelity_factor failed")
        return G[:R], w.value

    def close(self):
        if getattr(self, "h", None):
            self.lib.qcb_fidelity_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class FreePhaseFidelity:
    """F(x, theta) = |tr(U_goal(theta)' U)|^2 / n^2 with U_goal(theta) = diag(e^{i P theta}) U_goal
    (side 0) or U_goal diag(e^{i P theta}) (side 1) -- include/qcb200.h, qcb_free_phase_*."""

    def __init__(self, U_goal, phase_map, side: int = 0, device: int = 0):
        self.lib = _lib.load()
        U_goal = np.asarray(U_goal, dtype=complex)
        self.n = U_goal.shape[0]
        P = np.asarray(phase_map, dtype=np.float64)
        P = P.reshape(self.n, P.size // self.n)
        self.n_theta = P.shape[1]
        flat = U_goal.T.reshape(-1)
        gi = np.ascontiguousarray(np.stack([flat.real, flat.imag], axis=-1).reshape(-1))
        Pc = np.ascontiguousarray(P.T)                     # column-major n x n_theta
        h = ctypes.c_void_p()
        st = self.lib.qcb_free_phase_create(ctypes.byref(h), self.n, gi.ctypes.data_as(_lib.c_dp), self.n_theta,
                                            Pc.ctypes.data_as(_lib.c_dp) if self.n_theta else None, side, device)
        if st != 0:
            raise _lib.QcbError(st, self.lib.qcb_last_error(None).decode())
        self.h = h
        self.x_dim = 2 * self.n * self.n
        self.z_dim = self.x_dim + self.n_theta

    def eval(self, x, theta, want_grad=True, want_hess=False):
        """(F, grad over [x; theta], dense Hessian over [x; theta])."""
        x = np.ascontiguousarray(x, dtype=np.float64)
        theta = np.ascontiguousarray(theta, dtype=np.float64).reshape(self.n_theta)
        assert x.shape == (self.x_dim,)
        F = np.zeros(1)
        g = np.empty(self.z_dim) if want_grad else None
        H = np.empty((self.z_dim, self.z_dim)) if want_hess else None
        st = self.lib.qcb_free_phase_eval(self.h, _ptr(x), _ptr(theta) if self.n_theta else None, _ptr(F), _ptr(g), _ptr(H))
        if st != 0:
            raise _lib.QcbError(st, "free-phase fidelity evaluation failed")
        return float(F[0]), g, H

    def close(self):
        if getattr(self, "h", None):
            self.lib.qcb_free_phase_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Aux:
    """Derivative integrators and leakage terms on the GPU (include/qcb200.h, qcb_aux_*)."""

    def __init__(self, N: int, z_dim: int, dt_off: int, derivatives=(), leak_x_off: int = 0, leak_idx=(), device: int = 0):
        self.lib = _lib.load()
        self.N, self.z_dim = N, z_dim
        self._dx = np.asarray([d[0] for d in derivatives], dtype=np.int32)
        self._ddx = np.asarray([d[1] for d in derivatives], dtype=np.int32)
        self._dim = np.asarray([d[2] for d in derivatives], dtype=np.int32)
        self._leak = np.asarray(leak_idx, dtype=np.int32)
        p = lambda arr: arr.ctypes.data_as(_lib.c_ip) if len(arr) else None
        desc = _lib.QcbAuxDesc(struct_bytes=ctypes.sizeof(_lib.QcbAuxDesc), N=N, z_dim=z_dim, dt_off=dt_off, device=device,
                               n_deriv=len(self._dim), leak_x_off=leak_x_off, n_leak=len(self._leak),
                               deriv_x_off=p(self._dx), deriv_dx_off=p(self._ddx), deriv_dim=p(self._dim), leak_idx=p(self._leak))
        h = ctypes.c_void_p()
        st = self.lib.qcb_aux_create(ctypes.byref(h), ctypes.byref(desc))
        if st != 0:
            raise _lib.QcbError(st, self.lib.qcb_last_error(None).decode())
        self.h = h
        self.derivative_dim = self.lib.qcb_aux_derivative_dim(h)
        self.leakage_dim = self.lib.qcb_aux_leakage_dim(h)

    def derivative_structure(self, row_base=0):
        n = self.derivative_dim
        jr, jc = np.empty(4 * n, dtype=np.int64), np.empty(4 * n, dtype=np.int64)
        hr, hc = np.empty(n, dtype=np.int64), np.empty(n, dtype=np.int64)
        P = lambda x: x.ctypes.data_as(_lib.c_lp)
        st = self.lib.qcb_aux_derivative_structure(self.h, row_base, P(jr), P(jc), P(hr), P(hc))
        assert st == 0
        return (jr, jc), (hr, hc)

    def derivative_eval(self, Z, mu=None):
        Z = np.ascontiguousarray(Z, dtype=np.float64)
        n = self.derivative_dim
        delta, jac = np.empty(n), np.empty(4 * n)
        hess = np.empty(n) if mu is not None else None
        mu = None if mu is None else np.ascontiguousarray(mu, dtype=np.float64)
        st = self.lib.qcb_aux_derivative_eval(self.h, _ptr(Z), _ptr(mu), _ptr(delta), _ptr(jac), _ptr(hess))
        if st != 0:
            raise _lib.QcbError(st, "derivative integrator evaluation failed")
        return delta, jac, hess

    def leakage_eval(self, Z, Qs=None, value=0.0, mu=None):
        Z = np.ascontiguousarray(Z, dtype=np.float64)
        n = self.leakage_dim
        obj = ctypes.c_double(0.0)
        grad, con, cjac = np.empty(n), np.empty(n), np.empty(n)
        chess = np.empty(n) if mu is not None else None
        Qs = None if Qs is None else np.ascontiguousarray(Qs, dtype=np.float64)
        mu = None if mu is None else np.ascontiguousarray(mu, dtype=np.float64)
        st = self.lib.qcb_aux_leakage_eval(self.h, _ptr(Z), _ptr(Qs), float(value), _ptr(mu), ctypes.byref(obj),
                                           _ptr(grad), _ptr(con), _ptr(cjac), _ptr(chess))
        if st != 0:
            raise _lib.QcbError(st, "leakage evaluation failed")
        return obj.value, grad, con, cjac, chess

    @staticmethod
    def _terms(terms):
        off = np.asarray([t[0] for t in terms], dtype=np.int32)
        dim = np.asarray([t[1] for t in terms], dtype=np.int32)
        R = np.concatenate([np.broadcast_to(np.asarray(t[2], dtype=np.float64), (t[1],)) for t in terms]) if terms else np.zeros(0)
        return off, dim, np.ascontiguousarray(R)

    def regularizer_structure(self, terms):
        """(grad_cols, (hess_rows, hess_cols)): 1-based columns of Z of the compact gradient of
        `regularizer_eval` and the upper-triangle COO of its Hessian values (same orders)."""
        off, dim, _ = self._terms(terms)
        rows = int(dim.sum())
        gc = np.empty(self.N * (rows + 1), dtype=np.int64)
        hr, hc = np.empty(self.N * (2 * rows + 1), dtype=np.int64), np.empty(self.N * (2 * rows + 1), dtype=np.int64)
        ip = lambda a: a.ctypes.data_as(_lib.c_ip) if len(a) else None
        lp = lambda a: a.ctypes.data_as(_lib.c_lp)
        st = self.lib.qcb_aux_regularizer_structure(self.h, len(dim), ip(off), ip(dim), lp(gc), lp(hr), lp(hc))
        if st != 0:
            raise _lib.QcbError(st, "regulariser structure failed")
        return gc, (hr, hc)

    def regularizer_eval(self, Z, terms, min_time_D=0.0, want_hessian=True):
        """terms: [(offset in knot, dim, R scalar or [dim])].  J = sum 1/2 R (dt v)^2 + D sum_{t<N} dt.
        Returns (J, compact gradient values [N*(rows+1)], Hessian values [N*(2 rows+1)])."""
        Z = np.ascontiguousarray(Z, dtype=np.float64)
        off, dim, R = self._terms(terms)
        rows = int(dim.sum())
        obj = ctypes.c_double(0.0)
        grad = np.empty(self.N * (rows + 1))
        hess = np.empty(self.N * (2 * rows + 1)) if want_hessian else None
        ip = lambda a: a.ctypes.data_as(_lib.c_ip) if len(a) else None
        st = self.lib.qcb_aux_regularizer_eval(self.h, _ptr(Z), len(dim), ip(off), ip(dim), _ptr(R) if len(R) else None,
                                               float(min_time_D), ctypes.byref(obj), _ptr(grad), _ptr(hess))
        if st != 0:
            raise _lib.QcbError(st, "regulariser evaluation failed")
        return obj.value, grad, hess

    def regularizer_gradient(self, Z, terms, min_time_D=0.0):
        """Dense gradient over Z (length N*z_dim), scattered from the compact values."""
        _, g, _ = self.regularizer_eval(Z, terms, min_time_D, want_hessian=False)
        gc, _ = self.regularizer_structure(terms)
        dense = np.zeros(self.N * self.z_dim)
        np.add.at(dense, gc - 1, g)
        return dense

    def sensitivity_eval(self, Z, x_off, x_dim, times, Qs=None, scale=1.0, want_hessian=True):
        """UnitarySensitivityObjective: (J, grad [T, x_dim], hess [T, x_dim, x_dim]); times 1-based."""
        Z = np.ascontiguousarray(Z, dtype=np.float64)
        times = np.ascontiguousarray(times, dtype=np.int32)
        T = len(times)
        n = int(round(np.sqrt(x_dim / 2)))
        Qs = None if Qs is None else np.ascontiguousarray(Qs, dtype=np.float64)
        obj = ctypes.c_double(0.0)
        grad = np.empty((T, x_dim))
        hess = np.empty((T, x_dim, x_dim)) if want_hessian else None
        st = self.lib.qcb_aux_sensitivity_eval(self.h, _ptr(Z), x_off, x_dim, n, T, times.ctypes.data_as(_lib.c_ip) if T else None,
                                               _ptr(Qs), float(scale), ctypes.byref(obj), _ptr(grad), _ptr(hess))
        if st != 0:
            raise _lib.QcbError(st, "sensitivity objective evaluation failed")
        return obj.value, grad, hess

    def timesteps_equal(self, Z, row_base=0):
        """Rows dt_t - dt_N (t < N): (delta, (rows, cols), values)."""
        Z = np.ascontiguousarray(Z, dtype=np.float64)
        K = self.N - 1
        delta, vals = np.empty(K), np.empty(2 * K)
        jr, jc = np.empty(2 * K, dtype=np.int64), np.empty(2 * K, dtype=np.int64)
        st = self.lib.qcb_aux_timesteps_equal(self.h, _ptr(Z), row_base, _ptr(delta), jr.ctypes.data_as(_lib.c_lp),
                                              jc.ctypes.data_as(_lib.c_lp), _ptr(vals))
        if st != 0:
            raise _lib.QcbError(st, "timestep constraint evaluation failed")
        return delta, (jr, jc), vals

    def close(self):
        if getattr(self, "h", None):
            self.lib.qcb_aux_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Problem:
    """One call per Ipopt callback (include/qcb200.h, qcb_problem_*): the integrators, the derivative and
    timestep rows of an `Aux`, a terminal `Fidelity` with weight Q and the quadratic regularisers, laid out in
    Ipopt's full-length vectors.  The pieces are borrowed: keep them alive, close the problem first.

    integrators: `Evaluator`s owning the whole trajectory; fid_index: 0-based positions of the fidelity's
    state entries inside the last knot ([states, x_dim] for the systems of a sampling problem, fid_weights per state); reg_terms: [(offset in knot, dim, R scalar or [dim])]; leakage_rows /
    leakage_Q: the inequality rows x^2 - leakage_value <= 0 and the objective term of `aux`'s leakage indices."""

    def __init__(self, N: int, z_dim: int, integrators=(), aux: Optional["Aux"] = None, derivative_rows: bool = False,
                 dt_equal_rows: bool = False, fidelity: Optional["Fidelity"] = None, fid_index=None, Q: float = 100.0,
                 fid_hessian: bool = True, reg_terms=(), min_time_D: float = 0.0, leakage_rows: bool = False,
                 leakage_value: float = 0.0, leakage_Q: Optional[float] = None, fid_weights=None):
        self.lib = _lib.load()
        self._keep = (list(integrators), aux, fidelity)
        hs = (ctypes.c_void_p * max(1, len(integrators)))(*[ev.h for ev in integrators])
        self._fid_index = None if fid_index is None else np.ascontiguousarray(fid_index, dtype=np.int32).reshape(-1)
        n_states = 0
        if fidelity is not None and self._fid_index is not None:
            n_states = self._fid_index.size // int(self.lib.qcb_fidelity_x_dim(fidelity.h))
        self._fid_w = None if fid_weights is None else np.ascontiguousarray(fid_weights, dtype=np.float64)
        assert self._fid_w is None or self._fid_w.size == n_states
        self._off, self._dim, self._R = Aux._terms(list(reg_terms))
        ip = lambda a: a.ctypes.data_as(_lib.c_ip) if a is not None and len(a) else None
        desc = _lib.QcbProblemDesc(
            struct_bytes=ctypes.sizeof(_lib.QcbProblemDesc), N=N, z_dim=z_dim, n_integrators=len(integrators),
            use_derivative_rows=int(bool(derivative_rows)), use_dt_equal_rows=int(bool(dt_equal_rows)),
            fid_hessian=int(bool(fid_hessian)), n_reg=len(self._dim), Q=float(Q), min_time_D=float(min_time_D),
            use_leakage_rows=int(bool(leakage_rows)), use_leakage_objective=int(leakage_Q is not None),
            leak_Q=float(leakage_Q or 0.0), leak_value=float(leakage_value), n_fid_states=n_states,
            fid_weights=self._fid_w.ctypes.data_as(_lib.c_dp) if self._fid_w is not None else None,
            integrators=ctypes.cast(hs, ctypes.POINTER(ctypes.c_void_p)), aux=aux.h if aux is not None else None,
            fidelity=fidelity.h if fidelity is not None else None, fid_index=ip(self._fid_index),
            reg_off=ip(self._off), reg_dim=ip(self._dim),
            reg_R=self._R.ctypes.data_as(_lib.c_dp) if len(self._R) else None)
        h = ctypes.c_void_p()
        st = self.lib.qcb_problem_create(ctypes.byref(h), ctypes.byref(desc))
        if st != 0:
            raise _lib.QcbError(st, self.lib.qcb_problem_last_error(None).decode())
        self.h = h
        d = [ctypes.c_int64(0) for _ in range(4)]
        self.lib.qcb_problem_dims(h, *[ctypes.byref(x) for x in d])
        self.n, self.m, self.jacobian_nnz, self.hessian_nnz = (int(x.value) for x in d)

    def _check(self, st):
        if st != 0:
            raise _lib.QcbError(st, self.lib.qcb_problem_last_error(self.h).decode())

    def jacobian_structure(self):
        rows, cols = np.empty(self.jacobian_nnz, dtype=np.int64), np.empty(self.jacobian_nnz, dtype=np.int64)
        self._check(self.lib.qcb_problem_jacobian_structure(self.h, rows.ctypes.data_as(_lib.c_lp), cols.ctypes.data_as(_lib.c_lp)))
        return rows, cols

    def hessian_structure(self):
        rows, cols = np.empty(self.hessian_nnz, dtype=np.int64), np.empty(self.hessian_nnz, dtype=np.int64)
        self._check(self.lib.qcb_problem_hessian_structure(self.h, rows.ctypes.data_as(_lib.c_lp), cols.ctypes.data_as(_lib.c_lp)))
        return rows, cols

    def eval_f(self, Z, new_x: bool = True) -> float:
        Z = np.ascontiguousarray(Z, dtype=np.float64)
        f = ctypes.c_double(0.0)
        self._check(self.lib.qcb_problem_eval_f(self.h, _ptr(Z), 1 if new_x else 0, ctypes.byref(f)))
        return f.value

    def eval_grad_f(self, Z, new_x: bool = True, out=None):
        Z = np.ascontiguousarray(Z, dtype=np.float64)
        out = np.empty(self.n) if out is None else out
        self._check(self.lib.qcb_problem_eval_grad_f(self.h, _ptr(Z), 1 if new_x else 0, _ptr(out)))
        return out

    def eval_g(self, Z, new_x: bool = True, out=None):
        Z = np.ascontiguousarray(Z, dtype=np.float64)
        out = np.empty(self.m) if out is None else out
        self._check(self.lib.qcb_problem_eval_g(self.h, _ptr(Z), 1 if new_x else 0, _ptr(out)))
        return out

    def eval_jac_g(self, Z, new_x: bool = True, out=None):
        Z = np.ascontiguousarray(Z, dtype=np.float64)
        out = np.empty(self.jacobian_nnz) if out is None else out
        self._check(self.lib.qcb_problem_eval_jac_g(self.h, _ptr(Z), 1 if new_x else 0, _ptr(out)))
        return out

    def eval_h(self, Z, obj_factor: float, lam, new_x: bool = True, out=None):
        Z = np.ascontiguousarray(Z, dtype=np.float64)
        lam = np.ascontiguousarray(lam, dtype=np.float64)
        assert lam.size == self.m
        out = np.empty(self.hessian_nnz) if out is None else out
        self._check(self.lib.qcb_problem_eval_h(self.h, _ptr(Z), 1 if new_x else 0, float(obj_factor), _ptr(lam), _ptr(out)))
        return out

    def close(self):
        if getattr(self, "h", None):
            self.lib.qcb_problem_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
