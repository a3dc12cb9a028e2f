"""ctypes loader of the product library ``csrc/libqcb200.so`` (the C ABI of include/qcb200.h).

There is no fallback: if the library is missing or no CUDA device is usable, every
evaluation entry point raises.
"""
from __future__ import annotations

import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("QCB200_LIB") or os.path.join(_HERE, "csrc", "libqcb200.so")   # QCB200_LIB: another build of the same ABI

c_dp = ctypes.POINTER(ctypes.c_double)
c_ip = ctypes.POINTER(ctypes.c_int32)
c_lp = ctypes.POINTER(ctypes.c_int64)
c_vp = ctypes.c_void_p


class QcbDesc(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int32) for n in (
        "struct_bytes", "d", "c", "m", "N", "S", "z_dim", "u_off", "dt_off", "jac_mode", "device",
        "k_begin", "k_end", "force_path")] + [
        ("x_off", c_ip), ("G_drift", c_dp), ("G_drives", c_dp), ("x_index", c_ip), ("row_index", c_ip)]


class QcbAuxDesc(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int32) for n in (
        "struct_bytes", "N", "z_dim", "dt_off", "device", "n_deriv", "leak_x_off", "n_leak")] + [
        ("deriv_x_off", c_ip), ("deriv_dx_off", c_ip), ("deriv_dim", c_ip), ("leak_idx", c_ip)]


class QcbProblemDesc(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int32) for n in (
        "struct_bytes", "N", "z_dim", "n_integrators", "use_derivative_rows", "use_dt_equal_rows", "fid_hessian", "n_reg",
        "use_leakage_rows", "use_leakage_objective", "n_fid_states", "reserved0")] + [
        ("Q", ctypes.c_double), ("min_time_D", ctypes.c_double), ("leak_Q", ctypes.c_double), ("leak_value", ctypes.c_double),
        ("integrators", ctypes.POINTER(c_vp)), ("aux", c_vp),
        ("fidelity", c_vp), ("fid_index", c_ip), ("fid_weights", c_dp), ("reg_off", c_ip), ("reg_dim", c_ip), ("reg_R", c_dp)]


STATUS = {0: "QCB_OK", 1: "QCB_ERR_ARG", 2: "QCB_ERR_CUDA", 3: "QCB_ERR_RANGE", 4: "QCB_ERR_STRUCTURE", 5: "QCB_ERR_NOMEM"}

# every symbol include/qcb200.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "qcb_create": (ctypes.c_int, [ctypes.POINTER(c_vp), ctypes.POINTER(QcbDesc)]),
    "qcb_destroy": (None, [c_vp]),
    "qcb_last_error": (ctypes.c_char_p, [c_vp]),
    "qcb_version": (ctypes.c_char_p, []),
    "qcb_defect_dim": (ctypes.c_int64, [c_vp]),
    "qcb_jacobian_nnz": (ctypes.c_int64, [c_vp]),
    "qcb_hessian_nnz": (ctypes.c_int64, [c_vp]),
    "qcb_jacobian_structure": (ctypes.c_int, [c_vp, ctypes.c_int64, c_lp, c_lp]),
    "qcb_hessian_structure": (ctypes.c_int, [c_vp, c_lp, c_lp]),
    "qcb_set_structure": (ctypes.c_int, [c_vp, ctypes.c_int64, c_lp, c_lp, ctypes.c_int64, c_lp, c_lp, ctypes.c_int64]),
    "qcb_eval_defect": (ctypes.c_int, [c_vp, c_vp, c_vp]),
    "qcb_eval_jacobian": (ctypes.c_int, [c_vp, c_vp, c_vp]),
    "qcb_eval_hessian": (ctypes.c_int, [c_vp, c_vp, c_vp, c_vp]),
    "qcb_eval_all": (ctypes.c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "qcb_eval_defect_x": (ctypes.c_int, [c_vp, c_vp, ctypes.c_int32, c_vp]),
    "qcb_eval_jacobian_x": (ctypes.c_int, [c_vp, c_vp, ctypes.c_int32, c_vp]),
    "qcb_eval_hessian_x": (ctypes.c_int, [c_vp, c_vp, ctypes.c_int32, c_vp, c_vp]),
    "qcb_eval_all_x": (ctypes.c_int, [c_vp, c_vp, c_vp, ctypes.c_int32, c_vp, c_vp, c_vp]),
    "qcb_eval_all_device": (ctypes.c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "qcb_eval_all_device_x": (ctypes.c_int, [c_vp, c_vp, c_vp, ctypes.c_int32, c_vp, c_vp, c_vp]),
    "qcb_rollout": (ctypes.c_int, [c_vp, c_vp, c_vp, c_vp]),
    "qcb_arena_doubles": (ctypes.c_int64, [c_vp, ctypes.c_int32]),
    "qcb_eval_arenas_device": (ctypes.c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "qcb_assemble_device": (ctypes.c_int, [c_vp, c_vp, c_vp, ctypes.c_int64, ctypes.c_int32, c_vp, c_vp, ctypes.c_int32]),
    "qcb_set_arena_parts": (ctypes.c_int, [c_vp, ctypes.c_int32]),
    "qcb_eval_arenas_device_x": (ctypes.c_int, [c_vp, c_vp, c_vp, ctypes.c_int32, c_vp, c_vp, c_vp]),
    "qcb_set_async": (ctypes.c_int, [c_vp, ctypes.c_int32]),
    "qcb_wait": (ctypes.c_int, [c_vp, ctypes.c_int32]),
    "qcb_device_alloc": (ctypes.c_int, [ctypes.c_int32, ctypes.POINTER(c_vp), ctypes.c_size_t]),
    "qcb_device_free": (None, [ctypes.c_int32, c_vp]),
    "qcb_ipc_export": (ctypes.c_int, [ctypes.c_int32, c_vp, ctypes.c_char_p]),
    "qcb_ipc_open": (ctypes.c_int, [ctypes.c_int32, ctypes.c_char_p, ctypes.POINTER(c_vp)]),
    "qcb_ipc_close": (None, [ctypes.c_int32, c_vp]),
    "qcb_device_to_host": (ctypes.c_int, [ctypes.c_int32, c_vp, c_vp, ctypes.c_size_t]),
    "qcb_host_alloc": (ctypes.c_int, [ctypes.POINTER(c_vp), ctypes.c_size_t]),
    "qcb_host_free": (None, [c_vp]),
    "qcb_host_register": (ctypes.c_int, [c_vp, ctypes.c_size_t]),
    "qcb_host_unregister": (None, [c_vp]),
    "qcb_host_threads": (ctypes.c_int32, []),
    "qcb_host_stream_fill": (ctypes.c_int, [c_vp, ctypes.c_size_t, c_dp]),
    "qcb_kernel_launches": (ctypes.c_int64, [c_vp]),
    "qcb_get_timings": (ctypes.c_int, [c_vp, c_dp]),
    "qcb_get_transfer_bytes": (ctypes.c_int, [c_vp, c_lp]),
    "qcb_stream": (c_vp, [c_vp]),
    "qcb_profile_enable": (ctypes.c_int, [c_vp, ctypes.c_int32]),
    "qcb_profile_read": (ctypes.c_int, [c_vp, c_dp]),
    "qcb_fidelity_create": (ctypes.c_int, [ctypes.POINTER(c_vp), ctypes.c_int32, ctypes.c_int32, ctypes.c_int32,
                                           c_dp, c_ip, ctypes.c_int32, ctypes.c_int32]),
    "qcb_fidelity_destroy": (None, [c_vp]),
    "qcb_fidelity_x_dim": (ctypes.c_int64, [c_vp]),
    "qcb_fidelity_eval": (ctypes.c_int, [c_vp, c_vp, c_vp, c_vp, c_vp]),
    "qcb_fidelity_rank": (ctypes.c_int64, [c_vp]),
    "qcb_fidelity_factor": (ctypes.c_int, [c_vp, c_vp, c_dp]),
    "qcb_aux_create": (ctypes.c_int, [ctypes.POINTER(c_vp), ctypes.POINTER(QcbAuxDesc)]),
    "qcb_aux_destroy": (None, [c_vp]),
    "qcb_aux_derivative_dim": (ctypes.c_int64, [c_vp]),
    "qcb_aux_leakage_dim": (ctypes.c_int64, [c_vp]),
    "qcb_aux_derivative_structure": (ctypes.c_int, [c_vp, ctypes.c_int64, c_lp, c_lp, c_lp, c_lp]),
    "qcb_aux_derivative_eval": (ctypes.c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "qcb_aux_leakage_eval": (ctypes.c_int, [c_vp, c_vp, c_vp, ctypes.c_double, c_vp, c_dp, c_vp, c_vp, c_vp, c_vp]),
    "qcb_aux_regularizer_structure": (ctypes.c_int, [c_vp, ctypes.c_int32, c_ip, c_ip, c_lp, c_lp, c_lp]),
    "qcb_aux_regularizer_eval": (ctypes.c_int, [c_vp, c_vp, ctypes.c_int32, c_ip, c_ip, c_vp, ctypes.c_double, c_dp, c_vp, c_vp]),
    "qcb_aux_timesteps_equal": (ctypes.c_int, [c_vp, c_vp, ctypes.c_int64, c_vp, c_lp, c_lp, c_vp]),
    "qcb_free_phase_create": (ctypes.c_int, [ctypes.POINTER(c_vp), ctypes.c_int32, c_dp, ctypes.c_int32, c_dp,
                                             ctypes.c_int32, ctypes.c_int32]),
    "qcb_free_phase_destroy": (None, [c_vp]),
    "qcb_free_phase_eval": (ctypes.c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "qcb_aux_sensitivity_eval": (ctypes.c_int, [c_vp, c_vp, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32,
                                                c_ip, c_vp, ctypes.c_double, c_dp, c_vp, c_vp]),
    "qcb_fidelity_eval_batch_device_noalloc": (ctypes.c_int, [c_vp, c_vp, ctypes.c_int64, c_vp, c_vp]),
    "qcb_host_threads_in_use": (ctypes.c_int32, [c_vp, ctypes.c_int32]),
    "qcb_problem_create": (ctypes.c_int, [ctypes.POINTER(c_vp), ctypes.POINTER(QcbProblemDesc)]),
    "qcb_problem_destroy": (None, [c_vp]),
    "qcb_problem_last_error": (ctypes.c_char_p, [c_vp]),
    "qcb_problem_dims": (ctypes.c_int, [c_vp, c_lp, c_lp, c_lp, c_lp]),
    "qcb_problem_jacobian_structure": (ctypes.c_int, [c_vp, c_lp, c_lp]),
    "qcb_problem_hessian_structure": (ctypes.c_int, [c_vp, c_lp, c_lp]),
    "qcb_problem_eval_f": (ctypes.c_int, [c_vp, c_vp, ctypes.c_int32, c_dp]),
    "qcb_problem_eval_grad_f": (ctypes.c_int, [c_vp, c_vp, ctypes.c_int32, c_vp]),
    "qcb_problem_eval_g": (ctypes.c_int, [c_vp, c_vp, ctypes.c_int32, c_vp]),
    "qcb_problem_eval_jac_g": (ctypes.c_int, [c_vp, c_vp, ctypes.c_int32, c_vp]),
    "qcb_problem_eval_h": (ctypes.c_int, [c_vp, c_vp, ctypes.c_int32, ctypes.c_double, c_vp, c_vp]),
}

_lib = None


class QcbError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"{STATUS.get(status, status)}: {message}")
        self.status = status


def load():
    """Load libqcb200.so and bind every declared symbol.  Raises if the library is absent."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `make -C {os.path.dirname(LIB_PATH)}` "
                "(or __graft_entry__.build()).  There is no CPU fallback.")
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(lib, name)   # AttributeError if the symbol is not exported
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib
