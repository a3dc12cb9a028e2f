"""CPU tests of the drop-in boundary: the C-ABI library loads and exports every symbol that
include/qcb200.h declares, and fails loudly without a GPU (no compute calls here)."""
import ctypes
import os
import re

import pytest

import qcb200
from qcb200 import _lib
from conftest import ROOT


def declared_functions():
    hdr = open(os.path.join(ROOT, "include", "qcb200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(qcb_[a-z_0-9]+)\s*\(", hdr)))


def test_header_declares_expected_surface():
    names = declared_functions()
    for must in ("qcb_create", "qcb_destroy", "qcb_set_structure", "qcb_eval_defect", "qcb_eval_jacobian",
                 "qcb_eval_hessian", "qcb_eval_all", "qcb_eval_all_device", "qcb_fidelity_eval", "qcb_last_error"):
        assert must in names


def test_library_exports_every_declared_symbol(built_libs):
    lib = ctypes.CDLL(built_libs)
    for name in declared_functions():
        assert hasattr(lib, name), f"{name} declared in include/qcb200.h but not exported"


def test_python_binding_covers_header(built_libs):
    assert sorted(_lib.SYMBOLS) == declared_functions()
    lib = _lib.load()
    assert b"sm_100a" in lib.qcb_version()


def test_descriptor_struct_matches_header():
    hdr = open(os.path.join(ROOT, "include", "qcb200.h")).read()
    body = re.search(r"typedef struct \{(.*?)\} qcb_desc;", hdr, flags=re.S).group(1)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    fields = re.findall(r"\b(\w+);", body)
    assert fields == [f[0] for f in _lib.QcbDesc._fields_]
    assert ctypes.sizeof(_lib.QcbDesc) == 14 * 4 + 5 * 8
    body = re.search(r"typedef struct \{([^{}]*)\} qcb_problem_desc;", hdr, flags=re.S).group(1)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    assert re.findall(r"\b(\w+);", body) == [f[0] for f in _lib.QcbProblemDesc._fields_]
    assert ctypes.sizeof(_lib.QcbProblemDesc) == 12 * 4 + 4 * 8 + 8 * 8


def test_problem_create_checks_its_pieces_without_a_gpu(built_libs):
    """qcb_problem_create validates the descriptor before it touches any piece."""
    lib = _lib.load()
    h = ctypes.c_void_p()
    d = _lib.QcbProblemDesc(struct_bytes=ctypes.sizeof(_lib.QcbProblemDesc), N=10, z_dim=12, use_derivative_rows=1)
    assert lib.qcb_problem_create(ctypes.byref(h), ctypes.byref(d)) == 1 and not h.value
    assert b"qcb_aux" in lib.qcb_problem_last_error(None)
    d = _lib.QcbProblemDesc(struct_bytes=4, N=10, z_dim=12)
    assert lib.qcb_problem_create(ctypes.byref(h), ctypes.byref(d)) == 1
    assert b"struct_bytes" in lib.qcb_problem_last_error(None)
    # an empty problem is legal: n = N * z_dim variables, no rows, no entries
    d = _lib.QcbProblemDesc(struct_bytes=ctypes.sizeof(_lib.QcbProblemDesc), N=10, z_dim=12)
    assert lib.qcb_problem_create(ctypes.byref(h), ctypes.byref(d)) == 0
    dims = [ctypes.c_int64(-1) for _ in range(4)]
    assert lib.qcb_problem_dims(h, *[ctypes.byref(x) for x in dims]) == 0
    assert [x.value for x in dims] == [120, 0, 0, 0]
    lib.qcb_problem_destroy(h)


def test_create_rejects_bad_descriptors_and_has_no_cpu_fallback(built_libs):
    import numpy as np
    from qcb200 import KnotLayout, Evaluator
    import torch
    L = KnotLayout.smooth_pulse(4, 2, 2, 10)
    G0, Gd = np.zeros((1, 4, 4)), np.zeros((1, 2, 4, 4))
    if not torch.cuda.is_available():
        with pytest.raises(_lib.QcbError) as ei:
            Evaluator(L, G0, Gd)
        assert ei.value.status == 2 and "no CPU fallback" in str(ei.value)
    bad = KnotLayout.smooth_pulse(4, 2, 2, 10)
    bad.u_off = bad.z_dim          # u outside the knot
    with pytest.raises(_lib.QcbError) as ei:
        Evaluator(bad, G0, Gd)
    assert ei.value.status == 1
    with pytest.raises(_lib.QcbError):
        Evaluator(L, G0, Gd, k_begin=5, k_end=3)
