"""The allocation-free batched fidelity entry point, checked against the ordinary one.

Some users of this library cannot let it allocate GPU memory when they call it: their process has
handed all of the device to another memory pool (PyTorch's or CUDA.jl's caching allocator), or they
sit in a loop where the device-wide stall of a cudaMalloc is not acceptable.  For them
qcb_fidelity_eval_batch_device_noalloc keeps its scratch in a fixed array inside the library.  This
test calls it through the C ABI twice in a row (the array is shared and keeps old contents, so the
second call must not be disturbed by the first) and compares with qcb_fidelity_eval, state by
state, at the 1e-12 relative tolerance the other fidelity tests use -- and, so that the check is not
only one CUDA path against another, with the CPU oracle's value and gradient as well.

The library works on its own stream: the input and output tensors are made idle
(torch.cuda.synchronize) before they are handed over, as include/qcb200.h asks of every
device-buffer call.
"""
import ctypes

import numpy as np
import pytest

import qcb200
from qcb200 import Fidelity, workloads, _lib
import qc_oracle as O
from conftest import rel_err

pytestmark = pytest.mark.gpu


def _reference(f, X):
    out = [f.eval(x) for x in X]                       # existing entry point, one state at a time
    return np.array([o[0] for o in out]), np.stack([o[1] for o in out])


def _oracle(kind, X, goal, n_full=None, subspace=None):
    out = [O.fidelity_vgh(kind, x, goal, n_full=n_full, subspace=subspace) for x in X]
    return np.array([o[0] for o in out]), np.stack([o[1] for o in out])


def test_noalloc_batch_matches_existing_entry_point_twice(built_libs):
    import torch
    lib = _lib.load()
    rng = np.random.default_rng(42)
    goal = workloads.haar_unitary(3, rng)
    f = Fidelity(Fidelity.UNITARY, goal)
    n = f.x_dim
    for count in (5, 3):                               # two calls in a row, different sizes and data
        X = rng.normal(size=(count, n))
        dX = torch.from_numpy(X).cuda()
        dF = torch.full((count,), np.nan, dtype=torch.float64, device="cuda")
        dg = torch.full((count, n), np.nan, dtype=torch.float64, device="cuda")
        torch.cuda.synchronize()                       # buffers idle before the library's stream touches them
        st = lib.qcb_fidelity_eval_batch_device_noalloc(f.h, dX.data_ptr(), count, dF.data_ptr(), dg.data_ptr())
        assert st == 0
        F0, g0 = _reference(f, X)
        assert rel_err(dF.cpu().numpy(), F0) < 1e-12
        assert rel_err(dg.cpu().numpy(), g0) < 1e-12
        F1, g1 = _oracle(O.KIND_UNITARY, X, goal)
        assert rel_err(dF.cpu().numpy(), F1) < 1e-12 and rel_err(dg.cpu().numpy(), g1) < 1e-12
    # the subspace kind has many linear forms per state
    goal_s = workloads.haar_unitary(2, rng)
    fs = Fidelity(Fidelity.UNITARY_SUBSPACE, goal_s, n=3, subspace=[0, 1])
    X = rng.normal(size=(4, fs.x_dim))
    dX = torch.from_numpy(X).cuda()
    dF = torch.empty(4, dtype=torch.float64, device="cuda")
    dg = torch.empty(4, fs.x_dim, dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    assert lib.qcb_fidelity_eval_batch_device_noalloc(fs.h, dX.data_ptr(), 4, dF.data_ptr(), dg.data_ptr()) == 0
    F0, g0 = _reference(fs, X)
    assert rel_err(dF.cpu().numpy(), F0) < 1e-12 and rel_err(dg.cpu().numpy(), g0) < 1e-12
    F1, g1 = _oracle(O.KIND_UNITARY_SUBSPACE, X, goal_s, n_full=3, subspace=[0, 1])
    assert rel_err(dF.cpu().numpy(), F1) < 1e-12 and rel_err(dg.cpu().numpy(), g1) < 1e-12


def test_noalloc_rejects_batches_beyond_the_fixed_scratch(built_libs):
    import torch
    lib = _lib.load()
    f = Fidelity(Fidelity.KET, np.array([1, 0], dtype=complex))
    dX = torch.zeros(8, f.x_dim, dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    too_many = 65536 // 2 + 1                          # R = 2 forms per ket -> one past the limit
    st = lib.qcb_fidelity_eval_batch_device_noalloc(f.h, dX.data_ptr(), too_many, None, None)
    assert st == 1                                     # QCB_ERR_ARG, nothing launched
    assert lib.qcb_fidelity_eval_batch_device_noalloc(f.h, dX.data_ptr(), 0, None, None) == 1
