"""Multi-GPU drivers on real GPUs (skipped on a one-GPU box): NCCL broadcast + gather
(SlabCoordinator, what north_star prescribes), COO slabs stored over NVLink (PeerDelivery) and compact
arenas stored over NVLink then assembled on the Ipopt rank (ArenaDelivery), each bit-identical to a
single-handle evaluation.  One process per GPU under torch.distributed.run."""
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT, rel_err

pytestmark = pytest.mark.gpu


def test_two_gpu_drivers_match_single_handle(built_libs):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29731", os.path.join(ROOT, "tests", "helpers", "multi_gpu_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    for what in ("SlabCoordinator cfg5", "PeerDelivery cfg5", "ArenaDelivery cfg5 N=21 pipelined", "ArenaDelivery cfg5 N=21 plain",
                 "ArenaDelivery cfg5 N=34 pipelined", "ArenaDelivery cfg5 N=34 plain"):
        assert f"ok {what}" in r.stdout, r.stdout[-2000:]


def test_arena_delivery_on_one_gpu_is_bit_identical(built_libs):
    """qcb_eval_arenas_device + qcb_assemble_device (the two halves of the compact multi-GPU delivery) on
    one device: the COO vectors assembled from the delivered arenas equal those of qcb_eval_all_device."""
    import torch
    from qcb200 import Evaluator, workloads
    w = workloads.make("cfg5", N=9)
    L = w.layout
    ev = Evaluator(L, w.G_drift, w.G_drives)
    JA, HA = ev.arena_doubles()
    dZ, dmu = torch.from_numpy(w.Z).cuda(), torch.from_numpy(w.mu).cuda()
    z = lambda n: torch.full((n,), float("nan"), dtype=torch.float64, device="cuda")
    d0, j0, h0 = z(ev.defect_dim), z(ev.jacobian_nnz), z(ev.hessian_nnz)
    ev.eval_all_device(dZ, dmu, d0, j0, h0)
    d1, ja, ha = z(ev.defect_dim), z(L.K * JA), z(L.K * HA)
    ev.eval_arenas_device(dZ, dmu, d1, ja, ha)
    j1, h1 = z(ev.jacobian_nnz), z(ev.hessian_nnz)
    ev.assemble_device(ja, ha, L.K, L.K, j1, h1)
    assert torch.equal(d0, d1) and torch.equal(j0, j1) and torch.equal(h0, h1)
    # the same in the pipelined form: propagators first (E part), then the rest, then the Hessian side; the
    # Jacobian is assembled in two passes (entries copied from E, then all others)
    ja2, ha2, d2 = z(L.K * JA), z(L.K * HA), z(ev.defect_dim)
    ev.set_async(True)
    ev.set_arena_parts(1); ev.eval_arenas_device_x(dZ, None, True, d2, ja2, None)
    ev.set_arena_parts(2); ev.eval_arenas_device_x(dZ, None, False, d2, ja2, None)
    ev.set_arena_parts(3); ev.eval_arenas_device_x(dZ, dmu, False, d2, None, ha2)
    ev.wait(1)
    jp = z(ev.jacobian_nnz)
    ev.assemble_device(ja2, None, L.K, L.K, jp, None, parts=1)
    ev.wait(0)
    xblock = L.c * L.d * L.d                      # block-diagonal structure: the x-block comes first in every unit
    jp_v, j0_v = jp.view(L.K, -1), j0.view(L.K, -1)
    assert torch.equal(jp_v[:, :xblock], j0_v[:, :xblock]) and torch.isnan(jp_v[:, xblock:]).all()
    hp_ = z(ev.hessian_nnz)
    ev.assemble_device(ja2, None, L.K, L.K, jp, None, parts=2)
    ev.assemble_device(None, ha2, L.K, L.K, None, hp_)
    ev.wait(0)
    ev.set_async(False)
    assert torch.equal(jp, j0) and torch.equal(hp_, h0) and torch.equal(d2, d0)
    # part of the trajectory, as the Ipopt rank does for a foreign slab
    j2, h2 = z(3 * L.jac_nnz_unit), z(3 * L.hess_nnz_unit)
    ev.assemble_device(ja[2 * JA:], ha[2 * HA:], 3, 3, j2, h2)
    assert torch.equal(j2, j0[2 * L.jac_nnz_unit:5 * L.jac_nnz_unit]) and torch.equal(h2, h0[2 * L.hess_nnz_unit:5 * L.hess_nnz_unit])
    small = Evaluator(workloads.make("cfg2", N=5).layout, *[getattr(workloads.make("cfg2", N=5), a) for a in ("G_drift", "G_drives")])
    import qcb200
    with pytest.raises(qcb200.QcbError):
        small.eval_arenas_device(dZ, None, None, ja, None)
    ev.close(); small.close()
