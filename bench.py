#!/usr/bin/env python
"""bench.py -- knot-point evaluations per second (defect + Jacobian + Hessian-of-Lagrangian).

One "step" = one full evaluation pass of the hot path over the synthetic trajectory of the
named workload (default: BASELINE.json's headline, cfg5 = 6-qubit coupled-transmon unitary,
d = 128, 12 drives, T = 2000 -> 1999 units), i.e. what Ipopt triggers with one eval_g +
eval_jac_g + eval_h round.

  python bench.py --gpus 1 --steps K --warmup W                 # the CUDA library
  python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...
  python bench.py --impl reference ...                          # CPU port on the host cores

`value`  : device-resident throughput (inputs and outputs in HBM), CUDA events on the library's
           stream, max over ranks.  N > 1: the T = 2000 trajectory is cut into N contiguous time
           slabs, one per GPU ("strong" scaling), no data-path collective.
`e2e`    : same metric through the public host-buffer call (Evaluator.eval_all ->
           qcb_eval_all): pinned host Z / mu in, H2D, kernels, D2H of delta / Jacobian /
           Hessian values into pinned host buffers, every step.  N > 1: every rank copies its
           slab over its own PCIe link into one shared pinned host buffer.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOAD_DESC = {
    "cfg1": "single-qubit Hadamard, X/Y drives, T=50 (README example)",
    "cfg2": "3-level transmon X gate, T=200",
    "cfg3": "two transmon qutrits CZ (d=18) unitary, T=500",
    "cfg3_ket": "two transmon qutrits (d=18) ket, T=500",
    "cfg4": "robust single-qubit gate, 256 samples, T=400 (knot x sample batching)",
    "cfg5": "6-qubit coupled-transmon unitary (dim 64, real-iso 128), 12 drives, T=2000",
}


def algorithmic_bytes_per_unit(L):
    """8 * [(3 x_dim + m + 1) inputs + x_dim defect + Jacobian nnz + Hessian nnz]."""
    return 8 * (3 * L.x_dim + L.m + 1 + L.x_dim + L.jac_nnz_unit + L.hess_nnz_unit)


def survey_flops_per_unit(L):
    """SURVEY.md 8d convention: 2 d^3 * 10 * (1 + 6m) + 2 d^2 c (2m + 4)."""
    return 2 * L.d ** 3 * 10 * (1 + 6 * L.m) + 2 * L.d ** 2 * L.c * (2 * L.m + 4)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.tmp = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.proc = None
        self.gpu_index = gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu_index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "50"],
                stdout=self.tmp, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.tmp.flush()
        rows = [r.strip().split(",") for r in open(self.tmp.name) if r.strip()]
        os.unlink(self.tmp.name)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
                for nm, v in zip(names, r[5:9]):
                    if v.strip().lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_port_throughput(w, sample_units, threads, repeats=1, blas=True):
    """Oracle port (oracle/qc_port.c, built -march=native on this box) on the host cores; its matrix
    products go through numpy's bundled OpenBLAS (single-threaded per call, OpenMP over units) unless
    blas=False (the port's own loops)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import qc_port
    import qc_oracle
    L = w.layout
    OL = qc_oracle.Layout(L.d, L.c, L.m, L.N, L.S, L.z_dim, L.x_off, L.u_off, L.dt_off, L.jac_mode)
    try:
        qc_port.build(native=True)
        native = True
    except Exception:
        native = False
    blas_desc = ""
    try:
        blas_desc = qc_port.use_numpy_openblas(native, enable=blas)
    except Exception:
        blas_desc = ""
    nk = max(1, min(L.K, sample_units // L.S))
    best = None
    qc_port.slot_codes(OL)          # structure set-up is not part of an evaluation
    out = (np.empty(L.S * nk * L.x_dim), np.empty(L.S * nk * L.jac_nnz_unit), np.empty(L.S * nk * L.hess_nnz_unit))
    for o in out:
        o.fill(0.0)                 # touch the pages: allocation is not part of an evaluation either
    for _ in range(repeats):
        t = time.perf_counter()
        # defect + Jacobian + Hessian COO value vectors, as the reference's evaluator hands them to Ipopt
        qc_port.evaluate_coo(OL, w.G_drift, w.G_drives, w.Z, w.mu, (0, nk), threads, native, out)
        dt = time.perf_counter() - t
        best = dt if best is None else min(best, dt)
    return nk * L.S / best, nk * L.S, native, blas_desc


def port_flops_per_unit(L, squarings=0):
    """Matrix-product flops the port executes per unit (full storage, sparse drive products excluded):
    (30 + 3 s) per-unit products and (11 + 3 s) per-drive products of 2 d^3, plus the panel products."""
    return 2.0 * L.d ** 3 * ((30 + 3 * squarings) + L.m * (11 + 3 * squarings)) + 2.0 * L.d ** 2 * L.c * (2 * L.m + 4)


def cpu_sample_units(L, cores, override=None):
    """Units per CPU evaluation: the same for the cpu_baseline leg and for --impl reference."""
    return override or (32 * cores if L.d >= 64 else 8192 * cores)


def run_reference(args):
    """--impl reference: the CPU port of the path on the box's host cores (the reference's own
    Julia evaluator cannot run in this image: no Julia, un-vendored deps -- see DESIGN.md)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import qcb200
    from qcb200 import workloads
    cores = os.cpu_count() or 1
    w = workloads.make(args.workload, N=args.knots)
    L = w.layout
    per_step = max(cores, min(L.S * L.K, cpu_sample_units(L, cores, args.ref_units)))
    rates = []
    for i in range(args.warmup + args.steps):
        r, n_units, native, blas_desc = cpu_port_throughput(w, per_step, cores)
        if i >= args.warmup:
            rates.append(r)
    v = float(np.mean(rates))
    line = {
        "impl": "reference", "metric": "knot-pt evals/sec (defect+Jac+Hess)", "value": v, "unit": "knot-evals/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * n_units / v,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"{args.workload}: {WORKLOAD_DESC.get(args.workload, '')}", "units": L.S * L.K,
                   "units_per_step_sample": n_units, "d": L.d, "c": L.c, "m": L.m, "N": L.N, "S": L.S},
        "cpu_baseline": {"value": v, "unit": "knot-evals/s", "cores": cores, "kind": "port",
                         "gemm": blas_desc or "the port's own loops",
                         "gflops": v * port_flops_per_unit(L) / 1e9,
                         "sample": f"first {n_units} of {L.S * L.K} units per step (units are independent: a per-unit rate), "
                                   f"OpenMP over units, {'-march=native' if native else '-march=x86-64-v3'}"},
        "e2e": {"value": v, "unit": "knot-evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--workload", default="cfg5", choices=sorted(WORKLOAD_DESC))
    ap.add_argument("--knots", type=int, default=None, help="override the number of knots (testing)")
    ap.add_argument("--ref-units", type=int, default=None)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-mix", action="store_true", help="skip the per-callback timings (profiling runs)")
    ap.add_argument("--out-mem", default="auto", choices=["auto", "pinned", "pageable"])
    ap.add_argument("--dt-scale", type=float, default=1.0,
                    help="multiply every timestep of the synthetic trajectory (sensitivity of the headline to ||dt*G||: "
                         "each factor of 2 adds one squaring level)")
    ap.add_argument("--no-small", action="store_true", help="skip the small-d configs (cfg1-4) side table")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)

    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    import qcb200
    from qcb200 import workloads, Evaluator

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU port")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    w = workloads.make(args.workload, N=args.knots)
    L = w.layout
    if args.dt_scale != 1.0:
        w.Z.reshape(L.N, L.z_dim)[:, L.dt_off] *= args.dt_scale
    K = L.K
    k0, k1 = (K * rank) // world, (K * (rank + 1)) // world
    ev = Evaluator(L, w.G_drift, w.G_drives, device=local, k_begin=k0, k_end=k1)
    units_total = L.S * K

    # ---------------- device-resident throughput (`value`) ----------------
    dZ = torch.from_numpy(w.Z).to(dev)
    dmu = torch.from_numpy(w.mu).to(dev)
    d_delta = torch.empty(ev.defect_dim, dtype=torch.float64, device=dev)
    d_jac = torch.empty(ev.jacobian_nnz, dtype=torch.float64, device=dev)
    d_hess = torch.empty(ev.hessian_nnz, dtype=torch.float64, device=dev)
    st = torch.cuda.ExternalStream(ev.stream, device=dev)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()             # covers the warm-up evaluations and the timed region (same load)
    for _ in range(max(args.warmup, 3)):
        ev.eval_all_device(dZ, dmu, d_delta, d_jac, d_hess)
    # live FP64 GEMM peak (MEASURED_PEAKS.json has no FP64 figure)
    dgemm_tflops = None
    if rank == 0:
        a = torch.randn(8192, 8192, dtype=torch.float64, device=dev)
        b = torch.randn(8192, 8192, dtype=torch.float64, device=dev)
        for _ in range(2):
            torch.matmul(a, b)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            torch.matmul(a, b)
        e1.record(); e1.synchronize()
        dgemm_tflops = 3 * 2 * 8192 ** 3 / (e0.elapsed_time(e1) * 1e-3) / 1e12
        del a, b
    launches0 = ev.kernel_launches
    barrier()
    step_bytes = algorithmic_bytes_per_unit(L) * L.S * (k1 - k0)
    flush = None
    if step_bytes < 2 * 126e6:      # working set could sit in the 126 MB L2: flush it between steps
        flush = torch.empty(64 * 1024 * 1024, dtype=torch.float64, device=dev)   # 512 MB
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ms_sum = 0.0
    for _ in range(args.steps):
        if flush is not None:
            flush.add_(1.0)
            torch.cuda.synchronize()
        e0.record(st)
        ev.eval_all_device(dZ, dmu, d_delta, d_jac, d_hess)
        e1.record(st)
        e1.synchronize()
        ms_sum += e0.elapsed_time(e1)
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms_dev = max_over_ranks(ms_sum) / args.steps
    launches = ev.kernel_launches - launches0
    value = units_total / (ms_dev * 1e-3)
    # ---- per-launch timing of the dominant kernel: the same steps again, right after the timed region,
    #      with a CUDA-event pair around every GEMM launch on the library's stream.  The timed region
    #      itself overlaps the squaring classes on two streams (per-kernel durations are not separable
    #      there); with profiling on the library serialises them, so every kernel is timed alone. ----
    ev.profile(True)
    ms_prof = 0.0
    for _ in range(args.steps):
        if flush is not None:
            flush.add_(1.0)
            torch.cuda.synchronize()
        e0.record(st)
        ev.eval_all_device(dZ, dmu, d_delta, d_jac, d_hess)
        e1.record(st)
        e1.synchronize()
        ms_prof += e0.elapsed_time(e1)
    ms_prof /= args.steps
    prof = ev.profile_read()
    ev.profile(False)

    # ---- evaluator time per Ipopt iteration with the callback mix the reference's logs show
    #      (55 eval_g : 51 eval_jac_g : 50 eval_h per 50 iterations, multilevel_transmon.jl:147-153) ----
    def timed(fn, reps=2):
        fn()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(st)
        for _ in range(reps):
            fn()
        b.record(st); b.synchronize()
        return max_over_ranks(a.elapsed_time(b)) / reps
    callback_mix = None
    if not args.no_mix:
        # Ipopt's protocol on an iterate: eval_g(new_x = 1) first, then eval_jac_g and eval_h with new_x = 0 --
        # the library keeps what depends on Z alone between them (qcb_eval_all_device_x).  The three calls
        # are timed in that order, each iteration starting from a new_x = 1 call, so that nothing is reused
        # that a real iteration would not have.
        def protocol(reps=3):
            ev3 = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
            acc = [0.0, 0.0, 0.0]
            for it in range(reps + 1):
                ev3[0].record(st)
                ev.eval_all_device_x(dZ, None, True, d_delta, None, None)
                ev3[1].record(st)
                ev.eval_all_device_x(dZ, None, False, d_delta, d_jac, None)
                ev3[2].record(st)
                ev.eval_all_device_x(dZ, dmu, False, d_delta, None, d_hess)
                ev3[3].record(st); ev3[3].synchronize()
                if it:      # first round is a warm-up
                    for q in range(3):
                        acc[q] += ev3[q].elapsed_time(ev3[q + 1])
            return [max_over_ranks(a) / reps for a in acc]
        t_g, t_j, t_h = protocol()
        t_j1 = timed(lambda: ev.eval_all_device(dZ, None, d_delta, d_jac, None))
        t_h1 = timed(lambda: ev.eval_all_device(dZ, dmu, d_delta, None, d_hess))
        callback_mix = {"eval_g_ms": t_g, "eval_jac_g_ms": t_j, "eval_h_ms": t_h,
                        "evaluator_ms_per_ipopt_iteration": (55 * t_g + 51 * t_j + 50 * t_h) / 50.0,
                        "eval_jac_g_ms_recomputing": t_j1, "eval_h_ms_recomputing": t_h1,
                        "evaluator_ms_per_ipopt_iteration_recomputing": (55 * t_g + 51 * t_j1 + 50 * t_h1) / 50.0,
                        "protocol": "per iteration eval_g with new_x = 1, then eval_jac_g and eval_h with new_x = 0 (propagators, "
                                    "then tangents and the Jacobian arena, are reused), timed in that order; the *_recomputing "
                                    "figures are the same callbacks without reuse",
                        "mix": "55 g : 51 jac : 50 hess per 50 iterations (docs/literate/examples/multilevel_transmon.jl:147-153)"}
        # objective side of the same log (55 f : 51 grad f): terminal infidelity + the three quadratic
        # regularisers, through the host-buffer entry points (their transfers included)
        if rank == 0 and L.S == 1 and w.kind == "unitary" and L.z_dim >= L.u_off + 3 * L.m:
            try:
                from qcb200 import Aux, Fidelity
                fid = Fidelity(Fidelity.UNITARY, w.U_goal)
                aux = Aux(L.N, L.z_dim, L.dt_off)
                terms = [(L.u_off + i * L.m, L.m, 1e-2) for i in range(3)]
                xN = np.ascontiguousarray(w.Z.reshape(L.N, L.z_dim)[-1, L.x_off[0]:L.x_off[0] + L.x_dim])

                def objective():
                    F, g, _ = fid.eval(xN)
                    J, gz, _ = aux.regularizer_eval(w.Z, terms, want_hessian=False)
                    return 100.0 * abs(1.0 - F) + J
                objective()
                t0 = time.perf_counter()
                for _ in range(3):
                    objective()
                t_f = (time.perf_counter() - t0) / 3 * 1e3
                callback_mix["eval_f_and_grad_f_ms"] = t_f
                callback_mix["objective_ms_per_ipopt_iteration"] = 55 * t_f / 50.0
                callback_mix["objective_note"] = ("one call returns value and gradient (terminal infidelity on x_N + quadratic "
                                                  "regularisers on u, du, ddu; host buffers, wall clock)")
                aux.close()
            except Exception as exc:      # the objective side is auxiliary to the metric: never fail the bench on it
                callback_mix["objective_note"] = f"not measured: {exc}"

    # ---- N > 1: the data plane north_star prescribes -- Z and mu broadcast from the rank that hosts Ipopt,
    #      every value vector complete in that rank's HBM at the end of the step.  Three drivers:
    #      arena   compact value arenas stored over NVLink by the producing kernels, COO vectors assembled on
    #              the Ipopt rank (the product path; this is the N > 1 `value`);
    #      peer    the COO slabs themselves stored over NVLink (5x more bytes through one GPU's links);
    #      nccl    plain NCCL broadcast + gather (SlabCoordinator), the baseline the fused forms must beat. ----
    peer = plane = None
    if world > 1 and L.S == 1:
        from qcb200 import distributed as qdist

        def wall(fn, steps):
            fn()
            barrier()
            t0 = time.perf_counter()
            for _ in range(steps):
                fn()
            torch.cuda.synchronize()
            return max_over_ranks(time.perf_counter() - t0) / steps

        # inputs live on the Ipopt rank only; the broadcast is inside every timed step
        bZ = dZ if rank == 0 else torch.empty_like(dZ)
        bmu = dmu if rank == 0 else torch.empty_like(dmu)
        del d_jac, d_hess               # the drivers own their result buffers
        torch.cuda.empty_cache()
        # (a smaller slab on the Ipopt rank, to make room for its assembly work, was measured slower at 2, 4 and
        # 8 GPUs: the other ranks' compute + delivery chain is the critical path, not the root -- DESIGN.md 10)
        ad = qdist.ArenaDelivery(ev, L, local)

        def arena_step():
            ad.broadcast(bZ, bmu)
            ad.evaluate(bZ, bmu)
        dt_arena = wall(arena_step, args.steps)
        dt_bcast = wall(lambda: ad.broadcast(bZ, bmu), args.steps)
        plane = {"value": units_total / dt_arena, "unit": "knot-evals/s", "ms_per_step": 1e3 * dt_arena,
                 "broadcast_ms": 1e3 * dt_bcast,
                 "bytes_over_nvlink_per_step": int(8 * (K - (k1 - k0)) * (sum(ev.arena_doubles()) + L.x_dim)),
                 "what": "NCCL broadcast of Z and mu from rank 0 + slab evaluation + value arenas stored into rank 0's HBM by "
                         "the producing kernels (CUDA IPC peer mapping) + COO assembly on rank 0; ends with every value vector "
                         "complete in rank 0's HBM"}
        barrier()
        ad.close()
        pd = qdist.PeerDelivery(ev, L, local)
        dtp = wall(lambda: pd.evaluate(dZ, dmu), max(1, args.steps // 2))
        peer = {"value": units_total / dtp, "unit": "knot-evals/s", "ms_per_step": 1e3 * dtp,
                "what": "COO slabs stored into rank 0's HBM through IPC-mapped peer pointers (no broadcast in this figure)"}
        barrier()
        pd.close()
        try:        # NCCL baseline: broadcast + three padded gathers (needs the full vectors twice on rank 0)
            co = qdist.SlabCoordinator(L, qdist.gpu_compute(ev), dev)
            dtn = wall(lambda: co.evaluate(dZ, dmu) if rank == 0 else co.evaluate(), 1)
            plane["nccl_baseline"] = {"value": units_total / dtn, "ms_per_step": 1e3 * dtn,
                                      "what": "SlabCoordinator: dist.broadcast(Z, mu) + dist.gather of the COO slabs"}
        except Exception as exc:
            plane["nccl_baseline"] = {"value": None, "note": f"not measured: {type(exc).__name__}: {exc}"[:300]}
            torch.cuda.empty_cache()
        barrier()
        d_jac = d_hess = None

    # ---------------- end-to-end through the host-buffer API (`e2e`) ----------------
    e2e = None
    if not args.no_e2e:
        n_delta, n_jac, n_hess = units_total * L.x_dim, units_total * L.jac_nnz_unit, units_total * L.hess_nnz_unit
        if world == 1:
            hZ = torch.from_numpy(w.Z).pin_memory()
            hmu = torch.from_numpy(w.mu).pin_memory()
            # result buffers: pinned for the direct-DMA route (small-d path); ordinary (pageable,
            # huge-page eligible) memory when the library writes them with the host cores
            # (large-d path, compact transfer) -- which is also what Ipopt's own arrays are
            out_pinned = args.out_mem != "pageable"      # measured: pinned is faster on both routes
            # the whole smooth-pulse problem behind one call per Ipopt callback (qcb_problem_*): integrator +
            # derivative rows (u, du), (du, ddu) + timestep rows; terminal infidelity + regularisers on u, du, ddu.
            # Its vectors are the integrator's plus a few thousand entries: the result buffers below are sized
            # for it and shared with the plain e2e calls.
            prob = prob_pieces = None
            extra = (0, 0, 0)
            if (not args.no_mix and out_pinned and L.S == 1 and w.kind == "unitary" and L.z_dim >= L.u_off + 3 * L.m
                    and ev.defect_dim == n_delta):
                try:
                    from qcb200 import Aux, Fidelity, Problem
                    m_ = L.m
                    aux_p = Aux(L.N, L.z_dim, L.dt_off, derivatives=[(L.u_off, L.u_off + m_, m_), (L.u_off + m_, L.u_off + 2 * m_, m_)],
                                device=local)
                    fid_p = Fidelity(Fidelity.UNITARY, w.U_goal if w.U_goal is not None else np.eye(L.c, dtype=complex), device=local)
                    prob = Problem(L.N, L.z_dim, [ev], aux_p, derivative_rows=True, dt_equal_rows=True, fidelity=fid_p,
                                   fid_index=np.arange(L.x_off[0], L.x_off[0] + L.x_dim), Q=100.0,
                                   fid_hessian=L.x_dim <= 1024, reg_terms=[(L.u_off + i * m_, m_, 1e-2) for i in range(3)])
                    prob_pieces = (aux_p, fid_p)
                    extra = (prob.m - n_delta, prob.jacobian_nnz - n_jac, prob.hessian_nnz - n_hess)
                except Exception as exc:
                    prob = None
                    print(f"# problem-level callbacks not measured: {exc}", file=sys.stderr)
            if out_pinned:
                h_delta = torch.empty(n_delta + extra[0], dtype=torch.float64, pin_memory=True)
                h_jac = torch.empty(n_jac + extra[1], dtype=torch.float64, pin_memory=True)
                h_hess = torch.empty(n_hess + extra[2], dtype=torch.float64, pin_memory=True)
            else:
                h_delta = torch.from_numpy(np.zeros(n_delta))
                h_jac = torch.from_numpy(np.zeros(n_jac))
                h_hess = torch.from_numpy(np.zeros(n_hess))
            ptr = lambda t, off=0: t.data_ptr() + 8 * off
            pZ, pmu = ptr(hZ), ptr(hmu)
            p_delta, p_jac, p_hess = ptr(h_delta), ptr(h_jac), ptr(h_hess)
        else:
            # one shared pinned host buffer; every rank writes its slab over its own PCIe link
            from multiprocessing import shared_memory
            sizes = [len(w.Z), len(w.mu), n_delta, n_jac, n_hess]
            PG = (1 << 21) // 8                       # every array starts on a 2 MiB boundary
            starts, tot = [], 0
            for n_ in sizes:
                starts.append(tot)
                tot += (n_ + PG - 1) // PG * PG
            offs = starts + [tot]
            name = f"qcb200_bench_{os.environ.get('MASTER_PORT', '0')}"
            if rank == 0:
                shm = shared_memory.SharedMemory(create=True, size=tot * 8, name=name)
            dist.barrier()
            if rank != 0:
                shm = shared_memory.SharedMemory(name=name)
                try:    # only the creator (rank 0) unlinks: keep the other ranks' trackers quiet
                    from multiprocessing import resource_tracker
                    resource_tracker.unregister(shm._name, "shared_memory")
                except Exception:
                    pass
            buf = np.ndarray((tot,), dtype=np.float64, buffer=shm.buf)
            if rank == 0:
                buf[offs[0]:offs[0] + len(w.Z)] = w.Z
                buf[offs[1]:offs[1] + len(w.mu)] = w.mu
            base = buf.ctypes.data
            # pin only what this rank touches (inputs + its own output slabs), page-rounded
            assert L.S == 1, "slab-contiguous outputs need S == 1"
            regs = [(offs[0], offs[2]),
                    (offs[2] + k0 * L.x_dim, offs[2] + k1 * L.x_dim),
                    (offs[3] + k0 * L.jac_nnz_unit, offs[3] + k1 * L.jac_nnz_unit),
                    (offs[4] + k0 * L.hess_nnz_unit, offs[4] + k1 * L.hess_nnz_unit)]
            registered = []
            for lo, hi in regs:
                lo_b, hi_b = (8 * lo) // 4096 * 4096, (8 * hi + 4095) // 4096 * 4096
                rc = torch.cuda.cudart().cudaHostRegister(base + lo_b, hi_b - lo_b, 0)
                assert int(rc) == 0, f"cudaHostRegister failed: {int(rc)}"
                registered.append(base + lo_b)
            dist.barrier()
            pZ, pmu = base + 8 * int(offs[0]), base + 8 * int(offs[1])
            p_delta = base + 8 * (int(offs[2]) + k0 * L.x_dim)
            p_jac = base + 8 * (int(offs[3]) + k0 * L.jac_nnz_unit)
            p_hess = base + 8 * (int(offs[4]) + k0 * L.hess_nnz_unit)
        # warm-up: the first call allocates the staging buffers, the next eight let the handle choose its number of
        # host store threads (qcb_host_threads_in_use; small-d path: nothing to choose)
        for _ in range(10):
            ev.eval_all_raw(pZ, pmu, p_delta, p_jac, p_hess)
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            ev.eval_all_raw(pZ, pmu, p_delta, p_jac, p_hess)
        torch.cuda.synchronize()
        dt = max_over_ranks(time.perf_counter() - t0)
        barrier()
        tm = ev.timings()
        h2d, d2h = ev.transfer_bytes()            # what the last call really moved over PCIe
        host_written = 8 * (ev.defect_dim + ev.jacobian_nnz + ev.hessian_nnz)
        if world > 1:
            t_h2d = torch.tensor([h2d, d2h, host_written], dtype=torch.float64, device=dev)
            dist.all_reduce(t_h2d)
            h2d, d2h, host_written = int(t_h2d[0].item()), int(t_h2d[1].item()), int(t_h2d[2].item())
            # spot check: the shared buffer holds every rank's slab
            if rank == 0:
                chk = buf[offs[2]:offs[2] + n_delta]
                assert np.isfinite(chk).all() and np.abs(chk[-L.x_dim:]).sum() > 0 and np.abs(chk[:L.x_dim]).sum() > 0
            for ptr_ in registered:
                torch.cuda.cudart().cudaHostUnregister(ptr_)
            del buf
            dist.barrier()
            shm.close()
            if rank == 0:
                shm.unlink()
        # host-write roofline: the same streaming stores, the same thread pool, the same result buffer
        host_gbs = None
        if world == 1:
            try:
                lib = qcb200.load_library()
                import ctypes
                sec = ctypes.c_double(0.0)
                nbytes = min(8 * n_jac, 8 << 30)
                best = None
                for _ in range(3):
                    lib.qcb_host_stream_fill(p_jac, nbytes, ctypes.byref(sec))
                    best = sec.value if best is None else min(best, sec.value)
                host_gbs = nbytes / best / 1e9
            except Exception:
                host_gbs = None
        # the same three callbacks through HOST buffers in Ipopt's order (what the Julia shim calls)
        host_mix = None
        if world == 1 and not args.no_mix:
            try:
                lib = qcb200.load_library()
                for _ in range(9):      # lets the Jacobian-only calls settle their thread count too
                    lib.qcb_eval_jacobian_x(ev.h, pZ, 1, p_jac)
                acc = [0.0, 0.0, 0.0]
                for it in range(3):
                    t0 = time.perf_counter()
                    lib.qcb_eval_defect_x(ev.h, pZ, 1, p_delta)
                    t1 = time.perf_counter()
                    lib.qcb_eval_jacobian_x(ev.h, pZ, 0, p_jac)
                    t2 = time.perf_counter()
                    lib.qcb_eval_hessian_x(ev.h, pZ, 0, pmu, p_hess)
                    t3 = time.perf_counter()
                    if it:
                        acc = [acc[0] + t1 - t0, acc[1] + t2 - t1, acc[2] + t3 - t2]
                g_ms, j_ms, h_ms = (1e3 * a / 2 for a in acc)
                host_mix = {"eval_g_ms": g_ms, "eval_jac_g_ms": j_ms, "eval_h_ms": h_ms,
                            "evaluator_ms_per_ipopt_iteration": (55 * g_ms + 51 * j_ms + 50 * h_ms) / 50.0,
                            "what": "host buffers, wall clock, new_x = 1 / 0 / 0 (qcb_eval_defect_x, _jacobian_x, _hessian_x)"}
            except Exception as exc:
                host_mix = {"note": f"not measured: {exc}"[:200]}
        # the five Ipopt callbacks of the whole problem, one C call each, in Ipopt's order on one iterate
        problem_mix = None
        if world == 1 and prob is not None:
            try:
                lam = torch.zeros(prob.m, dtype=torch.float64, pin_memory=True)
                lam[:n_delta] = hmu
                lam[n_delta:] = 0.5
                grad = np.empty(prob.n)
                import ctypes
                lib = qcb200.load_library()
                fval = ctypes.c_double(0.0)
                acc = [0.0] * 5
                for it in range(3):
                    t = [time.perf_counter()]
                    lib.qcb_problem_eval_f(prob.h, pZ, 1, ctypes.byref(fval)); t.append(time.perf_counter())
                    lib.qcb_problem_eval_grad_f(prob.h, pZ, 0, grad.ctypes.data); t.append(time.perf_counter())
                    lib.qcb_problem_eval_g(prob.h, pZ, 0, p_delta); t.append(time.perf_counter())
                    lib.qcb_problem_eval_jac_g(prob.h, pZ, 0, p_jac); t.append(time.perf_counter())
                    lib.qcb_problem_eval_h(prob.h, pZ, 0, 1.0, lam.data_ptr(), p_hess); t.append(time.perf_counter())
                    if it:
                        acc = [a + 1e3 * (t[q + 1] - t[q]) / 2 for q, a in enumerate(acc)]
                problem_mix = {"eval_f_ms": acc[0], "eval_grad_f_ms": acc[1], "eval_g_ms": acc[2], "eval_jac_g_ms": acc[3],
                               "eval_h_ms": acc[4],
                               "ms_per_ipopt_iteration": (55 * acc[0] + 51 * acc[1] + 55 * acc[2] + 51 * acc[3] + 50 * acc[4]) / 50.0,
                               "n": prob.n, "m": prob.m, "nnz_jac": prob.jacobian_nnz, "nnz_hess": prob.hessian_nnz,
                               "what": "qcb_problem_eval_f / _grad_f / _g / _jac_g / _h: integrator + derivative rows + timestep "
                                       "rows, terminal infidelity + regularisers on u, du, ddu; host buffers, wall clock, "
                                       "new_x = 1 on eval_f and 0 after; mix 55 f : 51 grad f : 55 g : 51 jac : 50 hess per 50 "
                                       "iterations (docs/literate/examples/multilevel_transmon.jl:147-153)"}
                prob.close()
                for p_ in prob_pieces:
                    p_.close()
            except Exception as exc:
                problem_mix = {"note": f"not measured: {exc}"[:200]}
        e2e = {"value": units_total / (dt / args.steps), "unit": "knot-evals/s", "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": int(d2h), "host_result_bytes_per_step": int(host_written),
               "result_buffers": ("shared pinned" if world > 1 else ("pinned" if out_pinned else "pageable")),
               "ms_per_step": 1e3 * dt / args.steps,
               "last_call_rank0_ms": tm}
        if host_mix:
            e2e["callback_mix_host"] = host_mix
        if problem_mix:
            e2e["problem_callbacks"] = problem_mix
        if host_gbs:
            e2e["host_threads"] = int(qcb200.load_library().qcb_host_threads())
            e2e["host_threads_in_use"] = {"with_hessian": int(qcb200.load_library().qcb_host_threads_in_use(ev.h, 1)),
                                          "jacobian_only": int(qcb200.load_library().qcb_host_threads_in_use(ev.h, 0))}
            e2e["host_write_gbs"] = host_gbs
            e2e["achieved_host_write_gbs"] = host_written / (dt / args.steps) / 1e9
            e2e["frac_of_host_write_peak"] = e2e["achieved_host_write_gbs"] / host_gbs
            e2e["bound"] = ("host DRAM writes: the COO vectors (host_result_bytes_per_step) are written into the caller's "
                            "buffers by the host cores from the compact arenas; host_write_gbs is the same pool filling "
                            "the same buffer with streaming stores")

    # ---------------- small-d configs of BASELINE.json (cfg1-4), N = 1 only ----------------
    small_cfgs = None
    if rank == 0 and world == 1 and not args.no_small and args.workload == "cfg5" and args.knots is None:
        small_cfgs = {}
        try:
            pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", 6650.0)
        except Exception:
            pk = 6650.0
        flushbuf = torch.empty(64 * 1024 * 1024, dtype=torch.float64, device=dev)
        for nm in ("cfg1", "cfg2", "cfg3", "cfg3_ket", "cfg4"):
            try:
                ws = workloads.make(nm)
                Ls = ws.layout
                evs = Evaluator(Ls, ws.G_drift, ws.G_drives, device=local)
                sZ, smu = torch.from_numpy(ws.Z).to(dev), torch.from_numpy(ws.mu).to(dev)
                sd = torch.empty(evs.defect_dim, dtype=torch.float64, device=dev)
                sj = torch.empty(evs.jacobian_nnz, dtype=torch.float64, device=dev)
                sh = torch.empty(evs.hessian_nnz, dtype=torch.float64, device=dev)
                sst = torch.cuda.ExternalStream(evs.stream, device=dev)
                for _ in range(3):
                    evs.eval_all_device(sZ, smu, sd, sj, sh)
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                tot = 0.0
                reps = 10
                for _ in range(reps):
                    flushbuf.add_(1.0)
                    torch.cuda.synchronize()
                    a.record(sst)
                    evs.eval_all_device(sZ, smu, sd, sj, sh)
                    b.record(sst); b.synchronize()
                    tot += a.elapsed_time(b)
                ms = tot / reps
                un = Ls.S * Ls.K
                gbs = algorithmic_bytes_per_unit(Ls) * un / (ms * 1e-3) / 1e9
                small_cfgs[nm] = {"value": un / (ms * 1e-3), "unit": "knot-evals/s", "ms_per_step": ms, "units": un,
                                  "d": Ls.d, "c": Ls.c, "m": Ls.m, "S": Ls.S,
                                  "roofline": {"bound": "hbm", "achieved": gbs, "peak": pk, "unit": "GB/s", "frac": gbs / pk,
                                               "fp64_tflops_survey_convention": survey_flops_per_unit(Ls) * un / (ms * 1e-3) / 1e12},
                                  "timing": "CUDA events on the library's stream around one device-resident call (launch, status "
                                            "read-back and synchronisation included), L2 flushed between calls"}
                evs.close()
            except Exception as exc:
                small_cfgs[nm] = {"value": None, "note": f"{type(exc).__name__}: {exc}"[:200]}
        del flushbuf

    # ---------------- CPU baseline (rank 0, N = 1 only) ----------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cores = os.cpu_count() or 1
        sample = cpu_sample_units(L, cores, args.ref_units)
        rate, n_units, native, blas_desc = cpu_port_throughput(w, sample, cores)
        cpu = {"value": rate, "unit": "knot-evals/s", "cores": cores, "kind": "port",
               "gemm": blas_desc or "the port's own loops", "gflops": rate * port_flops_per_unit(L) / 1e9,
               "sample": f"first {n_units} of {units_total} units, one pass, OpenMP over units, "
                         f"{'-march=native' if native else '-march=x86-64-v3'}"}
        try:    # the same port at one thread, and with its own (un-blocked) loops instead of the BLAS
            r1, n1, _, _ = cpu_port_throughput(w, max(1, n_units // cores), 1)
            cpu["one_thread"] = {"value": r1, "units": n1, "gflops": r1 * port_flops_per_unit(L) / 1e9}
            if blas_desc:
                r2, n2, _, _ = cpu_port_throughput(w, sample, cores, blas=False)
                cpu["plain_loops"] = {"value": r2, "units": n2, "note": "round-1 baseline: column-axpy products, no BLAS"}
        except Exception as exc:
            cpu["one_thread"] = {"value": None, "note": str(exc)[:200]}

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        per_rank_units = L.S * (k1 - k0)
        if prof["launches"] > 0 and L.d >= 25:
            gemm_ms = prof["ms"] / prof["launches"]
            achieved = prof["flops"] / (prof["ms"] * 1e-3) / 1e12
            traffic, traffic_src = None, None
            try:    # DRAM bytes per tensor flop of this kernel from the committed ncu --set full capture
                tj = json.load(open(os.path.join(ROOT, "profiles", "r02_traffic_bgemm_iso.json")))
                traffic = tj["dram_bytes_per_flop_per_drive_launches"] * prof["flops"] / prof["launches"]
                traffic_src = "profiles/r02_traffic_bgemm_iso.json: (dram__bytes_read.sum + dram__bytes_write.sum) per tensor flop x flops of the average launch"
            except Exception:
                pass
            roofline = {
                "bound": "tensor", "kernel": "bgemm_iso_kernel / bgemm_kernel (batched FP64 DMMA GEMM, mma.sync m16n8k4 -> SASS DMMA.8x8x4)",
                "achieved": achieved, "peak": dgemm_tflops, "unit": "TFLOP/s", "frac": achieved / dgemm_tflops,
                "peak_source": "cuBLAS DGEMM 8192^3 timed live in this run (MEASURED_PEAKS.json has no FP64 entry; "
                               "raw DMMA issue peak measured 37.1 TFLOP/s, scratch/dmma_probe.cu)",
                "traffic": traffic, "traffic_source": traffic_src,
                "algorithmic_bytes_per_launch": 8.0 * (prof["flops"] / prof["launches"]) / (2.0 * L.d) * 2.5,
                "launches_per_step": prof["launches"] / args.steps, "avg_launch_ms": gemm_ms,
                "kernel_share_of_step": prof["ms"] / args.steps / ms_prof if world == 1 else None,
                "measured_in": f"{args.steps} profiled steps right after the timed region ({ms_prof:.2f} ms/step with every "
                               "kernel serialised on one stream and an event pair per GEMM launch)",
                "flops_per_unit_executed": prof["flops"] / args.steps / per_rank_units,
                "flops_per_unit_survey_convention": survey_flops_per_unit(L),
                "hbm_algorithmic_gbs": algorithmic_bytes_per_unit(L) * units_total / (ms_dev * 1e-3) / 1e9,
            }
        else:
            gbs = algorithmic_bytes_per_unit(L) * units_total / (ms_dev * 1e-3) / 1e9
            pk = peaks.get("hbm_gbs", 6650.0)
            roofline = {"bound": "hbm", "kernel": "small_unit_kernel (fused per-unit evaluator)", "achieved": gbs, "peak": pk,
                        "unit": "GB/s", "frac": gbs / pk,
                        "peak_source": "MEASURED_PEAKS.json hbm_gbs" if "hbm_gbs" in peaks else "fallback 6.65 TB/s",
                        "traffic": None, "bytes_per_unit": algorithmic_bytes_per_unit(L)}
        slab_local = None
        if plane is not None:
            # N > 1: the headline is the Ipopt-rank-complete number (broadcast and delivery inside the step);
            # the collective-free slab evaluation explains it
            slab_local = {"value": value, "unit": "knot-evals/s", "ms_per_step": ms_dev,
                          "what": "every rank evaluates its slab with inputs resident, results left on its own GPU "
                                  "(CUDA events, max over ranks): no data-path exchange"}
            value, ms_dev = plane["value"], plane["ms_per_step"]
        line = {
            "metric": "knot-pt evals/sec (defect+Jac+Hess)", "value": value, "unit": "knot-evals/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_dev, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"{args.workload}: {WORKLOAD_DESC[args.workload]}", "units": units_total, "d": L.d,
                       "c": L.c, "m": L.m, "N": L.N, "S": L.S,
                       "jac_structure": "block-diagonal" if L.jac_mode else "dense-x",
                       "parallelism": (f"time-slab x{world}; Z, mu broadcast from rank 0 (NCCL), value arenas delivered to rank 0 over "
                                       f"NVLink by the producing kernels, COO assembly on rank 0") if world > 1 else "single GPU",
                       "l2": "inputs+outputs per step exceed L2 (no flush needed)" if flush is None
                             else "L2 flushed between timed steps (512 MB read-modify-write)"},
            "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu,
            "callback_mix": callback_mix, "data_plane": plane, "slab_local": slab_local, "peer_gather": peer,
            "small_configs": small_cfgs,
        }
        if args.dt_scale != 1.0:
            line["config"]["dt_scale"] = args.dt_scale
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
