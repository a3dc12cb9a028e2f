/* qcb200.h -- C ABI of the B200-native collocation evaluator.
 *
 * Drop-in for the evaluator behind QuantumCollocation.jl's bilinear integrators
 * (reference seam: the `integrator=` keyword of the problem templates,
 * src/problem_templates/smooth_pulse_problem.jl:57,103-127; the objects built by
 * src/quantum_integrators.jl:28-134).  Julia reaches these entry points with `ccall`
 * (see INTEGRATION.md and julia/QCB200.jl); Python reaches them with ctypes.
 *
 * Conventions
 *  - every matrix is column-major FP64; every index array handed across the ABI is 1-based
 *    int64 (what Julia's Int arrays hold);
 *  - the caller owns all host arrays; the library owns device memory, streams and pinned
 *    staging; no host pointer is retained after a call returns; no callbacks;
 *  - every function returns a qcb_status; the text of the last error on a handle is
 *    available through qcb_last_error();
 *  - a handle is not thread-safe (Ipopt's callbacks are sequential on one host thread,
 *    src/quantum_control_problem.jl:163-164);
 *  - there is no CPU fallback: if no CUDA device is usable, qcb_create fails.
 */
#ifndef QCB200_H
#define QCB200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct qcb_handle_s *qcb_handle;
typedef struct qcb_fidelity_s *qcb_fidelity;

typedef enum {
    QCB_OK = 0,
    QCB_ERR_ARG = 1,        /* bad argument / descriptor */
    QCB_ERR_CUDA = 2,       /* CUDA runtime error (message has the detail) */
    QCB_ERR_RANGE = 3,      /* non-finite input or ||dt*G(u)||_1 beyond the supported range:
                               map to Ipopt's "evaluation failed" in the shim */
    QCB_ERR_STRUCTURE = 4,  /* a caller-supplied sparsity structure could not be mapped */
    QCB_ERR_NOMEM = 5
} qcb_status;

/* x-block layout of the defect Jacobian */
#define QCB_JAC_DENSE_X 0      /* d(delta)/d(x_k) as a dense x_dim^2 block (explicit zeros) */
#define QCB_JAC_BLOCK_DIAG 1   /* c diagonal d x d blocks */

/* What to evaluate */
#define QCB_WANT_DEFECT 0
#define QCB_WANT_JACOBIAN 1
#define QCB_WANT_HESSIAN 2

/* Problem descriptor.  Replaces the arguments of
 *   DirectTrajOpt.BilinearIntegrator(G_hat, x_sym, u_sym, traj)
 * as called at src/quantum_integrators.jl:32,44,55,71,112,123,133: the closure G_hat(u) =
 * G_drift + sum_j u_j G_j is passed as its matrices, the symbols as offsets into a knot. */
typedef struct {
    int32_t struct_bytes;   /* sizeof(qcb_desc), for ABI evolution */
    int32_t d;              /* rows of the generator (2n for kets / vectorised unitaries) */
    int32_t c;              /* columns of the state panel: n for I_n (x) G (unitary), 1 for a ket */
    int32_t m;              /* number of drives */
    int32_t N;              /* knots of the trajectory; K = N-1 intervals */
    int32_t S;              /* systems sharing the controls (SamplingTrajectory), else 1 */
    int32_t z_dim;          /* doubles per knot in Z */
    int32_t u_off;          /* 0-based offset of u inside a knot */
    int32_t dt_off;         /* 0-based offset of the timestep inside a knot */
    int32_t jac_mode;       /* QCB_JAC_* */
    int32_t device;         /* CUDA device ordinal */
    int32_t k_begin;        /* this handle evaluates intervals [k_begin, k_end), 0-based; */
    int32_t k_end;          /*   k_end = 0 means K (the multi-GPU time-slab partition)      */
    int32_t force_path;     /* 0 auto, 1 small-d kernel, 2 large-d DMMA path (testing) */
    const int32_t *x_off;   /* [S] 0-based offset of each system's state inside a knot */
    const double *G_drift;  /* [S][d*d]    */
    const double *G_drives; /* [S][m][d*d] */
    /* Optional state index map (both NULL: the state is the d*c contiguous entries at x_off[s]).
     * For integrators whose state is a concatenation of several named components --
     * VariationalKetIntegrator / VariationalUnitaryIntegrator, src/quantum_integrators.jl:140-169:
     * x = vcat(state, variations...) under the block generator var_G(G, G_vars).  The kernels work
     * on the d x c panel X (column-major, entry i = row + col*d); the maps say where each panel entry
     * lives on the caller's side:
     *   x_index   [S][d*c]  0-based offset inside a knot of panel entry i of system s (x_off is ignored);
     *   row_index [d*c]     0-based position of panel entry i among the unit's defect rows / multipliers
     *                       (NULL: identity).  For the unitary variant, whose vector is
     *                       [vec(U); vec(U_v1); ...] while the panel stacks the blocks row-wise, this is
     *                       the permutation (block b, row r, column q) -> b*d0*c + q*d0 + r.
     * A caller compiled against the older struct (struct_bytes up to G_drives) is still accepted. */
    const int32_t *x_index;
    const int32_t *row_index;
} qcb_desc;

/* ---- life cycle -------------------------------------------------------------------- */
qcb_status qcb_create(qcb_handle *out, const qcb_desc *desc);
void qcb_destroy(qcb_handle h);
const char *qcb_last_error(qcb_handle h);   /* h may be NULL: error of the last failed qcb_create */
const char *qcb_version(void);

/* ---- sizes ---------------------------------------------------------------------------
 * integrator.dim in the reference = x_dim*(N-1) per integrator
 * (src/problem_templates/smooth_pulse_problem.jl:513); here summed over the S systems and
 * restricted to the handle's slab. */
int64_t qcb_defect_dim(qcb_handle h);
int64_t qcb_jacobian_nnz(qcb_handle h);
int64_t qcb_hessian_nnz(qcb_handle h);

/* ---- sparsity structure ---------------------------------------------------------------
 * Native structure, order [system][interval][entries sorted by (column,row)].  rows of the
 * Jacobian are 1-based positions in this handle's defect vector plus row_base; columns are
 * 1-based positions in Z.  Replaces the jacobian_structure / hessian_structure queries the
 * MOI evaluator of DirectTrajOpt makes of an AbstractIntegrator. */
qcb_status qcb_jacobian_structure(qcb_handle h, int64_t row_base, int64_t *rows, int64_t *cols);
qcb_status qcb_hessian_structure(qcb_handle h, int64_t *rows, int64_t *cols);
/* Adopt the reference's own COO order instead: pass the (rows, cols) the Julia integrator
 * reports for its Jacobian block (rows relative to row_base) and its Hessian block.  Entries
 * must be grouped [system][interval] with the same pattern in every interval; entries the
 * model makes structurally zero are accepted and filled with 0.  Either pair may be NULL
 * (nnz 0) to keep the native order for it. */
qcb_status qcb_set_structure(qcb_handle h, int64_t row_base,
                             const int64_t *jac_rows, const int64_t *jac_cols, int64_t jac_nnz,
                             const int64_t *hess_rows, const int64_t *hess_cols, int64_t hess_nnz);

/* ---- evaluation, host buffers (what the Julia shim calls from Ipopt's callbacks) -------
 * Z: N*z_dim doubles.  mu: multipliers of this integrator set, [S][K][x_dim] with K = N-1
 * (full-trajectory indexing even for a slab handle).  Outputs cover the handle's slab.
 *   qcb_eval_defect   <- DirectTrajOpt.evaluate!(delta, integrator, traj)
 *                        (src/problem_templates/smooth_pulse_problem.jl:512-515)
 *   qcb_eval_jacobian <- the integrator's Jacobian-values callback (eval_jac_g)
 *   qcb_eval_hessian  <- the integrator's Hessian-of-Lagrangian-values callback (eval_h)
 *   qcb_eval_all      <- all three from one pass over the trajectory
 * Any output pointer of qcb_eval_all may be NULL. */
qcb_status qcb_eval_defect(qcb_handle h, const double *Z, double *delta);
qcb_status qcb_eval_jacobian(qcb_handle h, const double *Z, double *jac_vals);
qcb_status qcb_eval_hessian(qcb_handle h, const double *Z, const double *mu, double *hess_vals);
qcb_status qcb_eval_all(qcb_handle h, const double *Z, const double *mu,
                        double *delta, double *jac_vals, double *hess_vals);

/* The same four calls with Ipopt's `new_x` flag -- the argument Ipopt passes to eval_g / eval_jac_g / eval_h
 * (0: "x is the point of the previous callback").  new_x != 0: identical to the calls above.  new_x == 0: the
 * caller promises that Z holds what the previous host-buffer call on this handle was given; Z is then not
 * uploaded again and whatever depends on Z alone (propagators, forward tangents, the Jacobian arenas) is taken
 * from the handle's scratch, so that eval_g -> eval_jac_g -> eval_h on one iterate cost one pass; mu may
 * differ (new_lambda).  Results are bit-identical to the recomputing calls.  Reuse needs the large-d path with
 * the whole slab inside the scratch budget; otherwise the flag is ignored and the call recomputes. */
qcb_status qcb_eval_defect_x(qcb_handle h, const double *Z, int32_t new_x, double *delta);
qcb_status qcb_eval_jacobian_x(qcb_handle h, const double *Z, int32_t new_x, double *jac_vals);
qcb_status qcb_eval_hessian_x(qcb_handle h, const double *Z, int32_t new_x, const double *mu, double *hess_vals);
qcb_status qcb_eval_all_x(qcb_handle h, const double *Z, const double *mu, int32_t new_x,
                          double *delta, double *jac_vals, double *hess_vals);

/* ---- evaluation, device buffers (inputs/outputs already resident in HBM) ---------------
 * Same semantics; all pointers are device pointers on the handle's device; the work is
 * ordered on the handle's OWN stream (qcb_stream) and the call returns after it has completed.
 * Stream ordering is the caller's job on the way in: every input and output buffer must be idle
 * when the call is made -- synchronise the stream that produced the inputs (or last touched the
 * outputs) first.  The library's stream is non-blocking and does not wait for the legacy default
 * stream or for any other stream.  The same holds for qcb_fidelity_eval_batch_device_noalloc and
 * for device pointers handed to qcb_ipc_* peers. */
qcb_status qcb_eval_all_device(qcb_handle h, const double *dZ, const double *dmu,
                               double *d_delta, double *d_jac_vals, double *d_hess_vals);

/* The same call with Ipopt's `new_x` flag (Ipopt passes it to eval_g / eval_jac_g / eval_h: 0 means "x is
 * the point of the previous callback").  new_x != 0: identical to qcb_eval_all_device.  new_x == 0: the
 * caller promises that dZ holds the same values as in the previous call on this handle; what depends on
 * Z alone (the propagators, their forward tangents, the Jacobian arena) is then taken from the handle's
 * scratch instead of being recomputed, so that eval_g -> eval_jac_g -> eval_h on one iterate cost one
 * pass; dmu may differ (Ipopt's new_lambda).  Results are bit-identical to the recomputing call.  Only
 * whole-slab scratch keeps the intermediates (large-d path, slab within the scratch budget); otherwise,
 * and on the small-d path, the call silently recomputes. */
qcb_status qcb_eval_all_device_x(qcb_handle h, const double *dZ, const double *dmu, int32_t new_x,
                                 double *d_delta, double *d_jac_vals, double *d_hess_vals);

/* ---- rollout --------------------------------------------------------------
 * x_{k+1} = exp(dt_k G(u_k)) x_k for k = 1..N-1 with the controls and timesteps of Z held constant
 * over each interval -- the propagation behind sync_trajectory! / rollout fidelity
 * (src/quantum_control_problem.jl:137-150; piccolo_options.jl:30 `rollout_integrator = expv`).
 * x_init: [S][x_dim] initial states, or NULL to start from the states stored at knot 1 of Z.
 * states: [S][N][x_dim] (host).  All propagators are computed in parallel by the evaluator's
 * kernels; only the chain of N-1 panel products is sequential.  The handle must own the whole
 * trajectory (k_begin = 0, k_end = K). */
qcb_status qcb_rollout(qcb_handle h, const double *Z, const double *x_init, double *states);

/* Peer delivery for the multi-GPU time-slab partition (no counterpart in the
 * single-process reference).  The rank that hosts Ipopt allocates the full value buffers with
 * qcb_device_alloc and exports them (64-byte CUDA IPC handles, to be sent to the other ranks by any
 * means); every other rank maps them with qcb_ipc_open and passes `mapped + slab offset` as the
 * output pointers of qcb_eval_all_device: its kernels then store the slab straight into the host
 * rank's HBM over NVLink while they compute -- the gather is fused into the producing kernels. */
/* Compact delivery (the faster form of the same gather): a Jacobian of the unitary path repeats -E once
 * per state column, so the COO vectors are ~5x larger than the distinct numbers behind them.
 * qcb_eval_arenas_device evaluates the slab like qcb_eval_all_device but stores, per unit, the value
 * arenas (qcb_arena_doubles(h, 0) doubles for the Jacobian, (h, 1) for the Hessian; unit u of the slab at
 * offset u * that size) -- into the Ipopt rank's memory when the pointers are IPC mappings; delta is
 * delivered as before.  qcb_assemble_device then runs on the Ipopt rank: it writes the COO value vectors
 * of n_units units, laid out [system][nk_src intervals] as their source slab produced them, from such
 * arenas (any handle of the same problem holds the per-unit tables).  Large-d path only. */
int64_t qcb_arena_doubles(qcb_handle h, int32_t which);
qcb_status qcb_eval_arenas_device(qcb_handle h, const double *dZ, const double *dmu, double *d_delta,
                                  double *d_jac_arenas, double *d_hess_arenas);
qcb_status qcb_assemble_device(qcb_handle h, const double *d_jac_arenas, const double *d_hess_arenas,
                               int64_t n_units, int32_t nk_src, double *d_jac_vals, double *d_hess_vals,
                               int32_t parts);
/* Two-part delivery of the Jacobian arena.  The arena starts with the propagator E; on the unitary path
 * 90 % of the Jacobian's entries (the x-block, c copies of -E) come from that part alone, and E is known
 * long before the tangents.  qcb_set_arena_parts(h, 1): the next qcb_eval_arenas_device[_x] computes the
 * defect and the propagators only and delivers the E part; (h, 2): it computes the tangents and delivers
 * the rest of the Jacobian arena; (h, 3), the default: everything in one call.  qcb_assemble_device takes
 * the same `parts`: 1 = the entries copied from the E part, 2 = all other entries, 3 = all. */
qcb_status qcb_set_arena_parts(qcb_handle h, int32_t parts);
qcb_status qcb_eval_arenas_device_x(qcb_handle h, const double *dZ, const double *dmu, int32_t new_x,
                                    double *d_delta, double *d_jac_arenas, double *d_hess_arenas);
/* Asynchronous mode for the device-pointer calls (qcb_eval_all_device[_x], qcb_eval_arenas_device[_x],
 * qcb_assemble_device): with it on, a call returns as soon as its kernels are queued on the handle's
 * streams; qcb_wait(h, 0) blocks until everything queued so far has completed and returns the status the
 * synchronous call would have returned (QCB_ERR_RANGE for a non-finite input).  qcb_wait(h, 1) / (h, 2) /
 * (h, 3) block only until the last delivery of the E part / of the rest of the Jacobian arena / of the
 * Hessian arena by qcb_eval_arenas_device has left this GPU (deliveries run on a side stream behind the
 * kernels that produce the arenas).  This is what lets the multi-GPU
 * driver queue the Jacobian side and the Hessian side of one iterate back to back, so that the first
 * delivery crosses NVLink while the second half is still being computed, and lets the Ipopt rank assemble
 * foreign slabs while its own kernels run.  Buffers passed to a queued call must stay untouched until the
 * wait.  No counterpart in the reference. */
qcb_status qcb_set_async(qcb_handle h, int32_t on);
qcb_status qcb_wait(qcb_handle h, int32_t what);
qcb_status qcb_device_alloc(int32_t device, void **ptr, size_t bytes);
void qcb_device_free(int32_t device, void *ptr);
qcb_status qcb_ipc_export(int32_t device, void *ptr, unsigned char handle[64]);
qcb_status qcb_ipc_open(int32_t device, const unsigned char handle[64], void **ptr);
void qcb_ipc_close(int32_t device, void *ptr);
qcb_status qcb_device_to_host(int32_t device, void *dst_host, const void *src_device, size_t bytes);

/* Pinned host memory for callers that want full PCIe bandwidth on the host-buffer calls. */
qcb_status qcb_host_alloc(void **ptr, size_t bytes);
void qcb_host_free(void *ptr);
/* The same for memory the caller already owns (the value arrays Ipopt / MOI allocated, a Julia Vector):
 * page-locks [ptr, ptr + bytes) so that the host-buffer calls can DMA into it (on the large-d path the long
 * plain runs of the COO vectors then go straight to their place instead of through staging and the host
 * cores).  Register once per array, keep it alive and unregister before it is freed.  Already registered
 * memory is accepted.  No counterpart in the reference. */
qcb_status qcb_host_register(void *ptr, size_t bytes);
void qcb_host_unregister(void *ptr);

/* Host-side output formatting (no counterpart in the reference).  On the large-d path the COO
 * vectors are written into the caller's buffers by a pool of host threads from the compact arenas
 * the GPU sends.  qcb_host_threads: size of that pool -- QCB_HOST_THREADS, else OMP_NUM_THREADS
 * (ignored when it is 1 and LOCAL_WORLD_SIZE is set: torch.distributed.run's default), else (cores this
 * process may run on) / LOCAL_WORLD_SIZE; threads are pinned one per core
 * (QCB_HOST_PIN=0: not pinned).  qcb_host_stream_fill: writes `bytes` of dst with the same
 * streaming stores on the same pool and reports the wall time -- the measured host-write roofline
 * that bench.py prints beside the end-to-end number. */
int32_t qcb_host_threads(void);
/* Host-buffer calls on the large-d path that deliver the Jacobian are bound by the host: its cores' streaming
 * stores and the PCIe writes beside them share one memory system, and on some hosts fewer store threads give a
 * faster call.  Each handle therefore times its own calls: one settling call, two calls each with all pool
 * threads, 3/4, 1/2 and 3/8 of them, then the fastest stays (fewer threads only for a gain of at least 3 %); calls
 * with and without the Hessian are tuned separately.  QCB_HOST_TUNE=0 or an explicit QCB_HOST_THREADS turns
 * this off.  Returns the number of threads in use for that kind of call (the pool size until decided). */
int32_t qcb_host_threads_in_use(qcb_handle h, int32_t with_hessian);
qcb_status qcb_host_stream_fill(void *dst, size_t bytes, double *seconds);

/* Number of kernels this handle has launched since creation, and per-phase timings of the
 * last host-buffer call in milliseconds: [0] H2D, [1] kernels with the device->host copies and the
 * host-side expansion that overlap them, [2] always 0 (device->host is not separable: it is part
 * of [1]), [3] total. */
int64_t qcb_kernel_launches(qcb_handle h);
qcb_status qcb_get_timings(qcb_handle h, double ms[4]);
/* Bytes the last host-buffer call moved over PCIe: [0] host->device, [1] device->host.  On the
 * large-d path the device->host figure is the compact value arenas (every distinct number once);
 * the COO vectors are written from them by the host cores.  No counterpart in the reference. */
qcb_status qcb_get_transfer_bytes(qcb_handle h, int64_t bytes[2]);
/* The CUDA stream (cudaStream_t) the handle launches on, for callers that time with events. */
void *qcb_stream(qcb_handle h);
/* Per-launch CUDA-event timing of the dominant kernel of the large-d path (the batched FP64
 * DMMA GEMM).  After enabling, every evaluation accumulates into
 * out[0] = launches, out[1] = total milliseconds, out[2] = total algorithmic flops, out[3] = 0.
 * No counterpart in the reference (measurement only). */
qcb_status qcb_profile_enable(qcb_handle h, int32_t on);
qcb_status qcb_profile_read(qcb_handle h, double out[4]);

/* ---- terminal fidelities ---------------------------------------------------------------
 * F(x_N) for the losses of src/quantum_objectives.jl and the constraints of
 * src/quantum_constraints.jl.  goal: complex entries as interleaved (re, im) pairs,
 * column-major.  All of them are quadratic forms F = w * sum_r (g_r . x)^2. */
#define QCB_FID_UNITARY 0           /* unitary_fidelity_loss, quantum_objectives.jl:173-180 */
#define QCB_FID_KET 1               /* ket_fidelity_loss, :21-27 */
#define QCB_FID_COHERENT_KET 2      /* coherent_ket_fidelity, :88-103 (x = concatenated kets) */
#define QCB_FID_UNITARY_SUBSPACE 3  /* EmbeddedOperator variant, :182-191 */
#define QCB_FID_DENSITY_PURE 4      /* density_matrix_pure_state_infidelity_loss, :232-239: F = Re(psi' rho psi),
                                       LINEAR in x = [vec(Re rho); vec(Im rho)] (gradient constant, Hessian 0);
                                       goal = psi (n complex entries) */

/* n: Hilbert dimension; count: number of kets (COHERENT_KET) else 1; subspace: 0-based level
 * indices (UNITARY_SUBSPACE; goal is then n_sub x n_sub), NULL otherwise. */
qcb_status qcb_fidelity_create(qcb_fidelity *out, int32_t kind, int32_t n, int32_t count,
                               const double *goal_re_im, const int32_t *subspace, int32_t n_sub,
                               int32_t device);
void qcb_fidelity_destroy(qcb_fidelity f);
int64_t qcb_fidelity_x_dim(qcb_fidelity f);
/* x: x_dim doubles (host).  F, grad (x_dim) and hess (x_dim*x_dim, dense column-major) may be
 * NULL.  The infidelity objective Q*|1-F| (quantum_objectives.jl:199) and the constraint
 * F_min - F (quantum_constraints.jl:112) follow by the chain rule in the shim. */
qcb_status qcb_fidelity_eval(qcb_fidelity f, const double *x, double *F, double *grad, double *hess);
/* The Hessian in factored form: Hess F = 2 w G' G with G [rank][x_dim] row-major, constant in x (rank 2 for
 * unitary / ket / coherent-ket, 2 + 2 n_sub^2 for the subspace kind, 0 for the linear density kind).  For
 * large states the dense x_dim^2 block of qcb_fidelity_eval is impractical (cfg5: 537 MB); a caller that
 * needs Hessian-vector products or a few entries takes the factor instead.  G: rank * x_dim doubles (host). */
int64_t qcb_fidelity_rank(qcb_fidelity f);
qcb_status qcb_fidelity_factor(qcb_fidelity f, double *G, double *w);

/* Batched fidelity that never allocates.  It has no counterpart in the reference.
 *
 * For callers that must not (or cannot) obtain device memory while they run: a host program whose
 * GPU memory is already fully claimed by another allocator (PyTorch's or CUDA.jl's caching pool),
 * so that a cudaMalloc inside this library would fail; and latency-sensitive loops (the per-iteration
 * objective of a SamplingProblem, sampling_problem.jl:139-142) that cannot afford the device-wide
 * synchronisation cudaMalloc / cudaFree imply.  This call uses no cudaMalloc, cudaFree or any
 * allocating library routine: its only scratch -- the projections g_r . x_b -- is a fixed
 * 512 KiB array compiled into the library (resident on every GPU that loads it, until exit).
 *
 *   d_X    [count][x_dim]  terminal states, device memory (caller's)
 *   d_F    [count]         fidelities, device memory (caller's), may be NULL
 *   d_grad [count][x_dim]  gradients dF/dx, device memory (caller's), may be NULL
 *
 * Limit: count * R <= 65536 and count <= 65535, where R is the number of linear forms of the
 * fidelity kind (2 for unitary / ket / coherent-ket, 2 + 2 n_sub^2 for the subspace kind); larger
 * batches are rejected with QCB_ERR_ARG before anything is launched.  The first call on a device
 * also looks up the device address of the fixed array; make that first call before any stream
 * capture is begun.  All calls in the process share the one array and it keeps the previous
 * call's contents (each call overwrites every slot it later reads, so nothing needs clearing):
 * two calls whose kernels can overlap on the GPU (different handles, streams or host threads)
 * race with each other -- serialise them.  Returns after the work has completed. */
qcb_status qcb_fidelity_eval_batch_device_noalloc(qcb_fidelity f, const double *d_X, int64_t count,
                                                  double *d_F, double *d_grad);

/* ---- free-phase unitary fidelity -------------------------------------------
 * UnitaryFreePhaseInfidelityObjective (src/quantum_objectives.jl:203-226) and the theta-parametrised
 * FinalUnitaryFidelityConstraint (src/quantum_constraints.jl:123-144): the loss acts on z = [x_N; theta]
 * with theta a set of global variables and compares U with U_goal(theta).  The reference takes U_goal as
 * a Julia function; across a C ABI the family is the one free-phase problems use, a diagonal phase
 * factor:  U_goal(theta) = D(theta) U_goal (side 0) or U_goal D(theta) (side 1),
 * D = diag(exp(i phi)), phi = P theta, with P the n x n_theta matrix `phase_map` (column-major: the
 * diagonal of each commuting phase generator, e.g. of Z on one qubit).  Any other U_goal(theta) stays on
 * the Julia path.
 *   F(x, theta) = |tr(U_goal(theta)' U)|^2 / n^2
 * x: 2 n^2 doubles (iso-vec of U), theta: n_theta doubles (host).  grad: x_dim + n_theta entries over
 * z = [x; theta]; hess: dense (x_dim + n_theta)^2, column-major; F, grad, hess may be NULL.  The
 * objective Q |1 - F| and the constraint F_min - F follow by the chain rule in the shim. */
typedef struct qcb_free_phase_s *qcb_free_phase;
qcb_status qcb_free_phase_create(qcb_free_phase *out, int32_t n, const double *goal_re_im, int32_t n_theta,
                                 const double *phase_map, int32_t side, int32_t device);
void qcb_free_phase_destroy(qcb_free_phase f);
qcb_status qcb_free_phase_eval(qcb_free_phase f, const double *x, const double *theta, double *F, double *grad,
                               double *hess);

/* ---- auxiliary knot-point terms ------------------------------------------
 * The cheap terms every problem template adds around the bilinear integrator:
 *   derivative integrators  delta_k = x_{k+1} - x_k - dt_k * dx_k   (DirectTrajOpt.DerivativeIntegrator as
 *       used at src/problem_templates/smooth_pulse_problem.jl:150-152; "du[k] = (u[k+1] - u[k]) / dt",
 *       spline_pulse_problem.jl:176), 4 Jacobian entries and 1 Hessian entry per row;
 *   leakage objective  sum_t Q_t * sum_{i in idx} x_t[i]^2 / |idx|   (src/quantum_objectives.jl:291-307);
 *   leakage constraint x_t[i]^2 - value <= 0                          (src/quantum_constraints.jl:156-172).
 * Orders: derivative rows [k][integrator][component]; Jacobian values per row in the order
 * (x_k, dx_k, dt_k, x_{k+1}); leakage arrays [t][i] over all N knots. */
typedef struct qcb_aux_s *qcb_aux;
typedef struct {
    int32_t struct_bytes;       /* sizeof(qcb_aux_desc) */
    int32_t N, z_dim, dt_off, device;
    int32_t n_deriv;            /* number of derivative integrators */
    int32_t leak_x_off;         /* 0-based offset of the state the leakage indices refer to */
    int32_t n_leak;             /* number of leakage indices (0: none) */
    const int32_t *deriv_x_off;  /* [n_deriv] offset of x  (e.g. u,  du)  */
    const int32_t *deriv_dx_off; /* [n_deriv] offset of dx (e.g. du, ddu) */
    const int32_t *deriv_dim;    /* [n_deriv] */
    const int32_t *leak_idx;     /* [n_leak] 0-based indices inside the state component */
} qcb_aux_desc;
qcb_status qcb_aux_create(qcb_aux *out, const qcb_aux_desc *desc);
void qcb_aux_destroy(qcb_aux a);
int64_t qcb_aux_derivative_dim(qcb_aux a);      /* (N-1) * sum of dims */
int64_t qcb_aux_leakage_dim(qcb_aux a);         /* N * n_leak */
/* 1-based COO structure of the derivative rows: 4 Jacobian entries per row, 1 Hessian entry per row
 * (upper triangle).  Any pointer pair may be NULL. */
qcb_status qcb_aux_derivative_structure(qcb_aux a, int64_t row_base, int64_t *jac_rows, int64_t *jac_cols,
                                        int64_t *hess_rows, int64_t *hess_cols);
/* Host buffers; outputs may be NULL.  mu: multipliers of the derivative rows. */
qcb_status qcb_aux_derivative_eval(qcb_aux a, const double *Z, const double *mu, double *delta,
                                   double *jac_vals, double *hess_vals);
/* Qs: [N] knot weights (NULL = 1).  objective: scalar; grad_vals / con_vals / con_jac_vals /
 * con_hess_vals: [N][n_leak] (gradient entries dJ/dx_t[idx_i]; constraint values; their derivatives
 * 2 x; Hessian-of-Lagrangian diagonal 2 mu).  The objective's own Hessian is the constant diagonal
 * 2 Q_t / n_leak. */
qcb_status qcb_aux_leakage_eval(qcb_aux a, const double *Z, const double *Qs, double value, const double *mu,
                                double *objective, double *grad_vals, double *con_vals, double *con_jac_vals,
                                double *con_hess_vals);

/* Objective terms around the integrator: quadratic regularisers and the minimum-time
 * term.  Replaces DirectTrajOpt.QuadraticRegularizer (call sites smooth_pulse_problem.jl:139-152) and
 * MinimumTimeObjective (minimum_time_problem.jl:84-131):
 *     J = sum_terms 1/2 sum_{t=1..N} sum_i R_i (dt_t v_{t,i})^2  +  D sum_{t=1..N-1} dt_t
 * This form reproduces the objective values printed in the reference's solver logs to 1e-14
 * (docs/literate/examples/multilevel_transmon.jl:139, two_qubit_gates.jl:261) on the solved
 * trajectories the reference ships (docs/data), see tests/test_reference_fixtures.py.
 * Terms: component offsets off[i] / sizes dim[i] inside a knot, weights R concatenated; rows = sum of
 * the sizes.  Outputs are compact, in row order, for the caller to scatter:
 * grad_vals [N * (rows + 1)]: per knot dJ/dv for every regularised entry, then dJ/d(dt) (which
 * includes D for the first N-1 knots); their 1-based columns of Z come from
 * qcb_aux_regularizer_structure (grad_cols).  hess_vals [N * (2 rows + 1)]: per knot {(v,v), (v,dt)}
 * for every regularised entry, then (dt,dt); 1-based upper-triangle COO in hess_rows / hess_cols.
 * A component named by two terms yields two entries (Ipopt sums duplicates).  Any output may be NULL. */
qcb_status qcb_aux_regularizer_structure(qcb_aux a, int32_t n_terms, const int32_t *off, const int32_t *dim,
                                         int64_t *grad_cols, int64_t *hess_rows, int64_t *hess_cols);
qcb_status qcb_aux_regularizer_eval(qcb_aux a, const double *Z, int32_t n_terms, const int32_t *off,
                                    const int32_t *dim, const double *R, double min_time_D, double *objective,
                                    double *grad_vals, double *hess_vals);
/* TimeStepsAllEqualConstraint (applied by every template, "timesteps all equal constraint" in the
 * logs; 2 nnz per row = the 98 / 198 of SURVEY.md 8a-a5): rows t = 1..N-1: dt_t - dt_N = 0.
 * delta [N-1]; Jacobian COO (1-based) and values (+1, -1) [2 (N-1)].  Any output may be NULL. */
qcb_status qcb_aux_timesteps_equal(qcb_aux a, const double *Z, int64_t row_base, double *delta, int64_t *jac_rows,
                                   int64_t *jac_cols, double *jac_vals);

/* UnitarySensitivityObjective (src/quantum_objectives.jl:255-279):
 *     J = sum_{t in times} Q_t * scale^4 * |tr(U_t' U_t)|^2 / n^2,   U_t = iso_vec_to_operator(x_t),
 * on a (variational) state component of x_dim = 2 n^2 entries at offset x_off of a knot; times are
 * 1-based knot indices.  tr(U'U) = ||x_t||^2, so the term is a quartic in x_t.  Outputs (host, any may
 * be NULL): objective; grad_vals [n_times][x_dim]; hess_vals [n_times][x_dim][x_dim] dense. */
qcb_status qcb_aux_sensitivity_eval(qcb_aux a, const double *Z, int32_t x_off, int32_t x_dim, int32_t n,
                                    int32_t n_times, const int32_t *times, const double *Qs, double scale,
                                    double *objective, double *grad_vals, double *hess_vals);

/* ---- one call per Ipopt callback (second half) ------------------------------
 * The reference solves a QuantumControlProblem by handing Ipopt the five callbacks DirectTrajOpt's
 * MOI evaluator builds over the lists a problem template assembles (src/quantum_control_problem.jl:163-164;
 * src/problem_templates/smooth_pulse_problem.jl:139-165: objective = terminal infidelity +
 * QuadraticRegularizers on u, du, ddu; integrators = [BilinearIntegrator, DerivativeIntegrator(u, du),
 * DerivativeIntegrator(du, ddu)]; _problem_templates.jl:38-67 adds the timesteps-all-equal rows).
 * A qcb_problem is that list on this side of the ABI.  It BORROWS its pieces (the caller keeps and
 * destroys them, after the problem): integrator handles that own the whole trajectory, one qcb_aux
 * (derivative rows, dt_t - dt_N rows, regularisers, minimum-time term) and one qcb_fidelity for
 *     f(Z) = sum_s w_s Q |1 - F(x_N^s)| + sum_terms 1/2 sum_t sum_i R_i (dt_t v_{t,i})^2 + D sum_{t<N} dt_t
 * (one terminal state, or one per system of a sampling problem; all states are evaluated by one batched call).
 * Variables: Z, n = N * z_dim.  Constraint rows, in this order: every integrator's defect rows, the
 * derivative rows, the N-1 timestep rows, the leakage rows (m in total).  Each callback below is ONE call that writes
 * the caller's full-length array; new_x is Ipopt's flag, passed through unchanged: the problem
 * remembers which integrators have not yet seen the current iterate, so that
 * eval_f(new_x = 1) -> eval_g(0) -> eval_jac_g(0) -> eval_h(0) uploads Z once per integrator and reuses
 * the propagators and tangents between the three constraint callbacks (qcb_eval_*_x).
 * Free-phase and sensitivity objectives stay separate calls. */
typedef struct qcb_problem_s *qcb_problem;
typedef struct {
    int32_t struct_bytes;           /* sizeof(qcb_problem_desc) */
    int32_t N;                      /* knots */
    int32_t z_dim;                  /* doubles per knot */
    int32_t n_integrators;
    int32_t use_derivative_rows;    /* 1: aux's derivative integrators are constraint rows */
    int32_t use_dt_equal_rows;      /* 1: append the rows dt_t - dt_N = 0 */
    int32_t fid_hessian;            /* 1: dense upper-triangle Hessian block of the terminal term on x_N
                                       (x_dim <= 8192); 0: left out (limited-memory runs, large states) */
    int32_t n_reg;                  /* regulariser terms */
    int32_t use_leakage_rows;       /* 1: append the N * n_leak inequality rows x_t[i]^2 - leak_value <= 0 of aux's
                                       leakage indices (LeakageConstraint, src/quantum_constraints.jl:156-172); the
                                       caller gives Ipopt g_U = 0, g_L = -inf for them */
    int32_t use_leakage_objective;  /* 1: add leak_Q * sum_t sum_i x_t[i]^2 / n_leak (LeakageObjective,
                                       src/quantum_objectives.jl:291-307) */
    int32_t n_fid_states;           /* terminal states the infidelity is summed over (0 or 1: one; S for the systems
                                       of a sampling problem, sampling_problem.jl:139-142) */
    int32_t reserved0;              /* 0 */
    double Q;                       /* weight of the terminal infidelity */
    double min_time_D;              /* weight of sum dt (0: none) */
    double leak_Q;                  /* weight of the leakage objective */
    double leak_value;              /* bound of the leakage rows */
    const qcb_handle *integrators;  /* [n_integrators] */
    qcb_aux aux;                    /* may be NULL when nothing of it is used */
    qcb_fidelity fidelity;          /* may be NULL: no terminal term */
    const int32_t *fid_index;       /* [n_fid_states][x_dim of the fidelity] 0-based positions inside the LAST knot */
    const double *fid_weights;      /* [n_fid_states] weights of the terminal terms, NULL = 1 */
    const int32_t *reg_off;         /* [n_reg] component offsets inside a knot */
    const int32_t *reg_dim;         /* [n_reg] */
    const double *reg_R;            /* weights, concatenated over the terms */
} qcb_problem_desc;
qcb_status qcb_problem_create(qcb_problem *out, const qcb_problem_desc *desc);
void qcb_problem_destroy(qcb_problem p);
const char *qcb_problem_last_error(qcb_problem p);   /* p may be NULL: error of the last failed create */
/* n variables, m constraint rows, entries of the Jacobian and of the Hessian of the Lagrangian. */
qcb_status qcb_problem_dims(qcb_problem p, int64_t *n, int64_t *m, int64_t *nnz_jac, int64_t *nnz_hess);
/* 1-based COO structures in the order of the values (Hessian: upper triangle; an entry may occur more than
 * once, the solver sums duplicates). */
qcb_status qcb_problem_jacobian_structure(qcb_problem p, int64_t *rows, int64_t *cols);
qcb_status qcb_problem_hessian_structure(qcb_problem p, int64_t *rows, int64_t *cols);
/* Ipopt's callbacks; all buffers are host memory.  grad: n doubles, every entry written.  g: m doubles.
 * lambda: m multipliers in the row order of eval_g; values = obj_factor * Hess f + sum lambda_i Hess g_i.
 * |1 - F| is differentiated as the reference does (sign(1 - F), quantum_objectives.jl:199). */
qcb_status qcb_problem_eval_f(qcb_problem p, const double *Z, int32_t new_x, double *f);
qcb_status qcb_problem_eval_grad_f(qcb_problem p, const double *Z, int32_t new_x, double *grad);
qcb_status qcb_problem_eval_g(qcb_problem p, const double *Z, int32_t new_x, double *g);
qcb_status qcb_problem_eval_jac_g(qcb_problem p, const double *Z, int32_t new_x, double *jac_vals);
qcb_status qcb_problem_eval_h(qcb_problem p, const double *Z, int32_t new_x, double obj_factor, const double *lambda,
                              double *hess_vals);

#ifdef __cplusplus
}
#endif
#endif /* QCB200_H */
