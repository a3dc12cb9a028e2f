/* qc_port.c -- CPU port of the per-knot collocation evaluation (TEST / BASELINE INFRASTRUCTURE).
 *
 * Restates on the host cores the path that QuantumCollocation.jl hands to
 * DirectTrajOpt.BilinearIntegrator (reference: src/quantum_integrators.jl:28-45 builds the
 * generator closure; the defect x_{k+1} - exp(dt G(u)) x_k is docs/src/index.md:99-105):
 * defect, Jacobian blocks and Hessian-of-Lagrangian blocks for every unit (interval k of
 * system s).  It interprets the same matrix program (qcb_program.h) as the CUDA library,
 * with OpenMP over units, and is validated against the independent scipy oracle
 * (oracle/qc_oracle.py) by tests/test_oracle.py.
 *
 * PARITY STATUS: the reference's own evaluator (Julia, un-vendored DirectTrajOpt.jl 0.8 +
 * ExponentialAction.jl 0.2) cannot be run in this image.  The defect is pinned against the
 * evaluator's logged outputs on the solved trajectories the reference ships
 * (tests/test_reference_fixtures.py); Jacobian / Hessian values are PARITY UNPINNED (derivatives of
 * the pinned function, checked by finite differences); see oracle/qc_oracle.py header.
 *
 * Used only by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs.  Never linked into or loaded by the product library.
 *
 * Output: per-unit canonical arenas exactly as qcb_program.h describes them (dp = d, cp = c).
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "qcb_program.h"

typedef struct {
    int d, m;
    double *fixed[QR_NUM_FIXED];
    double *E[QCB_MAX_SQUARINGS + 1];
    double *W[QCB_MAX_SQUARINGS + 1];
    double **dpersist;   /* [QCB_NUM_DPERSIST * m] : DA2, DA4, DZ1, DZ2 of every direction */
    double *dtemp[QCB_NUM_DTEMP];
    double **L;      /* m + 1 rotating slots */
    int *Lidx;       /* L slot of direction j */
    int Lfree;
    double *tA, *tB; /* transposition temporaries */
    const double *Gj; /* current drive generator */
    int level, j;
    /* sparse view of the drive generators (the reference keeps them as SparseArrays): COO triplets
     * of every drive of the system `sp_src` points at, rebuilt when the system changes */
    const double *sp_src;
    int *sp_r, *sp_k, *sp_off;   /* triplets of drive j: sp_off[j] .. sp_off[j+1] */
    double *sp_v;
    int sp_ok;                   /* all drives sparse enough (nnz <= 8 d) */
} ws_t;

static void transpose(double *T, const double *A, int d) {
    for (int j = 0; j < d; ++j)
        for (int i = 0; i < d; ++i) T[j + (size_t)i * d] = A[i + (size_t)j * d];
}

/* Optional BLAS for the matrix products (a fair CPU denominator for the large-d shape): when the
 * harness hands over a cblas_dgemm with 64-bit integer arguments (the OpenBLAS that numpy bundles:
 * scipy_cblas_dgemm64_), products of dimension >= 32 go through it; OpenMP threads over units, so
 * that BLAS must run single-threaded.  NULL: the plain loops below. */
typedef void (*dgemm64_fn)(int order, int ta, int tb, int64_t M, int64_t N, int64_t K, double alpha,
                           const double *A, int64_t lda, const double *B, int64_t ldb, double beta,
                           double *C, int64_t ldc);
static dgemm64_fn g_dgemm = 0;
void qc_port_set_dgemm(void *fn) { g_dgemm = (dgemm64_fn)fn; }
int qc_port_has_dgemm(void) { return g_dgemm != 0; }

/* C (MxN, ld ldc) += A (MxK col-major, lda) * B (KxN col-major, ldb) */
static void gemm_nn_acc(double *restrict C, const double *restrict A, const double *restrict B,
                        int M, int N, int K, int lda, int ldb, int ldc) {
    if (g_dgemm && M >= 32 && K >= 32) {
        g_dgemm(102 /* CblasColMajor */, 111, 111 /* CblasNoTrans */, M, N, K, 1.0, A, lda, B, ldb, 1.0, C, ldc);
        return;
    }
    int j = 0;
    for (; j + 4 <= N; j += 4) {
        double *c0 = C + (size_t)j * ldc, *c1 = c0 + ldc, *c2 = c1 + ldc, *c3 = c2 + ldc;
        for (int k = 0; k < K; ++k) {
            const double *a = A + (size_t)k * lda;
            double b0 = B[k + (size_t)j * ldb], b1 = B[k + (size_t)(j + 1) * ldb];
            double b2 = B[k + (size_t)(j + 2) * ldb], b3 = B[k + (size_t)(j + 3) * ldb];
#pragma omp simd
            for (int i = 0; i < M; ++i) {
                double av = a[i];
                c0[i] += av * b0; c1[i] += av * b1; c2[i] += av * b2; c3[i] += av * b3;
            }
        }
    }
    for (; j < N; ++j) {
        double *c0 = C + (size_t)j * ldc;
        for (int k = 0; k < K; ++k) {
            const double *a = A + (size_t)k * lda;
            double b0 = B[k + (size_t)j * ldb];
#pragma omp simd
            for (int i = 0; i < M; ++i) c0[i] += a[i] * b0;
        }
    }
}

static double *resolve(ws_t *w, int r) {
    if (r < QR_NUM_FIXED) return w->fixed[r];
    switch (r) {
    case QR_EL: return w->E[w->level];
    case QR_EL1: return w->E[w->level + 1];
    case QR_WL: return w->W[w->level];
    case QR_WL1: return w->W[w->level + 1];
    case QR_E0: return w->E[0];
    case QR_W0: return w->W[0];
    case QR_GJ: return (double *)w->Gj;
    case QR_L: return w->L[w->Lidx[w->j]];
    case QR_LN: return w->L[w->Lfree];
    default:
        if (qcb_dpersist(r) >= 0) return w->dpersist[qcb_dpersist(r) * w->m + w->j];
        return w->dtemp[qcb_dtemp(r)];
    }
}

static void run_op(ws_t *w, const qcb_op *op) {
    int d = w->d;
    size_t dd = (size_t)d * d;
    double *dst = resolve(w, op->dst);
    memset(dst, 0, dd * sizeof(double));
    for (int t = 0; t < op->nterm; ++t) {
        const double *A = resolve(w, op->term[t].a), *B = resolve(w, op->term[t].b);
        if (w->sp_ok && op->term[t].a == QR_GJ && !op->term[t].tb) {       /* G_j * B or G_j^T * B */
            int lo = w->sp_off[w->j], hi = w->sp_off[w->j + 1], tr = op->term[t].ta;
            for (int q = 0; q < d; ++q) {
                double *cq = dst + (size_t)q * d;
                const double *bq = B + (size_t)q * d;
                for (int e = lo; e < hi; ++e) {
                    int r = tr ? w->sp_k[e] : w->sp_r[e], k = tr ? w->sp_r[e] : w->sp_k[e];
                    cq[r] += w->sp_v[e] * bq[k];
                }
            }
            continue;
        }
        if (w->sp_ok && op->term[t].b == QR_GJ && !op->term[t].ta && !op->term[t].tb) {   /* A * G_j */
            int lo = w->sp_off[w->j], hi = w->sp_off[w->j + 1];
            for (int e = lo; e < hi; ++e) {
                double *ck = dst + (size_t)w->sp_k[e] * d;
                const double *ar = A + (size_t)w->sp_r[e] * d;
                double v = w->sp_v[e];
#pragma omp simd
                for (int i = 0; i < d; ++i) ck[i] += v * ar[i];
            }
            continue;
        }
        if (op->term[t].ta) { transpose(w->tA, A, d); A = w->tA; }
        if (op->term[t].tb) { transpose(w->tB, B, d); B = w->tB; }
        gemm_nn_acc(dst, A, B, d, d, d, d, d, d);
    }
    for (int r = 0; r < op->nlin; ++r) {
        const double *R = resolve(w, op->lin[r].reg);
        double cf = QCB_TAYLOR[op->lin[r].coef];
#pragma omp simd
        for (size_t e = 0; e < dd; ++e) dst[e] += cf * R[e];
    }
    if (op->ident >= 0)
        for (int i = 0; i < d; ++i) dst[i + (size_t)i * d] += QCB_TAYLOR[op->ident];
}

static double dot(const double *a, const double *b, size_t n) {
    double s = 0;
#pragma omp simd reduction(+ : s)
    for (size_t e = 0; e < n; ++e) s += a[e] * b[e];
    return s;
}

static ws_t *ws_new(int d, int m) {
    ws_t *w = (ws_t *)calloc(1, sizeof(ws_t));
    size_t dd = (size_t)d * d;
    w->d = d; w->m = m;
    for (int i = 0; i < QR_NUM_FIXED; ++i) w->fixed[i] = (double *)malloc(dd * sizeof(double));
    for (int i = 0; i <= QCB_MAX_SQUARINGS; ++i) {
        w->E[i] = (double *)malloc(dd * sizeof(double));
        w->W[i] = (double *)malloc(dd * sizeof(double));
    }
    w->dpersist = (double **)malloc((size_t)QCB_NUM_DPERSIST * (m + 1) * sizeof(double *));
    for (int i = 0; i < QCB_NUM_DPERSIST * m; ++i) w->dpersist[i] = (double *)malloc(dd * sizeof(double));
    for (int i = 0; i < QCB_NUM_DTEMP; ++i) w->dtemp[i] = (double *)malloc(dd * sizeof(double));
    w->L = (double **)malloc((m + 1) * sizeof(double *));
    for (int i = 0; i <= m; ++i) w->L[i] = (double *)malloc(dd * sizeof(double));
    w->Lidx = (int *)malloc((m + 1) * sizeof(int));
    w->tA = (double *)malloc(dd * sizeof(double));
    w->tB = (double *)malloc(dd * sizeof(double));
    w->sp_r = (int *)malloc((size_t)8 * d * (m + 1) * sizeof(int));
    w->sp_k = (int *)malloc((size_t)8 * d * (m + 1) * sizeof(int));
    w->sp_v = (double *)malloc((size_t)8 * d * (m + 1) * sizeof(double));
    w->sp_off = (int *)malloc((m + 2) * sizeof(int));
    return w;
}
static void ws_free(ws_t *w) {
    for (int i = 0; i < QR_NUM_FIXED; ++i) free(w->fixed[i]);
    for (int i = 0; i <= QCB_MAX_SQUARINGS; ++i) { free(w->E[i]); free(w->W[i]); }
    for (int i = 0; i < QCB_NUM_DPERSIST * w->m; ++i) free(w->dpersist[i]);
    free(w->dpersist);
    for (int i = 0; i < QCB_NUM_DTEMP; ++i) free(w->dtemp[i]);
    for (int i = 0; i <= w->m; ++i) free(w->L[i]);
    free(w->sp_r); free(w->sp_k); free(w->sp_v); free(w->sp_off);
    free(w->L); free(w->Lidx); free(w->tA); free(w->tB); free(w);
}

/* one unit; returns 0 or -1 (norm out of range / non-finite) */
static int unit(ws_t *w, int d, int c, int m, const double *G0, const double *Gd,
                const double *x, const double *xn, const double *mu, const double *u, double dt,
                int want, double *delta, double *ja, double *ha) {
    size_t dd = (size_t)d * d, dc = (size_t)d * c;
    double *G = w->fixed[QR_G], *A = w->fixed[QR_A];
    memcpy(G, G0, dd * sizeof(double));
    for (int j = 0; j < m; ++j)
        for (size_t e = 0; e < dd; ++e) G[e] += u[j] * Gd[(size_t)j * dd + e];
    double nrm = 0;
    for (int q = 0; q < d; ++q) {
        double cs = 0;
        for (int r = 0; r < d; ++r) cs += fabs(dt * G[r + (size_t)q * d]);
        if (!(cs <= nrm)) nrm = cs; /* NaN-propagating max */
    }
    int s = qcb_squarings(nrm);
    if (s < 0) return -1;
    double tau = ldexp(dt, -s);
    for (size_t e = 0; e < dd; ++e) A[e] = tau * G[e];
    /* polynomial + squarings */
    w->level = 0;
    for (int i = 0; i < QCB_NOPS(QCB_P1); ++i) run_op(w, &QCB_P1[i]);
    for (int l = 0; l < s; ++l) { w->level = l; run_op(w, &QCB_P2[0]); }
    const double *E = w->E[s];
    /* delta = xn - E x */
    double *EX = w->tA;
    memset(EX, 0, dc * sizeof(double));
    gemm_nn_acc(EX, E, x, d, c, d, d, d, d);
    for (size_t e = 0; e < dc; ++e) delta[e] = xn[e] - EX[e];
    if (!want) return 0;
    int hess = (want & 2) && mu;
    double *H = NULL;
    if (hess) {
        /* W = M X^T */
        double *Wt = w->W[s];
        memset(Wt, 0, dd * sizeof(double));
        for (int b = 0; b < c; ++b)
            for (int q = 0; q < d; ++q) {
                double xv = x[q + (size_t)b * d];
                for (int r = 0; r < d; ++r) Wt[r + (size_t)q * d] += mu[r + (size_t)b * d] * xv;
            }
        for (int l = s - 1; l >= 0; --l) { w->level = l; run_op(w, &QCB_P3[0]); }
        w->level = 0;
        for (int i = 0; i < QCB_NOPS(QCB_P4); ++i) run_op(w, &QCB_P4[i]);
        H = (double *)calloc((size_t)m * m, sizeof(double));
    }
    for (int j = 0; j <= m; ++j) w->Lidx[j] = j;
    w->Lfree = m;
    if (w->sp_src != Gd) {      /* sparse view of this system's drive generators */
        int n = 0, cap = 8 * d;
        w->sp_src = Gd; w->sp_ok = getenv("QC_PORT_DENSE_DRIVES") ? 0 : 1;
        for (int j = 0; j < m && w->sp_ok; ++j) {
            const double *Gj = Gd + (size_t)j * dd;
            w->sp_off[j] = n;
            for (int k = 0; k < d && w->sp_ok; ++k)
                for (int r = 0; r < d; ++r)
                    if (Gj[r + (size_t)k * d] != 0.0) {
                        if (n - w->sp_off[j] >= cap) { w->sp_ok = 0; break; }
                        w->sp_r[n] = r; w->sp_k[n] = k; w->sp_v[n] = Gj[r + (size_t)k * d]; ++n;
                    }
        }
        w->sp_off[m] = n;
    }
    for (int j = 0; j < m; ++j) {
        w->j = j; w->Gj = Gd + (size_t)j * dd; w->level = 0;
        for (int i = 0; i < QCB_NOPS(QCB_P5); ++i) run_op(w, &QCB_P5[i]);
    }
    if (hess) {     /* second order at the polynomial level: one product per node and direction */
        for (int o = 0; o < QCB_NOPS(QCB_P6); ++o)
            for (int j = 0; j < m; ++j) {
                w->j = j; w->Gj = Gd + (size_t)j * dd; w->level = 0;
                run_op(w, &QCB_P6[o]);
                const double *U = resolve(w, QR_U);
                for (int i = 0; i < m; ++i) {
                    w->j = i; w->Gj = Gd + (size_t)i * dd;
                    H[j * m + i] += dot(U, resolve(w, QCB_P6_PARTNER[o]), dd);
                }
            }
    }
    for (int l = 0; l < s; ++l) {
        w->level = l;
        if (hess)
            for (int j = 0; j < m; ++j) {
                w->j = j;
                run_op(w, &QCB_P7U[0]);
                const double *U = resolve(w, QR_U);
                for (int i = 0; i < m; ++i) H[j * m + i] += dot(U, w->L[w->Lidx[i]], dd);
            }
        for (int j = 0; j < m; ++j) {
            w->j = j;
            run_op(w, &QCB_P7L[0]);
            int old = w->Lidx[j];
            w->Lidx[j] = w->Lfree;
            w->Lfree = old;
        }
    }
    /* final blocks */
    double *GE = w->fixed[QR_A2], *WE = w->fixed[QR_A3], *GW = w->fixed[QR_A4];
    memset(GE, 0, dd * sizeof(double));
    gemm_nn_acc(GE, G, E, d, d, d, d, d, d);
    /* Jacobian arena: E | JU_j | JT | 0 | 1 */
    memcpy(ja, E, dd * sizeof(double));
    double *JU = ja + dd, *JT = JU + (size_t)m * dc;
    memset(JU, 0, ((size_t)m + 1) * dc * sizeof(double));
    for (int j = 0; j < m; ++j) {
        gemm_nn_acc(JU + (size_t)j * dc, w->L[w->Lidx[j]], x, d, c, d, d, d, d);
        for (size_t e = 0; e < dc; ++e) JU[(size_t)j * dc + e] *= -tau;
    }
    gemm_nn_acc(JT, GE, x, d, c, d, d, d, d);
    for (size_t e = 0; e < dc; ++e) JT[e] = -JT[e];
    JT[dc] = 0.0; JT[dc + 1] = 1.0;
    if (hess) {
        const double *Wt = w->W[s];
        double *HXU = ha, *HXT = ha + (size_t)m * dc, *HUU = HXT + dc, *HUT = HUU + (size_t)m * m, *HTT = HUT + m;
        memset(ha, 0, ((size_t)m + 1) * dc * sizeof(double));
        for (int j = 0; j < m; ++j) {
            transpose(w->tA, w->L[w->Lidx[j]], d);
            gemm_nn_acc(HXU + (size_t)j * dc, w->tA, mu, d, c, d, d, d, d);
            for (size_t e = 0; e < dc; ++e) HXU[(size_t)j * dc + e] *= -tau;
        }
        transpose(w->tA, GE, d);
        gemm_nn_acc(HXT, w->tA, mu, d, c, d, d, d, d);
        for (size_t e = 0; e < dc; ++e) HXT[e] = -HXT[e];
        /* WE = W E^T, GW = G^T W */
        memset(WE, 0, dd * sizeof(double));
        transpose(w->tB, E, d);
        gemm_nn_acc(WE, Wt, w->tB, d, d, d, d, d, d);
        memset(GW, 0, dd * sizeof(double));
        transpose(w->tA, G, d);
        gemm_nn_acc(GW, w->tA, Wt, d, d, d, d, d, d);
        for (int i = 0; i < m; ++i)
            for (int j = 0; j < m; ++j) HUU[i * m + j] = -tau * tau * (H[i * m + j] + H[j * m + i]);
        for (int j = 0; j < m; ++j)
            HUT[j] = -(dot(WE, Gd + (size_t)j * dd, dd) + tau * dot(GW, w->L[w->Lidx[j]], dd));
        HTT[0] = -dot(GW, GE, dd);
        HTT[1] = 0.0;
        free(H);
    }
    return 0;
}

size_t qc_port_jac_arena(int d, int c, int m) { return (size_t)d * d + ((size_t)m + 1) * d * c + 2; }
size_t qc_port_hess_arena(int d, int c, int m) { return ((size_t)m + 1) * d * c + (size_t)m * m + m + 2; }

/* Whole slab [k_begin, k_end) of every system.  Layout of Z: knot-major, width z_dim.
 * mu: [S][K][x_dim] (K = N-1, full trajectory indexing).  Outputs indexed [s][k - k_begin].
 * want: bit0 Jacobian arena, bit1 Hessian arena.  Returns number of failed units. */
int qc_port_eval(int d, int c, int m, int N, int S, int z_dim, const int *x_off, int u_off, int dt_off,
                 const double *G0, const double *Gd, const double *Z, const double *mu,
                 int k_begin, int k_end, int want, double *delta, double *jarena, double *harena,
                 int nthreads) {
    int K = N - 1, nk = k_end - k_begin, xd = d * c;
    size_t dd = (size_t)d * d, JA = qc_port_jac_arena(d, c, m), HA = qc_port_hess_arena(d, c, m);
    int fails = 0;
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
#pragma omp parallel reduction(+ : fails)
    {
        ws_t *w = ws_new(d, m);
        double *jtmp = (double *)malloc(JA * sizeof(double));
        double *htmp = (double *)malloc(HA * sizeof(double));
#pragma omp for schedule(dynamic, 1)
        for (int un = 0; un < S * nk; ++un) {
            int s = un / nk, kl = un % nk, k = k_begin + kl;
            const double *zk = Z + (size_t)k * z_dim, *zn = zk + z_dim;
            const double *mk = mu ? mu + ((size_t)s * K + k) * xd : NULL;
            int rc = unit(w, d, c, m, G0 + (size_t)s * dd, Gd + (size_t)s * m * dd, zk + x_off[s], zn + x_off[s],
                          mk, zk + u_off, zk[dt_off], want, delta + (size_t)un * xd,
                          (want & 1) && jarena ? jarena + (size_t)un * JA : jtmp,
                          (want & 2) && harena ? harena + (size_t)un * HA : htmp);
            if (rc) ++fails;
        }
        free(jtmp); free(htmp);
        ws_free(w);
    }
    return fails;
}

/* Same evaluation, but delivering what the reference's evaluator must deliver to Ipopt: the COO value
 * vectors in the caller's sparsity order.  jcode / hcode: [S][jnnz] / [S][hnnz] slot codes (bit 31 =
 * negate, bits 0..30 = arena offset) as built by oracle/qc_port.py:arena_codes.  Outputs are indexed
 * [s][k - k_begin][entry]. */
int qc_port_eval_coo(int d, int c, int m, int N, int S, int z_dim, const int *x_off, int u_off, int dt_off,
                     const double *G0, const double *Gd, const double *Z, const double *mu,
                     int k_begin, int k_end, const unsigned *jcode, long long jnnz, const unsigned *hcode,
                     long long hnnz, double *delta, double *jac, double *hess, int nthreads) {
    int K = N - 1, nk = k_end - k_begin, xd = d * c;
    size_t dd = (size_t)d * d, JA = qc_port_jac_arena(d, c, m), HA = qc_port_hess_arena(d, c, m);
    int fails = 0;
    int want = (jac ? 1 : 0) | ((hess && mu) ? 2 : 0);
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
#pragma omp parallel reduction(+ : fails)
    {
        ws_t *w = ws_new(d, m);
        double *jtmp = (double *)malloc(JA * sizeof(double));
        double *htmp = (double *)malloc(HA * sizeof(double));
#pragma omp for schedule(dynamic, 1)
        for (int un = 0; un < S * nk; ++un) {
            int s = un / nk, kl = un % nk, k = k_begin + kl;
            const double *zk = Z + (size_t)k * z_dim, *zn = zk + z_dim;
            const double *mk = mu ? mu + ((size_t)s * K + k) * xd : NULL;
            int rc = unit(w, d, c, m, G0 + (size_t)s * dd, Gd + (size_t)s * m * dd, zk + x_off[s], zn + x_off[s],
                          mk, zk + u_off, zk[dt_off], want, delta + (size_t)un * xd, jtmp, htmp);
            if (rc) { ++fails; continue; }
            if (want & 1) {
                const unsigned *cd = jcode + (size_t)s * jnnz;
                double *o = jac + (size_t)un * jnnz;
                for (long long e = 0; e < jnnz; ++e) {
                    double v = jtmp[cd[e] & 0x7fffffffu];
                    o[e] = (cd[e] >> 31) ? -v : v;
                }
            }
            if (want & 2) {
                const unsigned *cd = hcode + (size_t)s * hnnz;
                double *o = hess + (size_t)un * hnnz;
                for (long long e = 0; e < hnnz; ++e) {
                    double v = htmp[cd[e] & 0x7fffffffu];
                    o[e] = (cd[e] >> 31) ? -v : v;
                }
            }
        }
        free(jtmp); free(htmp);
        ws_free(w);
    }
    return fails;
}
