// qcb_small.cuh -- small-d path: one CTA per unit (interval k of system s), the whole matrix
// program of qcb_program.h interpreted in a single launch with every virtual register in
// shared memory.  G_drift / G_j are read through L1/L2 (they are shared by every unit of a
// system); the unit's own state, multipliers and all intermediates never leave the SM; the
// COO values are written once, coalesced, straight in the caller's sparsity order.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "qcb_program.h"

namespace qcb {

__constant__ qcb_op c_P1[6], c_P2[1], c_P3[1], c_P4[5], c_P5[6], c_P6[6], c_P7S[1], c_P7L[1];
__constant__ double c_taylor[QCB_TAYLOR_DEGREE + 1];

struct SmallParams {
    int d, c, m, z_dim, u_off, dt_off, K, nk, k_begin, S;
    const int *x_off;
    const double *G0, *Gd;          // device, unpadded col-major [S][d*d], [S][m][d*d]
    const double *Z, *mu;
    double *delta, *jac, *hess;     // outputs (slab-local unit indexing)
    const uint32_t *jcode, *hcode;  // [S][nnz]
    long long jnnz, hnnz;
    int want;                       // bit0 jac, bit1 hess
    int smax;                       // chain capacity allocated in shared memory
    int *bad_flag;
};

struct SmallCtx {
    int d, dd, m, level, j;
    double *fixed;   // [QR_NUM_FIXED][dd]
    double *E, *W;   // [smax+1][dd]
    double *dslot;   // [QCB_NUM_DSLOTS][dd]
    double *L;       // [m+1][dd]
    const double *Gj;
    int *Lidx;       // [m+1] in shared memory (slot of direction j; entry m = free slot)
};

__device__ __forceinline__ double *sm_resolve(const SmallCtx &x, int r) {
    if (r < QR_NUM_FIXED) return x.fixed + r * x.dd;
    switch (r) {
    case QR_EL: return x.E + x.level * x.dd;
    case QR_EL1: return x.E + (x.level + 1) * x.dd;
    case QR_WL: return x.W + x.level * x.dd;
    case QR_WL1: return x.W + (x.level + 1) * x.dd;
    case QR_E0: return x.E;
    case QR_W0: return x.W;
    case QR_GJ: return const_cast<double *>(x.Gj);
    case QR_L: return x.L + x.Lidx[x.j] * x.dd;
    case QR_LN: return x.L + x.Lidx[x.m] * x.dd;
    default: return x.dslot + qcb_dslot(r) * x.dd;
    }
}

// dst = sum_t op(a)op(b) + sum_r c*reg + c*I ; whole CTA cooperates, one output element per
// thread per pass.  Ends with a block barrier.
__device__ void sm_run_op(const SmallCtx &x, const qcb_op &op) {
    const int d = x.d, dd = x.dd;
    double *dst = sm_resolve(x, op.dst);
    const double *A[QCB_MAX_TERMS], *B[QCB_MAX_TERMS], *R[QCB_MAX_LIN];
    for (int t = 0; t < op.nterm; ++t) { A[t] = sm_resolve(x, op.term[t].a); B[t] = sm_resolve(x, op.term[t].b); }
    for (int r = 0; r < op.nlin; ++r) R[r] = sm_resolve(x, op.lin[r].reg);
    for (int e = threadIdx.x; e < dd; e += blockDim.x) {
        const int r = e % d, q = e / d;
        double acc = 0.0;
        for (int t = 0; t < op.nterm; ++t) {
            const double *a = A[t] + (op.term[t].ta ? r * d : r);
            const int as = op.term[t].ta ? 1 : d;
            const double *b = B[t] + (op.term[t].tb ? q : q * d);
            const int bs = op.term[t].tb ? d : 1;
            for (int k = 0; k < d; ++k) acc = fma(a[k * as], b[k * bs], acc);
        }
        for (int l = 0; l < op.nlin; ++l) acc = fma(c_taylor[op.lin[l].coef], R[l][e], acc);
        if (op.ident >= 0 && r == q) acc += c_taylor[op.ident];
        dst[e] = acc;
    }
    __syncthreads();
}

// block-wide Frobenius inner product, result broadcast to all threads
__device__ double sm_dot(const double *a, const double *b, int n, double *red) {
    double v = 0.0;
    for (int e = threadIdx.x; e < n; e += blockDim.x) v = fma(a[e], b[e], v);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    double s = 0.0;
    for (int w = 0; w < (int)((blockDim.x + 31) >> 5); ++w) s += red[w];
    __syncthreads();
    return s;
}

// C (d x c) = op(A) (d x d) * P (d x c), scaled
__device__ void sm_panel(double *C, const double *A, int ta, const double *P, int d, int c, double scale) {
    for (int e = threadIdx.x; e < d * c; e += blockDim.x) {
        const int r = e % d, b = e / d;
        const double *a = A + (ta ? r * d : r);
        const int as = ta ? 1 : d;
        double acc = 0.0;
        for (int k = 0; k < d; ++k) acc = fma(a[k * as], P[k + b * d], acc);
        C[e] = scale * acc;
    }
}

static inline size_t small_smem_doubles(int d, int c, int m, int smax) {
    size_t dd = (size_t)d * d, dc = (size_t)d * c;
    size_t n = (QR_NUM_FIXED + 2 * (size_t)(smax + 1) + QCB_NUM_DSLOTS + (m + 1)) * dd;
    n += 3 * dc;                                   // X, Xn, M panels
    n += dd + (m + 1) * dc + 2;                    // Jacobian arena
    n += (m + 1) * dc + (size_t)m * m + m + 2;     // Hessian arena
    n += (size_t)m * m + 32;                       // H accumulator + reduction scratch
    n += (m + 2) / 2 + 2;                          // Lidx (ints)
    return n;
}

__global__ void small_unit_kernel(const __grid_constant__ SmallParams p) {
    extern __shared__ __align__(16) double sm[];
    const int d = p.d, c = p.c, m = p.m, dd = d * d, dc = d * c;
    const int ug = blockIdx.x;                       // slab-local unit index = sys*nk + kl
    const int sys = ug / p.nk, k = p.k_begin + ug % p.nk;
    const double *zk = p.Z + (long long)k * p.z_dim, *zn = zk + p.z_dim;
    const double *G0 = p.G0 + (long long)sys * dd, *Gd = p.Gd + (long long)sys * m * dd;

    SmallCtx x;
    x.d = d; x.dd = dd; x.m = m; x.level = 0; x.j = 0;
    double *ptr = sm;
    x.fixed = ptr; ptr += QR_NUM_FIXED * dd;
    x.E = ptr; ptr += (p.smax + 1) * dd;
    x.W = ptr; ptr += (p.smax + 1) * dd;
    x.dslot = ptr; ptr += QCB_NUM_DSLOTS * dd;
    x.L = ptr; ptr += (m + 1) * dd;
    double *X = ptr; ptr += dc;
    double *Xn = ptr; ptr += dc;
    double *Mu = ptr; ptr += dc;
    double *JA = ptr; ptr += dd + (m + 1) * dc + 2;
    double *HA = ptr; ptr += (m + 1) * dc + m * m + m + 2;
    double *Hacc = ptr; ptr += m * m;
    double *red = ptr; ptr += 32;
    x.Lidx = reinterpret_cast<int *>(ptr);
    x.Gj = Gd;

    const bool want_jac = p.want & 1, want_hess = (p.want & 2) && p.mu;
    double *G = x.fixed + QR_G * dd, *A = x.fixed + QR_A * dd;
    const double dt = zk[p.dt_off];

    // ---- load unit inputs, G = G0 + sum u_j G_j ----
    for (int e = threadIdx.x; e < dc; e += blockDim.x) {
        X[e] = zk[p.x_off[sys] + e];
        Xn[e] = zn[p.x_off[sys] + e];
        Mu[e] = want_hess ? p.mu[((long long)sys * p.K + k) * dc + e] : 0.0;
    }
    for (int e = threadIdx.x; e < dd; e += blockDim.x) {
        double v = G0[e];
        for (int j = 0; j < m; ++j) v = fma(zk[p.u_off + j], Gd[j * dd + e], v);
        G[e] = v;
    }
    if (threadIdx.x <= m) x.Lidx[threadIdx.x] = threadIdx.x;
    for (int e = threadIdx.x; e < m * m; e += blockDim.x) Hacc[e] = 0.0;
    __syncthreads();
    // ---- ||dt G||_1 -> number of squarings of this unit ----
    double nrm = 0.0;
    {
        double mine = 0.0;   // thread q sums column q (d <= blockDim guaranteed by the host)
        bool bad = false;
        for (int q = threadIdx.x; q < d; q += blockDim.x) {
            double cs = 0.0;
            for (int r = 0; r < d; ++r) cs += fabs(dt * G[r + q * d]);
            if (!isfinite(cs)) bad = true;
            mine = fmax(mine, cs);
        }
        if (bad) mine = INFINITY;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mine = fmax(mine, __shfl_xor_sync(0xffffffffu, mine, o));
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = mine;
        __syncthreads();
        for (int w = 0; w < (int)((blockDim.x + 31) >> 5); ++w) nrm = fmax(nrm, red[w]);
        __syncthreads();
    }
    int s = qcb_squarings(nrm);
    if (s < 0 || s > p.smax) {      // uniform across the CTA
        if (threadIdx.x == 0) atomicExch(p.bad_flag, 1);
        return;
    }
    const double tau = ldexp(dt, -s);
    for (int e = threadIdx.x; e < dd; e += blockDim.x) A[e] = tau * G[e];
    __syncthreads();

    // ---- polynomial + squarings ----
    for (int i = 0; i < 6; ++i) sm_run_op(x, c_P1[i]);
    for (int l = 0; l < s; ++l) { x.level = l; sm_run_op(x, c_P2[0]); }
    const double *E = x.E + s * dd;

    // ---- defect ----
    {
        double *o = p.delta + (long long)ug * dc;
        for (int e = threadIdx.x; e < dc; e += blockDim.x) {
            const int r = e % d, b = e / d;
            double acc = 0.0;
            for (int q = 0; q < d; ++q) acc = fma(E[r + q * d], X[q + b * d], acc);
            o[e] = Xn[e] - acc;
        }
    }
    if (!want_jac && !want_hess) return;

    double *Wt = x.W + s * dd;
    if (want_hess) {
        for (int e = threadIdx.x; e < dd; e += blockDim.x) {   // W = M X^T
            const int r = e % d, q = e / d;
            double acc = 0.0;
            for (int b = 0; b < c; ++b) acc = fma(Mu[r + b * d], X[q + b * d], acc);
            Wt[e] = acc;
        }
        __syncthreads();
        for (int l = s - 1; l >= 0; --l) { x.level = l; sm_run_op(x, c_P3[0]); }
        x.level = 0;
        for (int i = 0; i < 5; ++i) sm_run_op(x, c_P4[i]);
    }
    // ---- per-direction tangents ----
    for (int j = 0; j < m; ++j) {
        x.j = j; x.Gj = Gd + (long long)j * dd; x.level = 0;
        for (int i = 0; i < 6; ++i) sm_run_op(x, c_P5[i]);
        if (want_hess) {
            for (int i = 0; i < 6; ++i) sm_run_op(x, c_P6[i]);
            const double *dAb = sm_resolve(x, QR_DAB);
            for (int i = 0; i < m; ++i) {
                double v = sm_dot(dAb, Gd + (long long)i * dd, dd, red);
                if (threadIdx.x == 0) Hacc[j * m + i] = v;
            }
        }
    }
    // ---- squaring levels ----
    for (int l = 0; l < s; ++l) {
        x.level = l;
        if (want_hess) {
            for (int j = 0; j < m; ++j) {
                x.j = j;
                sm_run_op(x, c_P7S[0]);
                const double *Sj = sm_resolve(x, QR_S);
                for (int i = 0; i < m; ++i) {
                    double v = sm_dot(Sj, x.L + x.Lidx[i] * dd, dd, red);
                    if (threadIdx.x == 0) Hacc[j * m + i] += v;
                }
            }
        }
        for (int j = 0; j < m; ++j) {
            x.j = j;
            sm_run_op(x, c_P7L[0]);
            if (threadIdx.x == 0) { int old = x.Lidx[j]; x.Lidx[j] = x.Lidx[m]; x.Lidx[m] = old; }
            __syncthreads();
        }
    }
    // ---- final blocks ----
    double *GE = x.fixed + QR_A2 * dd, *WE = x.fixed + QR_A3 * dd, *GW = x.fixed + QR_A4 * dd;
    for (int e = threadIdx.x; e < dd; e += blockDim.x) {
        const int r = e % d, q = e / d;
        double g1 = 0.0, g2 = 0.0, g3 = 0.0;
        for (int kk = 0; kk < d; ++kk) {
            g1 = fma(G[r + kk * d], E[kk + q * d], g1);            // G E
            if (want_hess) {
                g2 = fma(Wt[r + kk * d], E[q + kk * d], g2);       // W E^T
                g3 = fma(G[kk + r * d], Wt[kk + q * d], g3);       // G^T W
            }
        }
        GE[e] = g1; WE[e] = g2; GW[e] = g3;
        JA[e] = E[e];
    }
    __syncthreads();
    double *JU = JA + dd, *JT = JU + m * dc;
    for (int j = 0; j < m; ++j) sm_panel(JU + j * dc, x.L + x.Lidx[j] * dd, 0, X, d, c, -tau);
    sm_panel(JT, GE, 0, X, d, c, -1.0);
    if (threadIdx.x == 0) { JT[dc] = 0.0; JT[dc + 1] = 1.0; }
    if (want_hess) {
        double *HXU = HA, *HXT = HA + m * dc, *HUU = HXT + dc, *HUT = HUU + m * m, *HTT = HUT + m;
        for (int j = 0; j < m; ++j) sm_panel(HXU + j * dc, x.L + x.Lidx[j] * dd, 1, Mu, d, c, -tau);
        sm_panel(HXT, GE, 1, Mu, d, c, -1.0);
        for (int e = threadIdx.x; e < m * m; e += blockDim.x) {
            const int i = e / m, j = e % m;
            HUU[e] = -0.5 * tau * tau * (Hacc[i * m + j] + Hacc[j * m + i]);
        }
        for (int j = 0; j < m; ++j) {
            double c1 = sm_dot(WE, Gd + (long long)j * dd, dd, red);
            double c2 = sm_dot(GW, x.L + x.Lidx[j] * dd, dd, red);
            if (threadIdx.x == 0) HUT[j] = -(c1 + tau * c2);
        }
        double c3 = sm_dot(GW, GE, dd, red);
        if (threadIdx.x == 0) { HTT[0] = -c3; HTT[1] = 0.0; }
    }
    __syncthreads();
    // ---- coalesced COO stores in the caller's order ----
    if (want_jac && p.jac) {
        const uint32_t *cd = p.jcode + (long long)sys * p.jnnz;
        double *o = p.jac + (long long)ug * p.jnnz;
        for (long long e = threadIdx.x; e < p.jnnz; e += blockDim.x) {
            uint32_t cc = cd[e];
            double v = JA[cc & 0x7fffffffu];
            o[e] = (cc >> 31) ? -v : v;
        }
    }
    if (want_hess && p.hess) {
        const uint32_t *cd = p.hcode + (long long)sys * p.hnnz;
        double *o = p.hess + (long long)ug * p.hnnz;
        for (long long e = threadIdx.x; e < p.hnnz; e += blockDim.x) {
            uint32_t cc = cd[e];
            double v = HA[cc & 0x7fffffffu];
            o[e] = (cc >> 31) ? -v : v;
        }
    }
}

}  // namespace qcb
