// qcb_small.cuh -- small-d path: the whole matrix program of qcb_program.h for a unit (interval k
// of system s) runs inside ONE launch with every virtual register in shared memory.  A CTA holds
// `upc` units side by side, `tpu` threads each (one output element per thread and pass); the
// squaring count is uniform per CTA (max over its units).  G_drift / G_j are read through L1/L2
// (shared by every unit of a system); the unit's state, multipliers and all intermediates never
// leave the SM; the COO values are written once, coalesced, in the caller's sparsity order.
// The kernel is instantiated for the common compile-time dimensions (D = 4 ... 18) so that the
// inner products unroll with immediate offsets; D = 0 is the run-time-d fallback.  The op
// interpreter is kept out of line on purpose: one copy of its code stays hot in the I-cache.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <utility>
#include "qcb_program.h"

// rows of the d = 4 register tile: measured on cfg4 -- 1: 0.352 ms, 2: 0.364 ms, 4: 0.330 ms.  (Running the m
// drive directions side by side on m x 4 threads per unit was also measured: 0.40 ms -- 254 registers per
// thread and idle halves in the per-unit ops cost more than the shorter dependency chain gains.)
#ifndef QCB_TILE4
#define QCB_TILE4 4
#endif
namespace qcb {

__constant__ qcb_op c_P1[6], c_P2[1], c_P3[1], c_P4[5], c_P5[6], c_P6[4], c_P7U[1], c_P7L[1];
__constant__ uint8_t c_P6_partner[4];
__constant__ double c_taylor[QCB_TAYLOR_DEGREE + 1];

struct SmallParams {
    int d, c, m, z_dim, u_off, dt_off, K, nk, k_begin, S;
    int units;                      // slab-local units = S * nk
    int upc, tpu;                   // units per CTA, threads per unit
    int unit_doubles;               // shared-memory doubles per unit
    const int *x_off;
    const double *G0, *Gd;          // device, col-major [S][d*d], [S][m][d*d]
    const double *Z, *mu;
    double *delta, *jac, *hess;     // outputs (slab-local unit indexing)
    const uint32_t *jcode, *hcode;  // [S][nnz]
    long long jnnz, hnnz;
    int want;                       // bit0 jac, bit1 hess
    double *E_out;                  // optional [units][d*d]: the propagator of every unit (rollouts)
    double *ja_out, *ha_out;        // optional [units][ja_stride / ha_stride]: the value arenas instead of the COO vectors
    long long ja_stride, ha_stride; //   (compact host transfer: the host cores expand them)
    int smax;                       // chain capacity allocated in shared memory
    int *bad_flag;
};

// Matrices live in shared memory with an odd leading dimension (ld = d + 1 for even d): column and
// row (transposed) walks are then both free of bank conflicts.
// The compile-time instances (register-tiled products, 16-byte shared-memory loads) use ld = d; the
// run-time-d interpreter keeps an odd leading dimension (scalar row and column walks without conflicts).
QCB_HD static inline bool small_static_d(int d) { return d == 4 || d == 6 || d == 8 || d == 10 || d == 12 || d == 16 || d == 18; }
// Rows of the register tile a thread owns in the products of the compile-time instances (1: one element
// per thread).  Measured: the tile form (a whole column per thread) wins for d = 4 (cfg1 / cfg4); for d = 6 and d = 18 the fewer
// threads per unit it implies slow the element-wise phases (loads, panels, COO stores) by more than the
// products gain (cfg2 0.056 -> 0.081 ms, cfg3 0.37 -> 0.77 ms), so those stay at one element per thread.
QCB_HD static inline int small_tile_rows(int d) { return d == 4 ? QCB_TILE4 : 1; }
QCB_HD static inline bool small_tiled(int d) { return small_tile_rows(d) > 1; }
QCB_HD static inline int small_ld(int d) { return small_tiled(d) ? d : ((d & 1) ? d : d + 1); }

// M_ / TPU_ > 0: drives and threads per unit known at compile time (the instances of the BASELINE shapes):
// loops over drives unroll and the index arithmetic folds; 0: read from the launch parameters.
template <int M_, int TPU_> struct SmallCtxT {
    int d, ld, msz, m_, level, j;
    int e0, tpu_;    // this thread's first element and the element stride
    double *fixed;   // [QR_NUM_FIXED][msz]
    double *E, *W;   // [smax+1][msz]
    double *dpersist;  // [QCB_NUM_DPERSIST][m][msz]: DA2, DA3, DA4, DZ1, DZ2 of every direction
    double *L;       // [2][m][msz]: L_j at the current / next squaring level
    int lcur;        // which half of L holds the current level
    const double *Gs;  // [m][msz] the unit's drive generators, staged in shared memory
    __device__ __forceinline__ int m() const { return M_ > 0 ? M_ : m_; }
    __device__ __forceinline__ int tpu() const { return TPU_ > 0 ? TPU_ : tpu_; }
};

template <class Ctx>
__device__ __forceinline__ double *sm_resolve(const Ctx &x, int r) {
    if (r < QR_NUM_FIXED) return x.fixed + r * x.msz;
    switch (r) {
    case QR_EL: return x.E + x.level * x.msz;
    case QR_EL1: return x.E + (x.level + 1) * x.msz;
    case QR_WL: return x.W + x.level * x.msz;
    case QR_WL1: return x.W + (x.level + 1) * x.msz;
    case QR_E0: return x.E;
    case QR_W0: return x.W;
    case QR_GJ: return const_cast<double *>(x.Gs) + x.j * x.msz;
    case QR_L: return x.L + (x.lcur * x.m() + x.j) * x.msz;
    case QR_LN: return x.L + ((x.lcur ^ 1) * x.m() + x.j) * x.msz;
    default: return x.dpersist + (qcb_dpersist(r) * x.m() + x.j) * x.msz;   // QR_U is never stored here
    }
}

// Barrier among the threads of one unit.  A unit of at most 32 threads lives inside one warp (tpu is a
// power of two, units are aligned), so its producer/consumer steps need a warp barrier only; the
// CTA-wide barrier is kept for larger units and for the one CTA-level exchange (the squaring count).
__device__ __forceinline__ void unit_barrier(int tpu) {
    if (tpu <= 32) __syncwarp();
    else __syncthreads();
}

// ---- compile-time ops (D > 0): one specialised routine per table entry -- register bases are
//      constants, the k loops unroll with immediate offsets, Taylor coefficients are immediates ----
template <int D> struct SmDims {
    static constexpr int RB = (D == 4) ? QCB_TILE4 : 1;   // == small_tile_rows(D)
    static constexpr int ld = RB > 1 ? D : ((D & 1) ? D : D + 1);
    static constexpr int msz = ld * D;
    static constexpr int NT = (D / RB) * D;       // tiles per matrix (D = 0: unused)
};

// N doubles (N even) from 16-byte aligned shared memory: LDS.128
template <int N> __device__ __forceinline__ void lds_vec(double (&v)[N], const double *p) {
#pragma unroll
    for (int i = 0; i < N; i += 2) {
        const double2 t = *reinterpret_cast<const double2 *>(p + i);
        v[i] = t.x; v[i + 1] = t.y;
    }
}
template <int N> __device__ __forceinline__ void sts_vec(double *p, const double (&v)[N]) {
#pragma unroll
    for (int i = 0; i < N; i += 2) *reinterpret_cast<double2 *>(p + i) = make_double2(v[i], v[i + 1]);
}
// acc[i] += sum_k op(A)(r0 + i, k) * op(B)(k, q), i < RB: the column of op(B) is read once into registers,
// op(A) arrives RB rows at a time -- (RB + 1) / (2 RB) 16-byte loads per FMA instead of two 8-byte loads
template <int D, int RB, bool TA, bool TB>
__device__ __forceinline__ void tile_mac(double (&acc)[RB], const double *A, const double *B, int r0, int q) {
    double b[D];
    if (!TB) lds_vec<D>(b, B + q * D);
    else {
#pragma unroll
        for (int k = 0; k < D; ++k) b[k] = B[q + k * D];
    }
    if (!TA) {
#pragma unroll
        for (int k = 0; k < D; ++k) {
            double a[RB];
            lds_vec<RB>(a, A + r0 + k * D);
#pragma unroll
            for (int i = 0; i < RB; ++i) acc[i] = fma(a[i], b[k], acc[i]);
        }
    } else {
#pragma unroll
        for (int i = 0; i < RB; ++i) {
            double a[D];
            lds_vec<D>(a, A + (r0 + i) * D);
            double p0 = 0.0, p1 = 0.0;      // two chains per row: RB rows x 2 keep the FP64 pipe fed
#pragma unroll
            for (int k = 0; k + 1 < D; k += 2) { p0 = fma(a[k], b[k], p0); p1 = fma(a[k + 1], b[k + 1], p1); }
            acc[i] += p0 + p1;
        }
    }
}
template <int D, int AS, int BS> __device__ __forceinline__ double dot_k(const double *a, const double *b, double acc, int d, int ld) {
    // AS / BS: 1 = unit stride, 0 = stride ld
    if (D > 0) {
        constexpr int LD = SmDims<D>::ld;
        // four independent partial sums: the FP64 FMA latency chain would otherwise bound an op
        double p0 = 0.0, p1 = 0.0, p2 = 0.0, p3 = 0.0;
#pragma unroll
        for (int k = 0; k + 3 < D; k += 4) {
            p0 = fma(a[(k + 0) * (AS ? 1 : LD)], b[(k + 0) * (BS ? 1 : LD)], p0);
            p1 = fma(a[(k + 1) * (AS ? 1 : LD)], b[(k + 1) * (BS ? 1 : LD)], p1);
            p2 = fma(a[(k + 2) * (AS ? 1 : LD)], b[(k + 2) * (BS ? 1 : LD)], p2);
            p3 = fma(a[(k + 3) * (AS ? 1 : LD)], b[(k + 3) * (BS ? 1 : LD)], p3);
        }
#pragma unroll
        for (int k = D & ~3; k < D; ++k) p0 = fma(a[k * (AS ? 1 : LD)], b[k * (BS ? 1 : LD)], p0);
        acc += (p0 + p1) + (p2 + p3);
    } else {
        const int as = AS ? 1 : ld, bs = BS ? 1 : ld;
        for (int k = 0; k < d; ++k) acc = fma(a[k * as], b[k * bs], acc);
    }
    return acc;
}

// dst = sum_t op(a)op(b) + sum_r c*reg + c*I for this thread's unit.  Ends with a CTA barrier.
template <int D, bool SYNC = true, class Ctx> __device__ __noinline__ void sm_run_op(const Ctx &x, const qcb_op &op) {
    const int d = D > 0 ? D : x.d, ld = D > 0 ? SmDims<D>::ld : x.ld, dd = d * d;
    double *dst = sm_resolve(x, op.dst);
    for (int e = x.e0; e < dd; e += x.tpu()) {
        const int r = e % d, q = e / d;
        double acc = 0.0;
        for (int t = 0; t < op.nterm; ++t) {
            const int ta = op.term[t].ta, tb = op.term[t].tb;
            const double *a = sm_resolve(x, op.term[t].a) + (ta ? r * ld : r);   // a(r,k): stride ld, or 1 if transposed
            const double *b = sm_resolve(x, op.term[t].b) + (tb ? q : q * ld);   // b(k,q): stride 1, or ld if transposed
            if (!ta && !tb) acc = dot_k<D, 0, 1>(a, b, acc, d, ld);
            else if (ta && !tb) acc = dot_k<D, 1, 1>(a, b, acc, d, ld);
            else if (!ta && tb) acc = dot_k<D, 0, 0>(a, b, acc, d, ld);
            else acc = dot_k<D, 1, 0>(a, b, acc, d, ld);
        }
        const int off = r + q * ld;
        for (int l = 0; l < op.nlin; ++l) acc = fma(c_taylor[op.lin[l].coef], sm_resolve(x, op.lin[l].reg)[off], acc);
        if (op.ident >= 0 && r == q) acc += c_taylor[op.ident];
        dst[off] = acc;
    }
    if (SYNC) unit_barrier(x.tpu());
}

template <int D, int R, class Ctx> __device__ __forceinline__ double *sm_resolve_s(const Ctx &x) {
    constexpr int MSZ = SmDims<D>::msz;
    if constexpr (R < QR_NUM_FIXED) return x.fixed + R * MSZ;
    else if constexpr (R == QR_EL) return x.E + x.level * MSZ;
    else if constexpr (R == QR_EL1) return x.E + (x.level + 1) * MSZ;
    else if constexpr (R == QR_WL) return x.W + x.level * MSZ;
    else if constexpr (R == QR_WL1) return x.W + (x.level + 1) * MSZ;
    else if constexpr (R == QR_E0) return x.E;
    else if constexpr (R == QR_W0) return x.W;
    else if constexpr (R == QR_GJ) return const_cast<double *>(x.Gs) + x.j * MSZ;
    else if constexpr (R == QR_L) return x.L + (x.lcur * x.m() + x.j) * MSZ;
    else if constexpr (R == QR_LN) return x.L + ((x.lcur ^ 1) * x.m() + x.j) * MSZ;
    else return x.dpersist + ((R - QR_DA2) * x.m() + x.j) * MSZ;
}
template <int D, const qcb_op *TAB, int I, int T, class Ctx>
__device__ __forceinline__ double sm_term_s(const Ctx &x, int r, int q) {
    constexpr qcb_term tm = TAB[I].term[T];
    constexpr int LD = SmDims<D>::ld;
    const double *a = sm_resolve_s<D, tm.a>(x) + (tm.ta ? r * LD : r);
    const double *b = sm_resolve_s<D, tm.b>(x) + (tm.tb ? q : q * LD);
    return dot_k<D, (tm.ta ? 1 : 0), (tm.tb ? 0 : 1)>(a, b, 0.0, D, LD);
}
template <int D, const qcb_op *TAB, int I, int T, class Ctx>
__device__ __forceinline__ void sm_term_tile(const Ctx &x, int r0, int q, double (&acc)[SmDims<D>::RB]) {
    constexpr qcb_term tm = TAB[I].term[T];
    tile_mac<D, SmDims<D>::RB, (tm.ta != 0), (tm.tb != 0)>(acc, sm_resolve_s<D, tm.a>(x), sm_resolve_s<D, tm.b>(x), r0, q);
}
template <int D, const qcb_op *TAB, int I, int L, class Ctx>
__device__ __forceinline__ void sm_lin_tile(const Ctx &x, int off, double (&acc)[SmDims<D>::RB]) {
    constexpr qcb_lin ln = TAB[I].lin[L];
    constexpr double cf = QCB_TAYLOR[ln.coef];
    double v[SmDims<D>::RB];
    lds_vec<SmDims<D>::RB>(v, sm_resolve_s<D, ln.reg>(x) + off);
#pragma unroll
    for (int i = 0; i < SmDims<D>::RB; ++i) acc[i] = fma(cf, v[i], acc[i]);
}
template <int D, const qcb_op *TAB, int I, int L, class Ctx>
__device__ __forceinline__ double sm_lin_s(const Ctx &x, int off) {
    constexpr qcb_lin ln = TAB[I].lin[L];
    constexpr double cf = QCB_TAYLOR[ln.coef];
    return cf * sm_resolve_s<D, ln.reg>(x)[off];
}
template <int D, const qcb_op *TAB, int I, bool SYNC, class Ctx, int... T, int... L>
__device__ __forceinline__ void sm_run_static_impl(const Ctx &x, std::integer_sequence<int, T...>,
                                                   std::integer_sequence<int, L...>) {
    constexpr qcb_op op = TAB[I];
    double *dst = sm_resolve_s<D, op.dst>(x);
    if constexpr (SmDims<D>::RB > 1) {          // register tiles: RB rows of one column per thread
        constexpr int RB = SmDims<D>::RB, NT = SmDims<D>::NT, NRB = D / RB;
        for (int t = x.e0; t < NT; t += x.tpu()) {
            const int q = t / NRB, r0 = (t - q * NRB) * RB, off = r0 + q * D;
            double acc[RB];
#pragma unroll
            for (int i = 0; i < RB; ++i) acc[i] = 0.0;
            (sm_term_tile<D, TAB, I, T>(x, r0, q, acc), ...);
            (sm_lin_tile<D, TAB, I, L>(x, off, acc), ...);
            if constexpr (op.ident >= 0) {
                constexpr double ci = QCB_TAYLOR[op.ident >= 0 ? op.ident : 0];
#pragma unroll
                for (int i = 0; i < RB; ++i) if (r0 + i == q) acc[i] += ci;
            }
            sts_vec<RB>(dst + off, acc);
        }
    } else {                                    // one element per thread
        constexpr int LD = SmDims<D>::ld;
        for (int e = x.e0; e < D * D; e += x.tpu()) {
            const int r = e % D, q = e / D, off = r + q * LD;
            double acc = (0.0 + ... + sm_term_s<D, TAB, I, T>(x, r, q));
            acc += (0.0 + ... + sm_lin_s<D, TAB, I, L>(x, off));
            if constexpr (op.ident >= 0) {
                constexpr double ci = QCB_TAYLOR[op.ident >= 0 ? op.ident : 0];
                if (r == q) acc += ci;
            }
            dst[off] = acc;
        }
    }
    if (SYNC) unit_barrier(x.tpu());
}
template <int D, const qcb_op *TAB, int I, bool SYNC = true, class Ctx> __device__ __forceinline__ void sm_run_static(const Ctx &x) {
    sm_run_static_impl<D, TAB, I, SYNC>(x, std::make_integer_sequence<int, TAB[I].nterm>{},
                                  std::make_integer_sequence<int, TAB[I].nlin>{});
}
template <int D, const qcb_op *TAB, class Ctx, int... I>
__device__ __forceinline__ void sm_run_table_impl(const Ctx &x, std::integer_sequence<int, I...>) {
    (sm_run_static<D, TAB, I>(x), ...);
}
// run ops 0..N-1 of a table: specialised code for D > 0, the interpreter otherwise
template <int D, const qcb_op *TAB, int N, class Ctx> __device__ __forceinline__ void sm_run_table(const Ctx &x, const qcb_op *ctab) {
    if constexpr (D > 0) sm_run_table_impl<D, TAB>(x, std::make_integer_sequence<int, N>{});
    else for (int i = 0; i < N; ++i) sm_run_op<0>(x, ctab[i]);
}
template <int D, const qcb_op *TAB, int I, bool SYNC = true, class Ctx> __device__ __forceinline__ void sm_run_one(const Ctx &x, const qcb_op *ctab) {
    if constexpr (D > 0) sm_run_static<D, TAB, I, SYNC>(x);
    else sm_run_op<0, SYNC>(x, ctab[I]);
}

// value of the product terms of op I at element (r, q) -- no store (fused contractions); run-time d only
template <class Ctx> __device__ __forceinline__ double sm_products_rt(const Ctx &x, const qcb_op &op, int r, int q) {
    const int d = x.d, ld = x.ld;
    double acc = 0.0;
    for (int t = 0; t < op.nterm; ++t) {
        const int ta = op.term[t].ta, tb = op.term[t].tb;
        const double *a = sm_resolve(x, op.term[t].a) + (ta ? r * ld : r);
        const double *b = sm_resolve(x, op.term[t].b) + (tb ? q : q * ld);
        const int as = ta ? 1 : ld, bs = tb ? ld : 1;
        for (int k = 0; k < d; ++k) acc = fma(a[k * as], b[k * bs], acc);
    }
    return acc;
}
// compile-time d, one element
template <int D, const qcb_op *TAB, int I, class Ctx, int... T>
__device__ __forceinline__ double sm_products_s(const Ctx &x, int r, int q, std::integer_sequence<int, T...>) {
    return (0.0 + ... + sm_term_s<D, TAB, I, T>(x, r, q));
}
template <int D, const qcb_op *TAB, int I, class Ctx>
__device__ __forceinline__ double sm_products_e(const Ctx &x, int r, int q) {
    return sm_products_s<D, TAB, I>(x, r, q, std::make_integer_sequence<int, TAB[I].nterm>{});
}
// the same for a register tile (compile-time d): u[i] = products of op I at rows r0 + i of column q
template <int D, const qcb_op *TAB, int I, class Ctx, int... T>
__device__ __forceinline__ void sm_products_tile_impl(const Ctx &x, int r0, int q, double (&u)[SmDims<D>::RB],
                                                      std::integer_sequence<int, T...>) {
#pragma unroll
    for (int i = 0; i < SmDims<D>::RB; ++i) u[i] = 0.0;
    (sm_term_tile<D, TAB, I, T>(x, r0, q, u), ...);
}
template <int D, const qcb_op *TAB, int I, class Ctx>
__device__ __forceinline__ void sm_products_tile(const Ctx &x, int r0, int q, double (&u)[SmDims<D>::RB]) {
    sm_products_tile_impl<D, TAB, I>(x, r0, q, u, std::make_integer_sequence<int, TAB[I].nterm>{});
}

// Sum over the threads of a unit, result in every thread of the unit.  tpu is a power of two <= 32
// (aligned sub-warp groups) or a multiple of 32; `red` has room for 16 doubles per unit.
__device__ __forceinline__ double unit_sum(double v, int tpu, double *red_unit, int lane_in_unit) {
    if (tpu <= 32) {
        for (int o = tpu >> 1; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        return v;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((lane_in_unit & 31) == 0) red_unit[lane_in_unit >> 5] = v;
    __syncthreads();
    double s = 0.0;
    for (int w = 0; w < (tpu >> 5); ++w) s += red_unit[w];
    __syncthreads();
    return s;
}
__device__ __forceinline__ double unit_max(double v, int tpu, double *red_unit, int lane_in_unit) {
    if (tpu <= 32) {
        for (int o = tpu >> 1; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
        return v;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    if ((lane_in_unit & 31) == 0) red_unit[lane_in_unit >> 5] = v;
    __syncthreads();
    double s = 0.0;
    for (int w = 0; w < (tpu >> 5); ++w) s = fmax(s, red_unit[w]);
    __syncthreads();
    return s;
}

// Frobenius inner product of two d x d matrices (leading dimension ld) over a unit
__device__ __noinline__ double sm_dot(const double *a, const double *b, int d, int ld, int e0, int tpu, double *red_unit) {
    double v = 0.0;
    for (int e = e0; e < d * d; e += tpu) {
        const int off = (e % d) + (e / d) * ld;
        v = fma(a[off], b[off], v);
    }
    return unit_sum(v, tpu, red_unit, e0);
}

// out[i] += <U, P_i> for i < m in one pass and one reduction (P_i = base + slot(i) * msz with
// slot(i) = idx ? idx[i] : i).  `red_unit` needs 4 doubles per warp of the unit.
__device__ __noinline__ void sm_dot_multi(const double *U, const double *base, int msz, const int *idx, int m,
                                          int d, int ld, int e0, int tpu, double *red_unit, double *out) {
    for (int i0 = 0; i0 < m; i0 += 4) {
        const int nn = min(4, m - i0);
        const double *P[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) P[q] = base + (long long)((q < nn) ? (idx ? idx[i0 + q] : i0 + q) : (idx ? idx[i0] : i0)) * msz;
        double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
        for (int e = e0; e < d * d; e += tpu) {
            const int off = (e % d) + (e / d) * ld;
            const double u = U[off];
            a0 = fma(u, P[0][off], a0); a1 = fma(u, P[1][off], a1);
            a2 = fma(u, P[2][off], a2); a3 = fma(u, P[3][off], a3);
        }
        if (tpu <= 32) {
            for (int o = tpu >> 1; o > 0; o >>= 1) {
                a0 += __shfl_xor_sync(0xffffffffu, a0, o); a1 += __shfl_xor_sync(0xffffffffu, a1, o);
                a2 += __shfl_xor_sync(0xffffffffu, a2, o); a3 += __shfl_xor_sync(0xffffffffu, a3, o);
            }
            if (e0 == 0) {
                out[i0] += a0;
                if (nn > 1) out[i0 + 1] += a1;
                if (nn > 2) out[i0 + 2] += a2;
                if (nn > 3) out[i0 + 3] += a3;
            }
        } else {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                a0 += __shfl_xor_sync(0xffffffffu, a0, o); a1 += __shfl_xor_sync(0xffffffffu, a1, o);
                a2 += __shfl_xor_sync(0xffffffffu, a2, o); a3 += __shfl_xor_sync(0xffffffffu, a3, o);
            }
            if ((e0 & 31) == 0) {
                double *r = red_unit + (e0 >> 5) * 4;
                r[0] = a0; r[1] = a1; r[2] = a2; r[3] = a3;
            }
            __syncthreads();
            if (e0 < nn) {
                double v = 0.0;
                for (int w = 0; w < (tpu >> 5); ++w) v += red_unit[w * 4 + e0];
                out[i0 + e0] += v;
            }
            __syncthreads();
        }
    }
}

// C (d x c, packed) = scale * op(A) (d x d, ld) * P (d x c, packed)
__device__ __noinline__ void sm_panel(double *C, const double *A, int ta, const double *P, int d, int ld, int c,
                                      double scale, int e0, int tpu) {
    for (int e = e0; e < d * c; e += tpu) {
        const int r = e % d, b = e / d;
        const double *a = A + (ta ? r * ld : r);
        const int as = ta ? 1 : ld;
        double acc = 0.0;
        for (int k = 0; k < d; ++k) acc = fma(a[k * as], P[k + b * d], acc);
        C[e] = scale * acc;
    }
}

// Add the unit-wide sums of a[0..nn) to out[0..nn).  red_unit: 8 doubles per warp of the unit.
__device__ __forceinline__ void unit_reduce_add8(double (&a)[8], int nn, int tpu, int e0, double *red_unit, double *out) {
    if (tpu <= 32) {
#pragma unroll
        for (int t = 0; t < 8; ++t)
            for (int o = tpu >> 1; o > 0; o >>= 1) a[t] += __shfl_xor_sync(0xffffffffu, a[t], o);
        if (e0 == 0)
#pragma unroll
            for (int t = 0; t < 8; ++t) if (t < nn) out[t] += a[t];
    } else {
#pragma unroll
        for (int t = 0; t < 8; ++t)
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) a[t] += __shfl_xor_sync(0xffffffffu, a[t], o);
        if ((e0 & 31) == 0)
#pragma unroll
            for (int t = 0; t < 8; ++t) red_unit[(e0 >> 5) * 8 + t] = a[t];
        __syncthreads();
        if (e0 < nn) {
            double v = 0.0;
            for (int w = 0; w < (tpu >> 5); ++w) v += red_unit[w * 8 + e0];
            out[e0] += v;
        }
        __syncthreads();
    }
}

// Second order at the polynomial level for direction j, all product nodes at once and nothing
// stored:  Hrow[i] += sum_nodes sum_e (Xdot_j^T Nbar)[e] * Ydot_i[e].
template <int D, class Ctx> __device__ __noinline__ void sm_p6_pass(Ctx &x, int j, const qcb_op *ctab, double *red_unit, double *Hrow) {
    const int d = D > 0 ? D : x.d, ld = D > 0 ? SmDims<D>::ld : x.ld, msz = ld * d, m = x.m();
    constexpr int pr0 = QCB_P6_PARTNER[0], pr1 = QCB_P6_PARTNER[1], pr2 = QCB_P6_PARTNER[2], pr3 = QCB_P6_PARTNER[3];
    auto partner_base = [&](int reg) {
        return (reg == QR_GJ) ? x.Gs : x.dpersist + (long long)qcb_dpersist(reg) * m * msz;
    };
    const double *part[4] = {partner_base(pr0), partner_base(pr1), partner_base(pr2), partner_base(pr3)};
    x.j = j; x.level = 0;
    for (int i0 = 0; i0 < m; i0 += 8) {
        const int nn = min(8, m - i0);
        double a[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if constexpr (D > 0 && SmDims<D>::RB > 1) {
            constexpr int RB = SmDims<D>::RB, NT = SmDims<D>::NT, NRB = D / RB;
            for (int t = x.e0; t < NT; t += x.tpu()) {
                const int q = t / NRB, r0 = (t - q * NRB) * RB, off = r0 + q * D;
                double u0[RB], u1[RB], u2[RB], u3[RB];
                sm_products_tile<D, QCB_P6, 0>(x, r0, q, u0);
                sm_products_tile<D, QCB_P6, 1>(x, r0, q, u1);
                sm_products_tile<D, QCB_P6, 2>(x, r0, q, u2);
                sm_products_tile<D, QCB_P6, 3>(x, r0, q, u3);
#pragma unroll
                for (int tt = 0; tt < 8; ++tt)
                    if (tt < nn) {
                        const long long po = (long long)(i0 + tt) * msz + off;
                        double pv[RB], acc2 = 0.0;       // one partner at a time: registers, not bandwidth, are short here
                        lds_vec<RB>(pv, part[0] + po);
#pragma unroll
                        for (int i = 0; i < RB; ++i) a[tt] = fma(u0[i], pv[i], a[tt]);
                        lds_vec<RB>(pv, part[1] + po);
#pragma unroll
                        for (int i = 0; i < RB; ++i) acc2 = fma(u1[i], pv[i], acc2);
                        lds_vec<RB>(pv, part[2] + po);
#pragma unroll
                        for (int i = 0; i < RB; ++i) a[tt] = fma(u2[i], pv[i], a[tt]);
                        lds_vec<RB>(pv, part[3] + po);
#pragma unroll
                        for (int i = 0; i < RB; ++i) acc2 = fma(u3[i], pv[i], acc2);
                        a[tt] += acc2;
                    }
            }
        } else {
            for (int e = x.e0; e < d * d; e += x.tpu()) {
                const int r = e % d, q = e / d, off = r + q * ld;
                double u0, u1, u2, u3;
                if constexpr (D > 0) {
                    u0 = sm_products_e<D, QCB_P6, 0>(x, r, q); u1 = sm_products_e<D, QCB_P6, 1>(x, r, q);
                    u2 = sm_products_e<D, QCB_P6, 2>(x, r, q); u3 = sm_products_e<D, QCB_P6, 3>(x, r, q);
                } else {
                    u0 = sm_products_rt(x, ctab[0], r, q); u1 = sm_products_rt(x, ctab[1], r, q);
                    u2 = sm_products_rt(x, ctab[2], r, q); u3 = sm_products_rt(x, ctab[3], r, q);
                }
#pragma unroll
                for (int t = 0; t < 8; ++t)
                    if (t < nn) {
                        const long long po = (long long)(i0 + t) * msz + off;
                        a[t] = fma(u0, part[0][po], fma(u1, part[1][po], fma(u2, part[2][po], fma(u3, part[3][po], a[t]))));
                    }
            }
        }
        unit_reduce_add8(a, nn, x.tpu(), x.e0, red_unit, Hrow + i0);
    }
}

// One squaring level for direction j:  Hrow[i] += sum_e (L_j^T W[l+1])[e] * L_i[e]
template <int D, class Ctx> __device__ __noinline__ void sm_level_pass(Ctx &x, int j, int level, const qcb_op *ctab, double *red_unit, double *Hrow) {
    const int d = D > 0 ? D : x.d, ld = D > 0 ? SmDims<D>::ld : x.ld, msz = ld * d, m = x.m();
    const double *Lbase = x.L + (long long)x.lcur * m * msz;
    x.j = j; x.level = level;
    for (int i0 = 0; i0 < m; i0 += 8) {
        const int nn = min(8, m - i0);
        double a[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if constexpr (D > 0 && SmDims<D>::RB > 1) {
            constexpr int RB = SmDims<D>::RB, NT = SmDims<D>::NT, NRB = D / RB;
            for (int t = x.e0; t < NT; t += x.tpu()) {
                const int q = t / NRB, r0 = (t - q * NRB) * RB, off = r0 + q * D;
                double u[RB];
                sm_products_tile<D, QCB_P7U, 0>(x, r0, q, u);
#pragma unroll
                for (int tt = 0; tt < 8; ++tt)
                    if (tt < nn) {
                        double l[RB];
                        lds_vec<RB>(l, Lbase + (long long)(i0 + tt) * msz + off);
#pragma unroll
                        for (int i = 0; i < RB; ++i) a[tt] = fma(u[i], l[i], a[tt]);
                    }
            }
        } else {
            for (int e = x.e0; e < d * d; e += x.tpu()) {
                const int r = e % d, q = e / d, off = r + q * ld;
                double u;
                if constexpr (D > 0) u = sm_products_e<D, QCB_P7U, 0>(x, r, q);
                else u = sm_products_rt(x, ctab[0], r, q);
#pragma unroll
                for (int t = 0; t < 8; ++t)
                    if (t < nn) a[t] = fma(u, Lbase[(long long)(i0 + t) * msz + off], a[t]);
            }
        }
        unit_reduce_add8(a, nn, x.tpu(), x.e0, red_unit, Hrow + i0);
    }
}

static inline size_t small_unit_doubles(int d, int c, int m, int smax) {
    size_t msz = (size_t)small_ld(d) * d, dd = (size_t)d * d, dc = (size_t)d * c;
    size_t n = (QR_NUM_FIXED + 2 * (size_t)(smax + 1) + (size_t)QCB_NUM_DPERSIST * m + 2 * (size_t)m + m) * msz;
    n += 3 * dc;                                   // X, Xn, M panels
    n += dd + (m + 1) * dc + 2;                    // Jacobian arena
    n += (m + 1) * dc + (size_t)m * m + m + 2;     // Hessian arena
    const size_t tiles = (size_t)(d / small_tile_rows(d)) * d;
    n += (size_t)m * m + (tiles <= 32 ? 0 : 128);  // H accumulator + reduction scratch (8 per warp; units of
                                                   // at most 32 threads reduce with shuffles and need none)
    return (n + 1) & ~(size_t)1;                   // keep every unit 16-byte aligned
}
static inline size_t small_smem_doubles(int d, int c, int m, int smax, int upc = 1) {
    return (size_t)upc * small_unit_doubles(d, c, m, smax) + (size_t)upc + 2;
}

// At most 512 threads per CTA (host: upc * tpu <= 512), so the compiler may use up to 128 registers.
// Without the bound the run-time-d instance took 167 and could not be launched with 416..512 threads
// (d = 20..23).  Capping D = 4 at 64 or 85 registers for more resident warps was measured slower
// (spills: 0.60 / 0.57 ms against 0.51 ms on cfg4).
// The compile-time instances run at most 256 threads per CTA (register tiles: a unit needs few threads,
// and the tiles want more than 128 registers).
// out[e] = +-arena[code[e]] for the entries e0, e0 + tpu, ... of one unit
__device__ __forceinline__ void coo_store(const uint32_t *__restrict__ code, const double *arena, double *__restrict__ out,
                                          long long nnz, int e0, int tpu) {
    constexpr int U = 8;
    long long e = e0;
    for (; e + (long long)(U - 1) * tpu < nnz; e += (long long)U * tpu) {
        uint32_t cc[U];
#pragma unroll
        for (int u = 0; u < U; ++u) cc[u] = __ldg(code + e + (long long)u * tpu);
        double v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) v[u] = arena[cc[u] & 0x7fffffffu];
#pragma unroll
        for (int u = 0; u < U; ++u) out[e + (long long)u * tpu] = (cc[u] >> 31) ? -v[u] : v[u];
    }
    for (; e < nnz; e += tpu) {
        const uint32_t cc = __ldg(code + e);
        const double v = arena[cc & 0x7fffffffu];
        out[e] = (cc >> 31) ? -v : v;
    }
}

// C, M, TPU > 0: state columns, drives and threads per unit fixed at compile time (the BASELINE shapes
// d/c/m = 4/2/2, 6/3/2, 18/9/4, 18/1/4): ncu on the generic d = 4 instance showed 60 % of the executed
// instructions to be integer arithmetic and branches of run-time loop bounds, 8 % DFMA.
template <int D, int C = 0, int M = 0, int TPU = 0>
__global__ void __launch_bounds__(SmDims<D>::RB > 1 ? 256 : 512, 1)
small_unit_kernel(const __grid_constant__ SmallParams p) {
    extern __shared__ __align__(16) double sm[];
    const int d = D > 0 ? D : p.d, c = C > 0 ? C : p.c, m = M > 0 ? M : p.m, dd = d * d, dc = d * c;
    const int ld = D > 0 ? SmDims<D>::ld : small_ld(d), msz = ld * d;
    const int tpu = TPU > 0 ? TPU : p.tpu, ul = threadIdx.x / tpu, e0 = threadIdx.x % tpu;
    int ug = blockIdx.x * p.upc + ul;                 // slab-local unit index = sys*nk + kl
    const bool live = ug < p.units;                   // tail CTA: spare slots shadow the last unit
    if (!live) ug = p.units - 1;
    const int sys = ug / p.nk, k = p.k_begin + ug % p.nk;
    const double *zk = p.Z + (long long)k * p.z_dim, *zn = zk + p.z_dim;
    const double *G0 = p.G0 + (long long)sys * dd, *Gd = p.Gd + (long long)sys * m * dd;

    SmallCtxT<M, TPU> x;
    x.d = d; x.ld = ld; x.msz = msz; x.m_ = m; x.level = 0; x.j = 0; x.e0 = e0; x.tpu_ = tpu;
    double *ptr = sm + (size_t)ul * p.unit_doubles;
    x.fixed = ptr; ptr += QR_NUM_FIXED * msz;
    x.E = ptr; ptr += (p.smax + 1) * msz;
    x.W = ptr; ptr += (p.smax + 1) * msz;
    x.dpersist = ptr; ptr += QCB_NUM_DPERSIST * m * msz;
    x.L = ptr; ptr += 2 * m * msz;
    x.lcur = 0;
    double *Gs = ptr; ptr += m * msz;
    double *X = ptr; ptr += dc;
    double *Xn = ptr; ptr += dc;
    double *Mu = ptr; ptr += dc;
    double *JA = ptr; ptr += dd + (m + 1) * dc + 2;
    double *HA = ptr; ptr += (m + 1) * dc + m * m + m + 2;
    double *Hacc = ptr; ptr += m * m;
    double *red = ptr;
    double *cta_tail = sm + (size_t)p.upc * p.unit_doubles;
    double *cta_nrm = cta_tail;                                       // [upc]
    x.Gs = Gs;

    const bool want_jac = p.want & 1, want_hess = (p.want & 2) && p.mu;
    double *G = x.fixed + QR_G * msz, *A = x.fixed + QR_A * msz;
    const double dt = zk[p.dt_off];

    // ---- load unit inputs; stage G_j in shared memory; G = G0 + sum u_j G_j.  Loads are issued in batches
    //      ahead of the shared-memory stores that consume them (a unit has few warps to hide them behind) ----
    {
        const double *xs = zk + p.x_off[sys], *xn = zn + p.x_off[sys];
        const double *ms = want_hess ? p.mu + ((long long)sys * p.K + k) * dc : nullptr;
        constexpr int U = 4;
        int e = e0;
        for (; e + (U - 1) * tpu < dc; e += U * tpu) {
            double a[U], b[U], c2[U];
#pragma unroll
            for (int u = 0; u < U; ++u) { a[u] = __ldg(xs + e + u * tpu); b[u] = __ldg(xn + e + u * tpu); c2[u] = ms ? __ldg(ms + e + u * tpu) : 0.0; }
#pragma unroll
            for (int u = 0; u < U; ++u) { X[e + u * tpu] = a[u]; Xn[e + u * tpu] = b[u]; Mu[e + u * tpu] = c2[u]; }
        }
        for (; e < dc; e += tpu) { X[e] = __ldg(xs + e); Xn[e] = __ldg(xn + e); Mu[e] = ms ? __ldg(ms + e) : 0.0; }
    }
    for (int e = e0; e < dd; e += tpu) {
        const int off = (e % d) + (e / d) * ld;
        double v = __ldg(G0 + e);
        if constexpr (M > 0) {
            double gj[M];
#pragma unroll
            for (int j = 0; j < M; ++j) gj[j] = __ldg(Gd + j * dd + e);
#pragma unroll
            for (int j = 0; j < M; ++j) { Gs[j * msz + off] = gj[j]; v = fma(zk[p.u_off + j], gj[j], v); }
        } else {
            for (int j = 0; j < m; ++j) {
                const double gj = __ldg(Gd + j * dd + e);
                Gs[j * msz + off] = gj;
                v = fma(zk[p.u_off + j], gj, v);
            }
        }
        G[off] = v;
    }
    for (int e = e0; e < m * m; e += tpu) Hacc[e] = 0.0;
    unit_barrier(tpu);
    // ---- ||dt G||_1 per unit, then one squaring count for the CTA ----
    {
        double mine = 0.0;
        bool bad = false;
        for (int q = e0; q < d; q += tpu) {
            double cs = 0.0;
            for (int r = 0; r < d; ++r) cs += fabs(dt * G[r + q * ld]);
            if (!isfinite(cs)) bad = true;
            mine = fmax(mine, cs);
        }
        if (bad || !isfinite(dt)) mine = INFINITY;
        mine = unit_max(mine, tpu, red, e0);
        if (e0 == 0) cta_nrm[ul] = mine;
        __syncthreads();
    }
    double nrm = 0.0;
    for (int u = 0; u < p.upc; ++u) nrm = fmax(nrm, cta_nrm[u]);
    const int s = qcb_squarings(nrm);
    if (s < 0 || s > p.smax) {      // uniform across the CTA
        // bad_flag[0]: bit0 = non-finite / out of range, bit1 = the chains of this launch are too
        // short; bad_flag[1] = squarings needed (the host re-launches with that capacity)
        if (threadIdx.x == 0) {
            atomicOr(p.bad_flag, s < 0 ? 1 : 2);
            if (s >= 0) atomicMax(p.bad_flag + 1, s);
        }
        return;
    }
    if (threadIdx.x == 0 && s > 0) atomicMax(p.bad_flag + 1, s);
    const double tau = ldexp(dt, -s);
    for (int e = e0; e < dd; e += tpu) { const int off = (e % d) + (e / d) * ld; A[off] = tau * G[off]; }
    unit_barrier(tpu);

    const bool tang = want_jac || want_hess;
    double *Wt = x.W + s * msz;
    if (want_hess) {
        for (int e = e0; e < dd; e += tpu) {   // W = M X^T (independent of E: ready before it is needed)
            const int r = e % d, q = e / d;
            double acc = 0.0;
            for (int b = 0; b < c; ++b) acc = fma(Mu[r + b * d], X[q + b * d], acc);
            Wt[r + q * ld] = acc;
        }
    }
    // ---- polynomial: the five dependency levels of the Paterson-Stockmeyer scheme; the forward
    //      tangents of ALL drive directions advance in the same barrier steps ----
    x.level = 0;
    sm_run_one<D, QCB_P1, 0, false>(x, c_P1);                                           // A2
    if (tang) for (int j = 0; j < m; ++j) { x.j = j; sm_run_one<D, QCB_P5, 0, false>(x, c_P5); }   // dA2_j
    unit_barrier(tpu);
    sm_run_one<D, QCB_P1, 1, false>(x, c_P1);                                           // A3
    sm_run_one<D, QCB_P1, 2, false>(x, c_P1);                                           // A4
    if (tang) for (int j = 0; j < m; ++j) {
        x.j = j;
        sm_run_one<D, QCB_P5, 1, false>(x, c_P5);                                       // dA3_j
        sm_run_one<D, QCB_P5, 2, false>(x, c_P5);                                       // dA4_j
    }
    unit_barrier(tpu);
    sm_run_one<D, QCB_P1, 3, false>(x, c_P1);                                           // Z2
    if (tang) for (int j = 0; j < m; ++j) { x.j = j; sm_run_one<D, QCB_P5, 3, false>(x, c_P5); }   // dZ2_j
    unit_barrier(tpu);
    sm_run_one<D, QCB_P1, 4, false>(x, c_P1);                                           // Z1
    if (tang) for (int j = 0; j < m; ++j) { x.j = j; sm_run_one<D, QCB_P5, 4, false>(x, c_P5); }   // dZ1_j
    unit_barrier(tpu);
    sm_run_one<D, QCB_P1, 5, false>(x, c_P1);                                           // E[0]
    if (tang) for (int j = 0; j < m; ++j) { x.j = j; sm_run_one<D, QCB_P5, 5, false>(x, c_P5); }   // L_j
    unit_barrier(tpu);
    // ---- squarings (Jacobian-only calls advance L_j in the same steps) ----
    for (int l = 0; l < s; ++l) {
        x.level = l;
        sm_run_one<D, QCB_P2, 0, false>(x, c_P2);
        if (tang && !want_hess) for (int j = 0; j < m; ++j) { x.j = j; sm_run_one<D, QCB_P7L, 0, false>(x, c_P7L); }
        unit_barrier(tpu);
        if (tang && !want_hess) x.lcur ^= 1;
    }
    const double *E = x.E + s * msz;

    // ---- defect ----
    if (live) {
        double *o = p.delta + (long long)ug * dc;
        for (int e = e0; e < dc; e += tpu) {
            const int r = e % d, b = e / d;
            double acc = 0.0;
            for (int q = 0; q < d; ++q) acc = fma(E[r + q * ld], X[q + b * d], acc);
            o[e] = Xn[e] - acc;
        }
    }
    if (p.E_out && live) {
        double *o = p.E_out + (long long)ug * dd;
        for (int e = e0; e < dd; e += tpu) o[e] = E[(e % d) + (e / d) * ld];
    }
    if (!tang) return;

    if (want_hess) {
        // adjoint weight chain, reverse sweep of the polynomial (AB4 and AB3 share a step)
        for (int l = s - 1; l >= 0; --l) { x.level = l; sm_run_one<D, QCB_P3, 0>(x, c_P3); }
        x.level = 0;
        sm_run_one<D, QCB_P4, 0>(x, c_P4);                                              // ZB1
        sm_run_one<D, QCB_P4, 1>(x, c_P4);                                              // ZB2
        sm_run_one<D, QCB_P4, 2, false>(x, c_P4);                                       // AB4
        sm_run_one<D, QCB_P4, 3>(x, c_P4);                                              // AB3
        sm_run_one<D, QCB_P4, 4>(x, c_P4);                                              // AB2
        // second order at the polynomial level: fused product + contraction, nothing stored
        for (int j = 0; j < m; ++j) sm_p6_pass<D>(x, j, c_P6, red, Hacc + j * m);
        // squaring levels: contraction against W[l+1], then all L_j advance in one step
        for (int l = 0; l < s; ++l) {
            for (int j = 0; j < m; ++j) sm_level_pass<D>(x, j, l, c_P7U, red, Hacc + j * m);
            x.level = l;
            for (int j = 0; j < m; ++j) { x.j = j; sm_run_one<D, QCB_P7L, 0, false>(x, c_P7L); }
            unit_barrier(tpu);
            x.lcur ^= 1;
        }
    }
    const double *Lfin = x.L + (long long)x.lcur * m * msz;     // L_j of the last level, j-major
    // ---- final blocks ----
    double *GE = x.fixed + QR_A2 * msz, *WE = x.fixed + QR_A3 * msz, *GW = x.fixed + QR_A4 * msz;
    for (int e = e0; e < dd; e += tpu) {
        const int r = e % d, q = e / d;
        double g1 = 0.0, g2 = 0.0, g3 = 0.0;
        for (int kk = 0; kk < d; ++kk) {
            g1 = fma(G[r + kk * ld], E[kk + q * ld], g1);            // G E
            if (want_hess) {
                g2 = fma(Wt[r + kk * ld], E[q + kk * ld], g2);       // W E^T
                g3 = fma(G[kk + r * ld], Wt[kk + q * ld], g3);       // G^T W
            }
        }
        const int off = r + q * ld;
        GE[off] = g1; WE[off] = g2; GW[off] = g3;
        JA[e] = E[off];
    }
    unit_barrier(tpu);
    double *JU = JA + dd, *JT = JU + m * dc;
    for (int j = 0; j < m; ++j) sm_panel(JU + j * dc, Lfin + j * msz, 0, X, d, ld, c, -tau, e0, tpu);
    sm_panel(JT, GE, 0, X, d, ld, c, -1.0, e0, tpu);
    if (e0 == 0) { JT[dc] = 0.0; JT[dc + 1] = 1.0; }
    if (want_hess) {
        double *HXU = HA, *HXT = HA + m * dc, *HUU = HXT + dc, *HUT = HUU + m * m, *HTT = HUT + m;
        for (int j = 0; j < m; ++j) sm_panel(HXU + j * dc, Lfin + j * msz, 1, Mu, d, ld, c, -tau, e0, tpu);
        sm_panel(HXT, GE, 1, Mu, d, ld, c, -1.0, e0, tpu);
        unit_barrier(tpu);                                  // Hacc complete (written by e0 == 0 lanes)
        for (int e = e0; e < m * m; e += tpu) {
            const int i = e / m, j = e % m;
            HUU[e] = -tau * tau * (Hacc[i * m + j] + Hacc[j * m + i]);
        }
        for (int j = 0; j < m; ++j) {
            double c1 = sm_dot(WE, Gs + j * msz, d, ld, e0, tpu, red);
            double c2 = sm_dot(GW, Lfin + j * msz, d, ld, e0, tpu, red);
            if (e0 == 0) HUT[j] = -(c1 + tau * c2);
        }
        double c3 = sm_dot(GW, GE, d, ld, e0, tpu, red);
        if (e0 == 0) { HTT[0] = -c3; HTT[1] = 0.0; }
    }
    unit_barrier(tpu);
    if (!live) return;
    // ---- compact host transfer: hand over the arenas themselves ----
    if (want_jac && p.ja_out) {
        double *o = p.ja_out + (long long)ug * p.ja_stride;
        for (int e = e0; e < dd + (m + 1) * dc + 2; e += tpu) o[e] = JA[e];
    }
    if (want_hess && p.ha_out) {
        double *o = p.ha_out + (long long)ug * p.ha_stride;
        for (int e = e0; e < (m + 1) * dc + m * m + m + 2; e += tpu) o[e] = HA[e];
    }
    // ---- coalesced COO stores in the caller's order.  The slot codes come from global memory (L1 / L2): a
    //      batch of them is fetched before any is used, so that their latency is paid once per batch and not
    //      once per entry (ncu: a third of the d = 4 kernel's stall samples sat on these loads) ----
    if (want_jac && p.jac) coo_store(p.jcode + (long long)sys * p.jnnz, JA, p.jac + (long long)ug * p.jnnz, p.jnnz, e0, tpu);
    if (want_hess && p.hess) coo_store(p.hcode + (long long)sys * p.hnnz, HA, p.hess + (long long)ug * p.hnnz, p.hnnz, e0, tpu);
}

}  // namespace qcb
