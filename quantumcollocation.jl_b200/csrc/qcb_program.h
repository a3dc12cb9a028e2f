/* qcb_program.h -- the matrix program evaluated for every unit (knot x system).
 *
 * One unit = one interval k of one bilinear integrator
 * (reference: src/quantum_integrators.jl:28-45, DirectTrajOpt.BilinearIntegrator call sites).
 * Everything the evaluator needs from exp(dt*G(u)) -- the exponential itself, its first
 * Frechet derivatives along every drive generator, and the W-contracted second
 * derivatives -- is written here ONCE as a list of "matrix ops"
 *
 *     dst = sum_t op(a_t) * op(b_t)  +  sum_r c[coef_r] * reg_r  +  c[ident] * I
 *
 * over virtual d x d registers.  c[k] = 1/k! (degree-12 truncated Taylor, evaluated
 * Paterson-Stockmeyer style in blocks of 4), followed by s squarings.  The small-d CUDA
 * kernel interprets the table inside one launch with registers in shared memory; the
 * large-d path turns every op into one batched FP64 DMMA GEMM launch with registers in
 * HBM scratch; the CPU port under oracle/ interprets the same table on the host.
 *
 * Plain C so that all three translation units can include it.
 */
#ifndef QCB_PROGRAM_H
#define QCB_PROGRAM_H
#include <stdint.h>
#ifdef __CUDACC__
#define QCB_HD __host__ __device__
#else
#define QCB_HD
#endif

/* tables are compile-time constants for the C++ / CUDA translation units (the small-d kernel
 * instantiates one specialised routine per op), plain const data for the C port */
#ifdef __cplusplus
#define QCB_TABLE static constexpr
#else
#define QCB_TABLE static const
#endif

#define QCB_TAYLOR_DEGREE 12
/* ||A/2^s||_1 <= theta before the polynomial is applied.  Remainder bounds of the degree-12 polynomial
 * at theta = 0.5: value theta^13/13! = 2e-14, first derivative theta^12/12! = 5e-13, second derivative
 * theta^11/11! = 1.2e-11 -- all inside the 1e-10 parity bar; measured against scipy on cfg2/3/5 the
 * block-relative error stays at 1e-15 .. 3e-15 (scratch/theta_study.py).  theta = 0.25 cost one more
 * squaring level (3 products per unit and 3 per drive) for nothing measurable. */
#define QCB_THETA 0.5
#define QCB_MAX_SQUARINGS 10    /* ||dt*G(u)||_1 <= theta * 2^10 = 512; beyond: QCB_ERR_RANGE */
#define QCB_MAX_TERMS 6
#define QCB_MAX_LIN 4

/* ---- virtual registers ---- */
enum {
    /* per unit (shared by all drive directions) */
    QR_G = 0, QR_A, QR_A2, QR_A3, QR_A4, QR_Z1, QR_Z2,
    QR_ZB1, QR_ZB2, QR_AB4, QR_AB3, QR_AB2,
    QR_NUM_FIXED,                 /* = 12 */
    /* level-relative chain registers (resolved with the current squaring level l) */
    QR_EL = 16,                   /* E[l]   */
    QR_EL1,                       /* E[l+1] */
    QR_WL,                        /* Wlev[l]   */
    QR_WL1,                       /* Wlev[l+1] */
    QR_E0,                        /* E[0]   */
    QR_W0,                        /* Wlev[0] */
    /* system constant: generator of the current drive direction j */
    QR_GJ = 24,
    /* per (unit, direction) registers (see qcb_dslot / qcb_dpersist) */
    QR_DA2 = 32, QR_DA3, QR_DA4, QR_DZ1, QR_DZ2,
    QR_U,                         /* X_dot^T * N_bar of the current product node */
    QR_L = 48,                    /* L_j at the current level */
    QR_LN                         /* L_j at the next level */
};
#define QCB_NUM_DSLOTS 6          /* one slot per register when every direction has its own set */
#define QCB_NUM_DPERSIST 5        /* DA2, DA3, DA4, DZ1, DZ2: kept per direction (all directions advance together) */
#define QCB_NUM_DTEMP 1           /* U: dead once its contraction is done */

QCB_HD static inline int qcb_is_dreg(int r) { return r >= QR_DA2 && r <= QR_U; }
QCB_HD static inline int qcb_dslot(int r) { return r - QR_DA2; }
/* index among the persistent per-direction registers, or -1 */
QCB_HD static inline int qcb_dpersist(int r) {
    return (r >= QR_DA2 && r <= QR_DZ2) ? r - QR_DA2 : -1;
}
QCB_HD static inline int qcb_dtemp(int r) { (void)r; return 0; }

typedef struct { uint8_t a, b, ta, tb; } qcb_term;
typedef struct { uint8_t reg, coef; } qcb_lin;
typedef struct {
    uint8_t dst, nterm, nlin;
    int8_t ident;                 /* Taylor coefficient index multiplying I, or -1 */
    qcb_term term[QCB_MAX_TERMS];
    qcb_lin lin[QCB_MAX_LIN];
} qcb_op;

#define T_(a, b, ta, tb) { a, b, ta, tb }
#define NOT_ { 0, 0, 0, 0 }
#define NOL_ { 0, 0 }

/* Phase 1: polynomial, forward (per unit).  A = tau*G is prepared by the set-up step. */
QCB_TABLE qcb_op QCB_P1[] = {
    { QR_A2, 1, 0, -1, { T_(QR_A, QR_A, 0, 0), NOT_, NOT_, NOT_, NOT_, NOT_ }, { NOL_, NOL_, NOL_, NOL_ } },
    { QR_A3, 1, 0, -1, { T_(QR_A, QR_A2, 0, 0), NOT_, NOT_, NOT_, NOT_, NOT_ }, { NOL_, NOL_, NOL_, NOL_ } },
    { QR_A4, 1, 0, -1, { T_(QR_A2, QR_A2, 0, 0), NOT_, NOT_, NOT_, NOT_, NOT_ }, { NOL_, NOL_, NOL_, NOL_ } },
    { QR_Z2, 0, 4, 8, { NOT_, NOT_, NOT_, NOT_, NOT_, NOT_ }, { { QR_A, 9 }, { QR_A2, 10 }, { QR_A3, 11 }, { QR_A4, 12 } } },
    { QR_Z1, 1, 3, 4, { T_(QR_A4, QR_Z2, 0, 0), NOT_, NOT_, NOT_, NOT_, NOT_ }, { { QR_A, 5 }, { QR_A2, 6 }, { QR_A3, 7 }, NOL_ } },
    { QR_E0, 1, 3, 0, { T_(QR_A4, QR_Z1, 0, 0), NOT_, NOT_, NOT_, NOT_, NOT_ }, { { QR_A, 1 }, { QR_A2, 2 }, { QR_A3, 3 }, NOL_ } },
};
/* Phase 2: squaring, level l = 0..s-1:  E[l+1] = E[l]^2 */
QCB_TABLE qcb_op QCB_P2[] = {
    { QR_EL1, 1, 0, -1, { T_(QR_EL, QR_EL, 0, 0), NOT_, NOT_, NOT_, NOT_, NOT_ }, { NOL_, NOL_, NOL_, NOL_ } },
};
/* Phase 3: adjoint weight chain, level l = s-1..0:  W[l] = W[l+1] E[l]^T + E[l]^T W[l+1] */
QCB_TABLE qcb_op QCB_P3[] = {
    { QR_WL, 2, 0, -1, { T_(QR_WL1, QR_EL, 0, 1), T_(QR_EL, QR_WL1, 1, 0), NOT_, NOT_, NOT_, NOT_ }, { NOL_, NOL_, NOL_, NOL_ } },
};
/* Phase 4: polynomial, reverse sweep seeded with W[0] (per unit) */
QCB_TABLE qcb_op QCB_P4[] = {
    { QR_ZB1, 1, 0, -1, { T_(QR_A4, QR_W0, 1, 0), NOT_, NOT_, NOT_, NOT_, NOT_ }, { NOL_, NOL_, NOL_, NOL_ } },
    { QR_ZB2, 1, 0, -1, { T_(QR_A4, QR_ZB1, 1, 0), NOT_, NOT_, NOT_, NOT_, NOT_ }, { NOL_, NOL_, NOL_, NOL_ } },
    { QR_AB4, 2, 1, -1, { T_(QR_W0, QR_Z1, 0, 1), T_(QR_ZB1, QR_Z2, 0, 1), NOT_, NOT_, NOT_, NOT_ }, { { QR_ZB2, 12 }, NOL_, NOL_, NOL_ } },
    { QR_AB3, 0, 3, -1, { NOT_, NOT_, NOT_, NOT_, NOT_, NOT_ }, { { QR_W0, 3 }, { QR_ZB1, 7 }, { QR_ZB2, 11 }, NOL_ } },
    { QR_AB2, 3, 3, -1, { T_(QR_AB4, QR_A2, 0, 1), T_(QR_A2, QR_AB4, 1, 0), T_(QR_A, QR_AB3, 1, 0), NOT_, NOT_, NOT_ }, { { QR_W0, 2 }, { QR_ZB1, 6 }, { QR_ZB2, 10 }, NOL_ } },
};
/* Phase 5: polynomial, forward tangent along G_j (per unit x direction) */
QCB_TABLE qcb_op QCB_P5[] = {
    { QR_DA2, 2, 0, -1, { T_(QR_GJ, QR_A, 0, 0), T_(QR_A, QR_GJ, 0, 0), NOT_, NOT_, NOT_, NOT_ }, { NOL_, NOL_, NOL_, NOL_ } },
    { QR_DA3, 2, 0, -1, { T_(QR_GJ, QR_A2, 0, 0), T_(QR_A, QR_DA2, 0, 0), NOT_, NOT_, NOT_, NOT_ }, { NOL_, NOL_, NOL_, NOL_ } },
    { QR_DA4, 2, 0, -1, { T_(QR_DA2, QR_A2, 0, 0), T_(QR_A2, QR_DA2, 0, 0), NOT_, NOT_, NOT_, NOT_ }, { NOL_, NOL_, NOL_, NOL_ } },
    { QR_DZ2, 0, 4, -1, { NOT_, NOT_, NOT_, NOT_, NOT_, NOT_ }, { { QR_GJ, 9 }, { QR_DA2, 10 }, { QR_DA3, 11 }, { QR_DA4, 12 } } },
    { QR_DZ1, 2, 3, -1, { T_(QR_DA4, QR_Z2, 0, 0), T_(QR_A4, QR_DZ2, 0, 0), NOT_, NOT_, NOT_, NOT_ }, { { QR_GJ, 5 }, { QR_DA2, 6 }, { QR_DA3, 7 }, NOL_ } },
    { QR_L, 2, 3, -1, { T_(QR_DA4, QR_Z1, 0, 0), T_(QR_A4, QR_DZ1, 0, 0), NOT_, NOT_, NOT_, NOT_ }, { { QR_GJ, 1 }, { QR_DA2, 2 }, { QR_DA3, 3 }, NOL_ } },
};
/* Phase 6: second-order terms at the polynomial level.  For a computation made of products
 * N = X*Y and linear combinations, d^2 <W, T> / du_i du_j = sum over product nodes of
 * <N_bar, X_dot_i Y_dot_j + X_dot_j Y_dot_i>, with N_bar the first-order adjoint of the node (phase
 * 4) and X_dot, Y_dot the forward tangents (phase 5).  Per node and direction j one product
 * U_j = X_dot_j^T N_bar, then C[j][i] += <U_j, Y_dot_i> for every i; H = C + C^T.  The nodes of
 * phase 1: A2 = A*A, A3 = A*A2, A4 = A2*A2, Y2 = A4*Z2, Y1 = A4*Z1 (the two that share the partner
 * dA2 are merged). */
QCB_TABLE qcb_op QCB_P6[] = {
    { QR_U, 1, 0, -1, { T_(QR_GJ, QR_AB2, 1, 0), NOT_, NOT_, NOT_, NOT_, NOT_ }, { NOL_, NOL_, NOL_, NOL_ } },
    { QR_U, 2, 0, -1, { T_(QR_GJ, QR_AB3, 1, 0), T_(QR_DA2, QR_AB4, 1, 0), NOT_, NOT_, NOT_, NOT_ }, { NOL_, NOL_, NOL_, NOL_ } },
    { QR_U, 1, 0, -1, { T_(QR_DA4, QR_ZB1, 1, 0), NOT_, NOT_, NOT_, NOT_, NOT_ }, { NOL_, NOL_, NOL_, NOL_ } },
    { QR_U, 1, 0, -1, { T_(QR_DA4, QR_W0, 1, 0), NOT_, NOT_, NOT_, NOT_, NOT_ }, { NOL_, NOL_, NOL_, NOL_ } },
};
/* the register whose direction-i instance is contracted with U_j after each op of phase 6 */
QCB_TABLE uint8_t QCB_P6_PARTNER[] = { QR_GJ, QR_DA2, QR_DZ2, QR_DZ1 };
/* Phase 7 (per level l = 0..s-1, per direction): the squaring E[l+1] = E[l]*E[l] is a product node
 * with adjoint W[l+1]:  U_j = L_j^T W[l+1] ;  C[j][i] += <U_j, L_i> ;  L_j <- E[l] L_j + L_j E[l] */
QCB_TABLE qcb_op QCB_P7U[] = {
    { QR_U, 1, 0, -1, { T_(QR_L, QR_WL1, 1, 0), NOT_, NOT_, NOT_, NOT_, NOT_ }, { NOL_, NOL_, NOL_, NOL_ } },
};
QCB_TABLE qcb_op QCB_P7L[] = {
    { QR_LN, 2, 0, -1, { T_(QR_EL, QR_L, 0, 0), T_(QR_L, QR_EL, 0, 0), NOT_, NOT_, NOT_, NOT_ }, { NOL_, NOL_, NOL_, NOL_ } },
};
#undef T_
#undef NOT_
#undef NOL_

#define QCB_NOPS(t) ((int)(sizeof(t) / sizeof((t)[0])))

QCB_TABLE double QCB_TAYLOR[QCB_TAYLOR_DEGREE + 1] = {
    1.0, 1.0, 1.0 / 2, 1.0 / 6, 1.0 / 24, 1.0 / 120, 1.0 / 720, 1.0 / 5040, 1.0 / 40320,
    1.0 / 362880, 1.0 / 3628800, 1.0 / 39916800, 1.0 / 479001600 };

/* number of squarings for a given 1-norm of dt*G(u); -1 if out of range / non-finite */
QCB_HD static inline int qcb_squarings(double norm1) {
    if (!(norm1 == norm1) || norm1 > QCB_THETA * (double)(1 << QCB_MAX_SQUARINGS)) return -1;
    int s = 0;
    double t = QCB_THETA;
    while (norm1 > t) { t *= 2.0; ++s; }
    return s;
}

/* ---- canonical per-unit value arenas (sources of the COO value vectors) ----
 * Jacobian arena : [ E (dp*dp) | JU_0..JU_{m-1} (dp*cp each) | JT (dp*cp) | 0.0 | 1.0 ]
 * Hessian arena  : [ HXU_0..HXU_{m-1} (dp*cp each) | HXT (dp*cp) | HUU (m*m) | HUT (m) | HTT (1) | 0.0 ]
 * (HUU = -tau^2 (C + C^T) with C the accumulated node contractions of phases 6 and 7)
 * E is stored with its own sign (the x_k block of the Jacobian is -E: the slot code carries
 * the sign); every other entry is stored with the sign of the derivative of delta.
 * Slot code: bit 31 = negate, bits 0..30 = offset into the unit's arena. */
#define QCB_CODE_NEG 0x80000000u

#endif /* QCB_PROGRAM_H */
