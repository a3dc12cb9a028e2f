// qcb_large.cuh -- large-d path: every op of qcb_program.h becomes one batched FP64 DMMA GEMM
// launch over (unit x direction), registers in HBM scratch.
//
// tcgen05 has no FP64 kind, so the tensor-core route for this path is mma.sync DMMA
// (SASS DMMA.8x8x4 on sm_100a; measured 37.1 TFLOP/s on B200, scratch/dmma_probe.cu).
// Operand slabs are staged global -> shared with 16-byte cp.async (LDGSTS) in a
// multi-stage ring; fragments are read from padded shared tiles without bank conflicts;
// the epilogue fuses the linear terms / identity / per-unit scale of the matrix op.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace qcb {

struct Opnd {
    const double *p;
    long long su, sd, ss;   // strides (doubles) per local unit, per direction, per system
    int ld;
    int pad_;
};
struct GemmTerm { Opnd a, b; int ta, tb; };
struct GemmArgs {
    int M, N, K;            // padded: M % BM == 0, N % BN == 0, K % 16 == 0
    int d_ident;            // identity is added on the first d_ident diagonal entries
    int nterm, nlin;
    GemmTerm term[6];
    Opnd lin[4];
    double lin_coef[4];
    double ident;
    double alpha;           // scale on the product sum
    const double *unit_scale;  // optional per-local-unit scale on the product sum
    Opnd dst;
    int ndir;               // batch = units * ndir, direction fastest
    int unit0_global, nk;   // system of a unit = (unit0_global + unit) / nk
};

__device__ __forceinline__ void cp_async16(double *smem_dst, const double *gsrc) {
    unsigned sa = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(sa), "l"(gsrc));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N)); }

__device__ __forceinline__ void dmma_16x8x4(double (&c)[4], double a0, double a1, double b) {
    asm volatile("mma.sync.aligned.m16n8k4.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};\n"
                 : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3])
                 : "d"(a0), "d"(a1), "d"(b));
}

constexpr int GEMM_BK = 16;
template <int BM, int BN, int BKT = GEMM_BK> struct GemmSmem {
    static constexpr int A_ST = (BKT * (BM + 4) > BM * (BKT + 4)) ? BKT * (BM + 4) : BM * (BKT + 4);
    static constexpr int B_ST = (BKT * (BN + 4) > BN * (BKT + 4)) ? BKT * (BN + 4) : BN * (BKT + 4);
    static constexpr size_t bytes(int stages) {
        size_t pipe = (size_t)stages * (A_ST + B_ST) * sizeof(double);
        size_t ctile = (size_t)BN * (BM + 2) * sizeof(double);
        return pipe > ctile ? pipe : ctile;
    }
};

template <int BM, int BN, int WM, int WN, int STAGES>
__global__ void __launch_bounds__((BM / WM) * (BN / WN) * 32)
bgemm_kernel(const __grid_constant__ GemmArgs g) {
    constexpr int BK = GEMM_BK;
    constexpr int NTHR = (BM / WM) * (BN / WN) * 32;
    constexpr int MT = WM / 16, NT = WN / 8;
    constexpr int A_ST = GemmSmem<BM, BN>::A_ST, B_ST = GemmSmem<BM, BN>::B_ST;
    extern __shared__ __align__(16) double smem[];
    double *As = smem, *Bs = smem + STAGES * A_ST;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int gid = lane >> 2, tig = lane & 3;
    const int wm = warp % (BM / WM), wn = warp / (BM / WM);
    const int tiles_n = g.N / BN, tiles = (g.M / BM) * tiles_n;
    const int batch = blockIdx.x / tiles, tile = blockIdx.x % tiles;
    const int m0 = (tile / tiles_n) * BM, n0 = (tile % tiles_n) * BN;
    const int unit = batch / g.ndir, dir = batch % g.ndir;
    const int sys = (g.unit0_global + unit) / g.nk;
    const long long ou = unit, od = dir, os = sys;

    double acc[MT][NT][4];
#pragma unroll
    for (int i = 0; i < MT; ++i)
#pragma unroll
        for (int j = 0; j < NT; ++j)
#pragma unroll
            for (int q = 0; q < 4; ++q) acc[i][j][q] = 0.0;

    const int kslabs = g.K / BK;
    const int nit = g.nterm * kslabs;

    auto load_stage = [&](int it, int stage) {
        const GemmTerm &t = g.term[it / kslabs];
        const int k0 = (it % kslabs) * BK;
        const double *pa = t.a.p + ou * t.a.su + od * t.a.sd + os * t.a.ss;
        const double *pb = t.b.p + ou * t.b.su + od * t.b.sd + os * t.b.ss;
        double *as = As + stage * A_ST, *bs = Bs + stage * B_ST;
        if (!t.ta) {
            for (int ch = tid; ch < BK * (BM / 2); ch += NTHR) {
                int kc = ch / (BM / 2), r2 = ch % (BM / 2);
                cp_async16(as + kc * (BM + 4) + 2 * r2, pa + (m0 + 2 * r2) + (long long)(k0 + kc) * t.a.ld);
            }
        } else {
            for (int ch = tid; ch < BM * (BK / 2); ch += NTHR) {
                int r = ch / (BK / 2), k2 = ch % (BK / 2);
                cp_async16(as + r * (BK + 4) + 2 * k2, pa + (k0 + 2 * k2) + (long long)(m0 + r) * t.a.ld);
            }
        }
        if (!t.tb) {
            for (int ch = tid; ch < BN * (BK / 2); ch += NTHR) {
                int n = ch / (BK / 2), k2 = ch % (BK / 2);
                cp_async16(bs + n * (BK + 4) + 2 * k2, pb + (k0 + 2 * k2) + (long long)(n0 + n) * t.b.ld);
            }
        } else {
            for (int ch = tid; ch < BK * (BN / 2); ch += NTHR) {
                int kc = ch / (BN / 2), n2 = ch % (BN / 2);
                cp_async16(bs + kc * (BN + 4) + 2 * n2, pb + (n0 + 2 * n2) + (long long)(k0 + kc) * t.b.ld);
            }
        }
    };

#pragma unroll
    for (int s = 0; s < STAGES - 1; ++s) {
        if (s < nit) load_stage(s, s);
        cp_async_commit();
    }
    for (int it = 0; it < nit; ++it) {
        cp_async_wait<STAGES - 2>();
        __syncthreads();
        {
            int nx = it + STAGES - 1;
            if (nx < nit) load_stage(nx, nx % STAGES);
            cp_async_commit();
        }
        const GemmTerm &t = g.term[it / kslabs];
        const int a_rs = t.ta ? (BK + 4) : 1, a_ks = t.ta ? 1 : (BM + 4);
        const int b_ks = t.tb ? (BN + 4) : 1, b_ns = t.tb ? 1 : (BK + 4);
        const double *as = As + (it % STAGES) * A_ST + (wm * WM + gid) * a_rs + tig * a_ks;
        const double *bs = Bs + (it % STAGES) * B_ST + (wn * WN + gid) * b_ns + tig * b_ks;
#pragma unroll
        for (int kk = 0; kk < BK; kk += 4) {
            double a0[MT], a1[MT], b[NT];
#pragma unroll
            for (int i = 0; i < MT; ++i) {
                a0[i] = as[(i * 16) * a_rs + kk * a_ks];
                a1[i] = as[(i * 16 + 8) * a_rs + kk * a_ks];
            }
#pragma unroll
            for (int j = 0; j < NT; ++j) b[j] = bs[(j * 8) * b_ns + kk * b_ks];
#pragma unroll
            for (int i = 0; i < MT; ++i)
#pragma unroll
                for (int j = 0; j < NT; ++j) dmma_16x8x4(acc[i][j], a0[i], a1[i], b[j]);
        }
    }
    cp_async_wait<0>();

    // ---- epilogue: accumulators -> shared tile -> coalesced 16-byte pass that fuses
    //      alpha*scale*acc + sum_l coef_l*lin_l + ident*I and stores dst ----
    __syncthreads();                      // every warp is done with the operand stages
    constexpr int LDC = BM + 2;           // == 2 (mod 16) words: conflict-free fragment stores
    double *Cs = smem;
#pragma unroll
    for (int i = 0; i < MT; ++i)
#pragma unroll
        for (int j = 0; j < NT; ++j)
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                int r = wm * WM + i * 16 + gid + (q >> 1) * 8;
                int c = wn * WN + j * 8 + tig * 2 + (q & 1);
                Cs[c * LDC + r] = acc[i][j][q];
            }
    __syncthreads();
    double sc = g.alpha;
    if (g.unit_scale) sc *= g.unit_scale[unit];
    double *pd = const_cast<double *>(g.dst.p) + ou * g.dst.su + od * g.dst.sd + os * g.dst.ss;
    const double *pl[4];
    double cf[4];
    long long ldl[4];
#pragma unroll
    for (int l = 0; l < 4; ++l) {
        const bool on = l < g.nlin;
        pl[l] = on ? g.lin[l].p + ou * g.lin[l].su + od * g.lin[l].sd + os * g.lin[l].ss : nullptr;
        cf[l] = on ? g.lin_coef[l] : 0.0;
        ldl[l] = on ? g.lin[l].ld : 0;
    }
    // four output vectors per pass: all linear-term loads of a pass are issued before any is used
    constexpr int EPI_U = 4;
    static_assert(((BM / 2) * BN) % (NTHR * EPI_U) == 0, "epilogue tiling");
    for (int base = tid; base < (BM / 2) * BN; base += NTHR * EPI_U) {
        double2 t[EPI_U][4];
#pragma unroll
        for (int u = 0; u < EPI_U; ++u) {
            const int idx = base + u * NTHR, r = 2 * (idx % (BM / 2)), c = idx / (BM / 2);
#pragma unroll
            for (int l = 0; l < 4; ++l)
                t[u][l] = pl[l] ? *reinterpret_cast<const double2 *>(pl[l] + (m0 + r) + (long long)(n0 + c) * ldl[l])
                                : make_double2(0.0, 0.0);
        }
#pragma unroll
        for (int u = 0; u < EPI_U; ++u) {
            const int idx = base + u * NTHR, r = 2 * (idx % (BM / 2)), c = idx / (BM / 2);
            double2 v = *reinterpret_cast<const double2 *>(Cs + c * LDC + r);
            v.x *= sc; v.y *= sc;
#pragma unroll
            for (int l = 0; l < 4; ++l) { v.x = fma(cf[l], t[u][l].x, v.x); v.y = fma(cf[l], t[u][l].y, v.y); }
            const int gr = m0 + r, gc = n0 + c;
            if (gc == gr && gr < g.d_ident) v.x += g.ident;
            if (gc == gr + 1 && gc < g.d_ident) v.y += g.ident;
            *reinterpret_cast<double2 *>(pd + gr + (long long)gc * g.dst.ld) = v;
        }
    }
}

// ---------------------------------------------------------------------------------------
// Real-iso half-storage GEMM.
//
// Every matrix the program touches for a quantum generator G = iso(-iH) (closed or Lindblad)
// has the block form  M = [[P, -Q], [Q, P]]  (n x n blocks, d = 2n), and so do products, sums and
// transposes of such matrices.  Only the left block column M_L = [P; Q] (d x n, i.e. the first
// d*n doubles of the column-major matrix) is stored and computed:
//     [op(A) op(B)]_L = op(A)_full * op(B)_L          (d x d) * (d x n): half the flops of d^3.
// The full left operand is reconstructed while loading tiles:
//     X_full(i, j) = X_L(i, j)                      j <  n
//                  = -X_L(i + n, j - n)             j >= n, i <  n
//                  =  X_L(i - n, j - n)             j >= n, i >= n
// (rows rotated by n, sign on the top half), and  (Y^T)_L(k, c) = Y_L(c, k) for k < n,
// -Y_L(c + n, k - n) for k >= n.  The sign is uniform per (warp row-half, K slab) and is applied to
// the A fragments with one integer XOR each.  The right operand may also be a plain d x N panel
// (tb = 0 reads it as is), which is how E X, L_j X, L_j^T M are formed without expanding E, L_j.
// Tile 128 x 64, 8 warps (32 x 32 each), 3-stage cp.async ring -> 2 CTAs per SM, so one CTA's
// epilogue overlaps the other's DMMA main loop.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ double flip_sign(double v, unsigned mask) {
    return __hiloint2double(__double2hiint(v) ^ (int)mask, __double2loint(v));
}

// one staged K slab (16 deep) of the 128 x 64 tile: fragments from shared memory, 64 DMMA per warp
template <int TA, int TB, int MT, int NT, int BK>
__device__ __forceinline__ void iso_slab_mma(double (&acc)[MT][NT][4], const double *as0, const double *bs0,
                                             int arow, int bcol, int tig) {
    constexpr int BM = 128, BN = 64;
    constexpr int a_rs = TA ? (BK + 4) : 1, a_ks = TA ? 1 : (BM + 4);
    constexpr int b_ks = TB ? (BN + 4) : 1, b_ns = TB ? 1 : (BK + 4);
    const double *as = as0 + arow * a_rs + tig * a_ks;
    const double *bs = bs0 + bcol * b_ns + tig * b_ks;
#pragma unroll
    for (int kk = 0; kk < BK; kk += 4) {
        double a0[MT], a1[MT], b[NT];
#pragma unroll
        for (int i = 0; i < MT; ++i) {
            a0[i] = as[(i * 16) * a_rs + kk * a_ks];
            a1[i] = as[(i * 16 + 8) * a_rs + kk * a_ks];
        }
#pragma unroll
        for (int j = 0; j < NT; ++j) b[j] = bs[(j * 8) * b_ns + kk * b_ks];
#pragma unroll
        for (int i = 0; i < MT; ++i)
#pragma unroll
            for (int j = 0; j < NT; ++j) dmma_16x8x4(acc[i][j], a0[i], a1[i], b[j]);
    }
}

template <int STAGES, int BK = GEMM_BK>
__global__ void __launch_bounds__(256, 2)
bgemm_iso_kernel(const __grid_constant__ GemmArgs g) {
    constexpr int BM = 128, BN = 64, WM = 32, WN = 32, NTHR = 256;
    constexpr int MT = WM / 16, NT = WN / 8;
    constexpr int A_ST = GemmSmem<BM, BN, BK>::A_ST, B_ST = GemmSmem<BM, BN, BK>::B_ST;
    constexpr int NA = BK / 4, NB = BK / 8;          // 16-byte chunks per thread and slab: A 128 x BK, B BK x 64
    constexpr int KT = BK / 2;                       // threads that cover the K extent of a slab in pairs
    extern __shared__ __align__(16) double smem[];
    double *As = smem, *Bs = smem + STAGES * A_ST;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int gid = lane >> 2, tig = lane & 3;
    const int wm = warp & 3, wn = warp >> 2;
    const int tiles_n = g.N / BN, tiles = (g.M / BM) * tiles_n;
    const int batch = blockIdx.x / tiles, tile = blockIdx.x % tiles;
    const int m0 = (tile / tiles_n) * BM, n0 = (tile % tiles_n) * BN;
    const int unit = batch / g.ndir, dir = batch % g.ndir;
    const int sys = (g.unit0_global + unit) / g.nk;
    const long long ou = unit, od = dir, os = sys;
    const int dfull = g.K, nh = g.K >> 1;          // d and n = d/2 (K is always the full dimension)
    const bool warp_top = (m0 + wm * WM) < nh;     // rows of this warp lie in the top half

    double acc[MT][NT][4];
#pragma unroll
    for (int i = 0; i < MT; ++i)
#pragma unroll
        for (int j = 0; j < NT; ++j)
#pragma unroll
            for (int q = 0; q < 4; ++q) acc[i][j][q] = 0.0;

    const int kslabs = g.K / BK, half = kslabs >> 1;
    const int nit = g.nterm * kslabs;

    // ---- loader state: element offsets of this thread's 4 + 2 sixteen-byte chunks inside the two
    //      operands of the current term; advanced by a constant per slab, corrected once when the
    //      slab crosses k = n (the rotated half of the virtual full operand) ----
    const double *l_pa = nullptr, *l_pb = nullptr;
    int offA[NA], adjA[NA], offB[NB], stepA = 0, stepB = 0, adjB = 0;
    int sA = 0, sAstep = 0, sB = 0, sBstep = 0;     // shared-memory offsets of chunk 0 and the chunk stride
    int l_term = 0, l_slab = 0, l_stage = 0;
    auto loader_init = [&](int term) {
        const GemmTerm &t = g.term[term];
        l_pa = t.a.p + ou * t.a.su + od * t.a.sd + os * t.a.ss;
        l_pb = t.b.p + ou * t.b.su + od * t.b.sd + os * t.b.ss;
        const int lda = t.a.ld, ldb = t.b.ld;
        if (!t.ta) {        // X_full(i, k): column k mod n, rows rotated by n for k >= n
            const int r = m0 + 2 * (tid & 63), kc = tid >> 6;
            const int rrot = (r < nh) ? r + nh : r - nh;
#pragma unroll
            for (int i = 0; i < NA; ++i) { offA[i] = r + (kc + 4 * i) * lda; adjA[i] = (rrot - r) - nh * lda; }
            stepA = BK * lda;
            sA = kc * (BM + 4) + 2 * (tid & 63); sAstep = 4 * (BM + 4);
        } else {            // X_full(k, i): i < n -> X_L(k, i);  i >= n -> +-X_L((k + n) mod d, i - n)
            const int k2 = 2 * (tid % KT);
#pragma unroll
            for (int i = 0; i < NA; ++i) {
                const int irow = m0 + (tid / KT) + (NTHR / KT) * i;
                const bool ihi = irow >= nh;
                offA[i] = (ihi ? k2 + nh : k2) + (ihi ? irow - nh : irow) * lda;
                adjA[i] = ihi ? -dfull : 0;
            }
            stepA = BK;
            sA = (tid / KT) * (BK + 4) + k2; sAstep = (NTHR / KT) * (BK + 4);
        }
        if (!t.tb) {        // plain K x N right operand
            const int k2 = 2 * (tid % KT);
#pragma unroll
            for (int i = 0; i < NB; ++i) offB[i] = k2 + (n0 + (tid / KT) + (NTHR / KT) * i) * ldb;
            stepB = BK; adjB = 0;
            sB = (tid / KT) * (BK + 4) + k2; sBstep = (NTHR / KT) * (BK + 4);
        } else {            // (Y^T)_L(k, c) = Y_L(c, k) (k < n) or -Y_L(c + n, k - n)
            const int n2 = 2 * (tid & 31), kc = tid >> 5;
#pragma unroll
            for (int i = 0; i < NB; ++i) offB[i] = (n0 + n2) + (kc + 8 * i) * ldb;
            stepB = BK * ldb; adjB = nh - nh * ldb;
            sB = kc * (BN + 4) + n2; sBstep = 8 * (BN + 4);
        }
    };
    auto loader_issue = [&]() {
        double *as = As + l_stage * A_ST + sA, *bs = Bs + l_stage * B_ST + sB;
#pragma unroll
        for (int i = 0; i < NA; ++i) cp_async16(as + i * sAstep, l_pa + offA[i]);
#pragma unroll
        for (int i = 0; i < NB; ++i) cp_async16(bs + i * sBstep, l_pb + offB[i]);
        l_stage = (l_stage + 1 == STAGES) ? 0 : l_stage + 1;
        if (++l_slab == kslabs) {
            l_slab = 0;
            if (++l_term < g.nterm) loader_init(l_term);
        } else {
            const bool cross = (l_slab == half);
#pragma unroll
            for (int i = 0; i < NA; ++i) offA[i] += stepA + (cross ? adjA[i] : 0);
#pragma unroll
            for (int i = 0; i < NB; ++i) offB[i] += stepB + (cross ? adjB : 0);
        }
    };

    if (nit > 0) loader_init(0);
#pragma unroll
    for (int s = 0; s < STAGES - 1; ++s) {
        if (s < nit) loader_issue();
        cp_async_commit();
    }
    // ---- consumer state ----
    int c_term = 0, c_slab = 0, c_stage = 0;
    int a_rs = 1, a_ks = BM + 4, b_ks = 1, b_ns = BK + 4, c_ta = 0, c_tb = 0;
    bool cur_neg = false;      // accumulators currently hold minus the running sum
    auto consumer_term = [&](int term) {
        const GemmTerm &t = g.term[term];
        c_ta = t.ta; c_tb = t.tb;
        a_rs = t.ta ? (BK + 4) : 1; a_ks = t.ta ? 1 : (BM + 4);
        b_ks = t.tb ? (BN + 4) : 1; b_ns = t.tb ? 1 : (BK + 4);
    };
    if (nit > 0) consumer_term(0);
    for (int it = 0; it < nit; ++it) {
        cp_async_wait<STAGES - 2>();
        __syncthreads();
        if (it + STAGES - 1 < nit) loader_issue();
        cp_async_commit();
        // sign of this (warp, slab): left operand ta=0: -(khi && top); ta=1: -(bottom && !khi);
        // right operand tb=1: -khi.  Instead of negating fragments, the accumulators are negated
        // whenever the sign changes (at most twice per term): acc -= a*b  ==  -( (-acc) + a*b ).
        const bool khi = c_slab >= half;
        bool neg = c_ta ? (!warp_top && !khi) : (warp_top && khi);
        if (c_tb && khi) neg = !neg;
        if (neg != cur_neg) {
#pragma unroll
            for (int i = 0; i < MT; ++i)
#pragma unroll
                for (int j = 0; j < NT; ++j)
#pragma unroll
                    for (int q = 0; q < 4; ++q) acc[i][j][q] = -acc[i][j][q];
            cur_neg = neg;
        }
        // four compile-time layouts of the staged tiles -> LDS with immediate offsets
        const double *as0 = As + c_stage * A_ST, *bs0 = Bs + c_stage * B_ST;
        if (!c_ta && !c_tb) iso_slab_mma<0, 0, MT, NT, BK>(acc, as0, bs0, wm * WM + gid, wn * WN + gid, tig);
        else if (c_ta && !c_tb) iso_slab_mma<1, 0, MT, NT, BK>(acc, as0, bs0, wm * WM + gid, wn * WN + gid, tig);
        else if (!c_ta && c_tb) iso_slab_mma<0, 1, MT, NT, BK>(acc, as0, bs0, wm * WM + gid, wn * WN + gid, tig);
        else iso_slab_mma<1, 1, MT, NT, BK>(acc, as0, bs0, wm * WM + gid, wn * WN + gid, tig);
        c_stage = (c_stage + 1 == STAGES) ? 0 : c_stage + 1;
        if (++c_slab == kslabs) {
            c_slab = 0;
            if (++c_term < g.nterm) consumer_term(c_term);
        }
    }
    if (cur_neg) {
#pragma unroll
        for (int i = 0; i < MT; ++i)
#pragma unroll
            for (int j = 0; j < NT; ++j)
#pragma unroll
                for (int q = 0; q < 4; ++q) acc[i][j][q] = -acc[i][j][q];
    }
    cp_async_wait<0>();
    (void)dfull;

    if (g.nlin == 0) {
        // No linear terms to read: store the accumulator fragments straight from registers (each
        // warp store covers full 32-byte sectors: 8 consecutive rows x 8 bytes per column) -- no
        // shared-memory staging, no barrier, and the CTA's slot frees up ~2 us earlier.
        double sc0 = g.alpha;
        if (g.unit_scale) sc0 *= g.unit_scale[unit];
        double *pd0 = const_cast<double *>(g.dst.p) + ou * g.dst.su + od * g.dst.sd + os * g.dst.ss;
#pragma unroll
        for (int i = 0; i < MT; ++i)
#pragma unroll
            for (int j = 0; j < NT; ++j)
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int gr = m0 + wm * WM + i * 16 + gid + (q >> 1) * 8;
                    const int gc = n0 + wn * WN + j * 8 + tig * 2 + (q & 1);
                    double v = sc0 * acc[i][j][q];
                    if (gr == gc && gr < g.d_ident) v += g.ident;
                    pd0[gr + (long long)gc * g.dst.ld] = v;
                }
        return;
    }
    __syncthreads();
    constexpr int LDC = BM + 2;
    double *Cs = smem;
#pragma unroll
    for (int i = 0; i < MT; ++i)
#pragma unroll
        for (int j = 0; j < NT; ++j)
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                int r = wm * WM + i * 16 + gid + (q >> 1) * 8;
                int c = wn * WN + j * 8 + tig * 2 + (q & 1);
                Cs[c * LDC + r] = acc[i][j][q];
            }
    __syncthreads();
    double sc = g.alpha;
    if (g.unit_scale) sc *= g.unit_scale[unit];
    double *pd = const_cast<double *>(g.dst.p) + ou * g.dst.su + od * g.dst.sd + os * g.dst.ss;
    const double *pl[4];
    double cf[4];
    long long ldl[4];
#pragma unroll
    for (int l = 0; l < 4; ++l) {
        const bool on = l < g.nlin;
        pl[l] = on ? g.lin[l].p + ou * g.lin[l].su + od * g.lin[l].sd + os * g.lin[l].ss : nullptr;
        cf[l] = on ? g.lin_coef[l] : 0.0;
        ldl[l] = on ? g.lin[l].ld : 0;
    }
    // four output vectors per pass: all linear-term loads of a pass are issued before any is used
    constexpr int EPI_U = 4;
    static_assert(((BM / 2) * BN) % (NTHR * EPI_U) == 0, "epilogue tiling");
    for (int base = tid; base < (BM / 2) * BN; base += NTHR * EPI_U) {
        double2 t[EPI_U][4];
#pragma unroll
        for (int u = 0; u < EPI_U; ++u) {
            const int idx = base + u * NTHR, r = 2 * (idx % (BM / 2)), c = idx / (BM / 2);
#pragma unroll
            for (int l = 0; l < 4; ++l)
                t[u][l] = pl[l] ? *reinterpret_cast<const double2 *>(pl[l] + (m0 + r) + (long long)(n0 + c) * ldl[l])
                                : make_double2(0.0, 0.0);
        }
#pragma unroll
        for (int u = 0; u < EPI_U; ++u) {
            const int idx = base + u * NTHR, r = 2 * (idx % (BM / 2)), c = idx / (BM / 2);
            double2 v = *reinterpret_cast<const double2 *>(Cs + c * LDC + r);
            v.x *= sc; v.y *= sc;
#pragma unroll
            for (int l = 0; l < 4; ++l) { v.x = fma(cf[l], t[u][l].x, v.x); v.y = fma(cf[l], t[u][l].y, v.y); }
            const int gr = m0 + r, gc = n0 + c;
            if (gc == gr && gr < g.d_ident) v.x += g.ident;
            if (gc == gr + 1 && gc < g.d_ident) v.y += g.ident;
            *reinterpret_cast<double2 *>(pd + gr + (long long)gc * g.dst.ld) = v;
        }
    }
}

// E(:, n:d) from E_L = E(:, 0:n) of an iso matrix stored at the head of each unit's arena
__global__ void __launch_bounds__(256) iso_expand_kernel(double *arena, long long stride, int d) {
    const int n = d >> 1;
    double *E = arena + (long long)blockIdx.x * stride;
    for (int e = threadIdx.x; e < d * n; e += blockDim.x) {
        const int i = e % d, j = e / d;            // target (i, j + n)
        const double v = (i < n) ? -E[(i + n) + (long long)j * d] : E[(i - n) + (long long)j * d];
        E[i + (long long)(j + n) * d] = v;
    }
}

// W_L = left block column of the projection of a general d x d matrix onto the iso algebra:
//   P = (W11 + W22)/2,  Q = (W21 - W12)/2.   (<W, F> = <proj W, F> for every iso matrix F.)
__global__ void __launch_bounds__(256)
iso_project_kernel(const double *Wfull, long long w_stride, double *WL, long long wl_stride, int d) {
    const int n = d >> 1;
    const double *W = Wfull + (long long)blockIdx.x * w_stride;
    double *o = WL + (long long)blockIdx.x * wl_stride;
    for (int e = threadIdx.x; e < d * n; e += blockDim.x) {
        const int i = e % d, j = e / d;
        double v;
        if (i < n) v = 0.5 * (W[i + (long long)j * d] + W[(i + n) + (long long)(j + n) * d]);
        else v = 0.5 * (W[i + (long long)j * d] - W[(i - n) + (long long)(j + n) * d]);
        o[e] = v;
    }
}

// dst = sum_l coef_l * R_l + ident * I for ops without products (no tensor work): plain streaming
// kernel, 16-byte accesses, one CTA per (batch item, column block).
__global__ void __launch_bounds__(256) lincomb_kernel(const __grid_constant__ GemmArgs g) {
    const int cols_per_cta = 8;
    const int cblocks = g.N / cols_per_cta;
    const int batch = blockIdx.x / cblocks, c0 = (blockIdx.x % cblocks) * cols_per_cta;
    const int unit = batch / g.ndir, dir = batch % g.ndir;
    const int sys = (g.unit0_global + unit) / g.nk;
    const long long ou = unit, od = dir, os = sys;
    double *pd = const_cast<double *>(g.dst.p) + ou * g.dst.su + od * g.dst.sd + os * g.dst.ss;
    const double *pl[4];
    double cf[4];
#pragma unroll
    for (int l = 0; l < 4; ++l) {
        const bool on = l < g.nlin;
        pl[l] = on ? g.lin[l].p + ou * g.lin[l].su + od * g.lin[l].sd + os * g.lin[l].ss : nullptr;
        cf[l] = on ? g.lin_coef[l] : 0.0;
    }
    const int half_rows = g.M / 2;
    for (int idx = threadIdx.x; idx < half_rows * cols_per_cta; idx += blockDim.x) {
        const int r = 2 * (idx % half_rows), c = c0 + idx / half_rows;
        double2 v = make_double2(0.0, 0.0);
#pragma unroll
        for (int l = 0; l < 4; ++l)
            if (pl[l]) {
                const double2 t = *reinterpret_cast<const double2 *>(pl[l] + r + (long long)c * g.lin[l].ld);
                v.x = fma(cf[l], t.x, v.x); v.y = fma(cf[l], t.y, v.y);
            }
        if (c == r && r < g.d_ident) v.x += g.ident;
        if (c == r + 1 && c < g.d_ident) v.y += g.ident;
        *reinterpret_cast<double2 *>(pd + r + (long long)c * g.dst.ld) = v;
    }
}

// ---------------------------------------------------------------------------------------
// Products with a sparse drive generator.  The drive Hamiltonians of real devices are Paulis /
// ladder operators: G_j has a handful of nonzeros per row (the reference itself keeps them as
// SparseArrays), so G_j*X, G_j^T*X and X*G_j are O(d^2) gathers, not O(d^3) GEMMs.  ELL format:
// ell_idx / ell_val [S][m][d][W] hold the nonzeros of every row of G_j, ellT_* those of every
// column (= rows of G_j^T); unused slots have value 0 and index 0.
// dst = sum of up to two such terms (iso half storage or full storage); a following GEMM adds the
// op's dense terms on top through a linear term that reads dst back.
// ---------------------------------------------------------------------------------------
struct SparseArgs {
    Opnd dst;
    Opnd xl;                // left term  T * X  : T = G_j (tables ell_*) or G_j^T (tables ellT_*), chosen by the host
    Opnd xr;                // right term X * G_j (only sparse_drive_kernel<W, true>)
    const int *l_idx;       // [S][m][d][W] rows of T
    const double *l_val;
    const int *r_idx;       // [S][m][d][W] columns of G_j
    const double *r_val;
    int d, ncols, iso;      // rows, stored columns (d or d/2), iso half storage?
    int m, ndir, unit0_global, nk;
};
// dst = T X (+ X G_j).  Block (64, 4): x walks row PAIRS (one 16-byte store per pair and column), y the
// columns; the pair's ELL entries are loaded once, eight columns per pass keep the gathers in flight.
// The term kinds are compile-time so that the loop body is a handful of instructions per element
// (the first version spent ~60 and was issue bound at a third of the HBM write rate).
template <int W, bool RIGHT>
__global__ void __launch_bounds__(256) sparse_drive_kernel(const __grid_constant__ SparseArgs g) {
    const int unit = blockIdx.x / g.ndir, dir = blockIdx.x - unit * g.ndir;
    const int sys = (g.unit0_global + unit) / g.nk;
    const long long ou = unit, od = dir, os = sys;
    double *__restrict__ dst = const_cast<double *>(g.dst.p) + ou * g.dst.su + od * g.dst.sd + os * g.dst.ss;
    const double *__restrict__ XL = g.xl.p + ou * g.xl.su + od * g.xl.sd + os * g.xl.ss;
    const double *__restrict__ XR = RIGHT ? g.xr.p + ou * g.xr.su + od * g.xr.sd + os * g.xr.ss : nullptr;
    const int d = g.d, n = d >> 1, ncols = g.ncols, ldd = g.dst.ld, ldl = g.xl.ld, ldr = g.xr.ld;
    const int ell_base = (sys * g.m + dir) * d * W;
    const int *__restrict__ lidx = g.l_idx + ell_base;
    const double *__restrict__ lval = g.l_val + ell_base;
    const int *__restrict__ ridx = RIGHT ? g.r_idx + ell_base : nullptr;
    const double *__restrict__ rval = RIGHT ? g.r_val + ell_base : nullptr;
    const bool iso = g.iso != 0;
    constexpr int UNR = 8;
    const int cstep = blockDim.y;
    for (int r0 = 2 * threadIdx.x; r0 < d; r0 += 2 * blockDim.x) {
        int ia[W], ib[W];
        double va[W], vb[W];
#pragma unroll
        for (int w = 0; w < W; ++w) {
            ia[w] = __ldg(lidx + r0 * W + w); va[w] = __ldg(lval + r0 * W + w);
            ib[w] = __ldg(lidx + (r0 + 1) * W + w); vb[w] = __ldg(lval + (r0 + 1) * W + w);
        }
        // right term, iso half storage: X_full[r, k] for k >= n is -X[r + n, k - n] (r < n) or X[r - n, k - n]
        const int ra = (r0 < n) ? r0 + n : r0 - n, rb = (r0 + 1 < n) ? r0 + 1 + n : r0 + 1 - n;
        const double sa = (r0 < n) ? -1.0 : 1.0, sb = (r0 + 1 < n) ? -1.0 : 1.0;
        for (int c0 = threadIdx.y; c0 < ncols; c0 += UNR * cstep) {
            double a0[UNR], a1[UNR];
#pragma unroll
            for (int u = 0; u < UNR; ++u) {
                const int c = c0 + u * cstep;
                a0[u] = a1[u] = 0.0;
                if (c < ncols) {
                    const double *__restrict__ xc = XL + c * ldl;
#pragma unroll
                    for (int w = 0; w < W; ++w) {
                        a0[u] = fma(va[w], xc[ia[w]], a0[u]);
                        a1[u] = fma(vb[w], xc[ib[w]], a1[u]);
                    }
                    if (RIGHT) {
#pragma unroll
                        for (int w = 0; w < W; ++w) {
                            const int k = __ldg(ridx + c * W + w);
                            const double v = __ldg(rval + c * W + w);
                            if (!iso || k < n) {
                                a0[u] = fma(v, XR[r0 + k * ldr], a0[u]);
                                a1[u] = fma(v, XR[r0 + 1 + k * ldr], a1[u]);
                            } else {
                                a0[u] = fma(v * sa, XR[ra + (k - n) * ldr], a0[u]);
                                a1[u] = fma(v * sb, XR[rb + (k - n) * ldr], a1[u]);
                            }
                        }
                    }
                }
            }
#pragma unroll
            for (int u = 0; u < UNR; ++u) {
                const int c = c0 + u * cstep;
                if (c < ncols) *reinterpret_cast<double2 *>(dst + r0 + c * ldd) = make_double2(a0[u], a1[u]);
            }
        }
    }
}

// ---------------------------------------------------------------------------------------
// Frobenius contractions  out[unit][j*ni + i] (+)= < P(unit, j), Q(unit|sys, i) >
// ---------------------------------------------------------------------------------------
constexpr int CONTRACT_MAXI = 32;
struct ContractArgs {
    Opnd P, Q;        // P: sd = stride per j;  Q: sd = stride per i
    int nj, ni;
    long long nelem;  // dp*dp (contiguous, ld == dp)
    double *out;      // [units][nj*ni]
    int accumulate;
    int unit0_global, nk;
};
__global__ void __launch_bounds__(256) contract_kernel(const __grid_constant__ ContractArgs g) {
    const int unit = blockIdx.x / g.nj, j = blockIdx.x % g.nj;
    const int sys = (g.unit0_global + unit) / g.nk;
    const double *P = g.P.p + (long long)unit * g.P.su + (long long)j * g.P.sd + (long long)sys * g.P.ss;
    const double *Q = g.Q.p + (long long)unit * g.Q.su + (long long)sys * g.Q.ss;
    double acc[CONTRACT_MAXI];
#pragma unroll
    for (int i = 0; i < CONTRACT_MAXI; ++i) acc[i] = 0.0;
    for (long long e = threadIdx.x; e < g.nelem; e += blockDim.x) {
        double p = P[e];
#pragma unroll
        for (int i = 0; i < CONTRACT_MAXI; ++i)
            if (i < g.ni) acc[i] = fma(p, Q[(long long)i * g.Q.sd + e], acc[i]);
    }
    __shared__ double red[8][CONTRACT_MAXI];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int i = 0; i < CONTRACT_MAXI; ++i) {
        if (i < g.ni) {
            double v = acc[i];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if (lane == 0) red[warp][i] = v;
        }
    }
    __syncthreads();
    if (threadIdx.x < g.ni) {
        double v = 0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) v += red[w][threadIdx.x];
        double *o = g.out + ((long long)unit * g.nj + j) * g.ni + threadIdx.x;
        *o = g.accumulate ? (*o + v) : v;
    }
}

// Tensor-core variant for nj, ni <= 16: H = P Q^T with K = nelem on DMMA, one CTA per unit, every
// matrix read exactly once (32-byte sectors fully used).  The sum over elements is order-free, so
// lane (gid, tig) feeds elements e0 + 4*tig + t of rows gid / gid+8 into step t of both operands.
__global__ void __launch_bounds__(256) contract_mma_kernel(const __grid_constant__ ContractArgs g) {
    const int unit = blockIdx.x;
    const int sys = (g.unit0_global + unit) / g.nk;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, gid = lane >> 2, tig = lane & 3;
    const double *P = g.P.p + (long long)unit * g.P.su + (long long)sys * g.P.ss;
    const double *Q = g.Q.p + (long long)unit * g.Q.su + (long long)sys * g.Q.ss;
    const bool pa0 = gid < g.nj, pa1 = gid + 8 < g.nj, pb0 = gid < g.ni, pb1 = gid + 8 < g.ni;
    const double *A0 = P + (long long)gid * g.P.sd + 4 * tig, *A1 = P + (long long)(gid + 8) * g.P.sd + 4 * tig;
    const double *B0 = Q + (long long)gid * g.Q.sd + 4 * tig, *B1 = Q + (long long)(gid + 8) * g.Q.sd + 4 * tig;
    double c0[4] = {0, 0, 0, 0}, c1[4] = {0, 0, 0, 0};
    const double2 z2 = make_double2(0.0, 0.0);
    for (long long e = (long long)warp * 16; e < g.nelem; e += 8 * 16) {
        double2 a0x = z2, a0y = z2, a1x = z2, a1y = z2, b0x = z2, b0y = z2, b1x = z2, b1y = z2;
        if (pa0) { a0x = *reinterpret_cast<const double2 *>(A0 + e); a0y = *reinterpret_cast<const double2 *>(A0 + e + 2); }
        if (pa1) { a1x = *reinterpret_cast<const double2 *>(A1 + e); a1y = *reinterpret_cast<const double2 *>(A1 + e + 2); }
        if (pb0) { b0x = *reinterpret_cast<const double2 *>(B0 + e); b0y = *reinterpret_cast<const double2 *>(B0 + e + 2); }
        if (pb1) { b1x = *reinterpret_cast<const double2 *>(B1 + e); b1y = *reinterpret_cast<const double2 *>(B1 + e + 2); }
        dmma_16x8x4(c0, a0x.x, a1x.x, b0x.x); dmma_16x8x4(c1, a0x.x, a1x.x, b1x.x);
        dmma_16x8x4(c0, a0x.y, a1x.y, b0x.y); dmma_16x8x4(c1, a0x.y, a1x.y, b1x.y);
        dmma_16x8x4(c0, a0y.x, a1y.x, b0y.x); dmma_16x8x4(c1, a0y.x, a1y.x, b1y.x);
        dmma_16x8x4(c0, a0y.y, a1y.y, b0y.y); dmma_16x8x4(c1, a0y.y, a1y.y, b1y.y);
    }
    __shared__ double red[8][16][17];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const int r = gid + (q >> 1) * 8, cc = tig * 2 + (q & 1);
        red[warp][r][cc] = c0[q];
        red[warp][r][cc + 8] = c1[q];
    }
    __syncthreads();
    const int j = threadIdx.x >> 4, i = threadIdx.x & 15;
    if (j < g.nj && i < g.ni) {
        double v = 0;
#pragma unroll
        for (int w = 0; w < 8; ++w) v += red[w][j][i];
        double *o = g.out + ((long long)unit * g.nj + j) * g.ni + i;
        *o = g.accumulate ? (*o + v) : v;
    }
}

// ---------------------------------------------------------------------------------------
// Set-up kernels
// ---------------------------------------------------------------------------------------
struct UnitGeom {
    int d, c, m, dp, cp, z_dim, u_off, dt_off, K, nk, k_begin, S;
    int gcols;               // stored columns of every generator-like matrix: dp, or dp/2 in iso half storage
    long long gsz;           // doubles per stored matrix = dp * gcols
    const int *x_off;        // device [S]
    const double *G0, *Gd;   // device, padded: [S][dp*dp], [S][m][dp*dp]
};

// G = G0 + sum u_j G_j for every local unit in [u0, u0+nu) (padded dp x dp), plus ||dt*G||_1
// folded into a global max (norm as ordered-uint64 bits; NaN/Inf raise the flag).
__global__ void __launch_bounds__(256)
build_G_kernel(UnitGeom q, const double *Z, int u0_global, double *Gout, long long g_stride,
               unsigned long long *norm_max_bits, int *bad_flag, double *unit_norm = nullptr,
               const int *perm = nullptr) {
    // perm: units sorted by squaring count (single system only); position -> slab-local unit
    const int unit = blockIdx.x, ug = perm ? perm[u0_global + unit] : u0_global + unit;
    const int sys = ug / q.nk, k = q.k_begin + ug % q.nk;
    const double *zk = Z + (long long)k * q.z_dim;
    const long long dd = q.gsz;
    const double *G0 = q.G0 + sys * dd, *Gd = q.Gd + (long long)sys * q.m * dd;
    double *G = Gout ? Gout + (long long)unit * g_stride : nullptr;
    const double dt = zk[q.dt_off];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
    __shared__ double wmax[8];
    __shared__ double su[CONTRACT_MAXI];
    if (threadIdx.x < q.m) su[threadIdx.x] = zk[q.u_off + threadIdx.x];
    __syncthreads();
    // one warp per column: lanes stride the rows, a shuffle reduction gives the column's 1-norm
    // (iso half storage: the right half has the same column sums)
    double worst = 0.0;
    bool bad = !isfinite(dt);
    for (int col = warp; col < q.gcols; col += nwarp) {
        double part = 0.0;
        for (int r = lane; r < q.dp; r += 32) {
            const long long e = r + (long long)col * q.dp;
            double v = G0[e];
            for (int j = 0; j < q.m; ++j) v = fma(su[j], Gd[j * dd + e], v);
            if (G) G[e] = v;
            part += fabs(dt * v);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
        if (!isfinite(part)) bad = true;
        worst = fmax(worst, part);
    }
    if (bad) worst = INFINITY;
    if (lane == 0) wmax[warp] = worst;
    __syncthreads();
    if (threadIdx.x == 0) {
        double w = 0.0;
        for (int i = 0; i < nwarp; ++i) w = fmax(w, wmax[i]);
        if (!isfinite(w)) atomicExch(bad_flag, 1);
        else atomicMax(norm_max_bits, (unsigned long long)__double_as_longlong(w));
        if (unit_norm) unit_norm[unit] = w;
    }
}

// A = tau*G, padded panels XP / MP from Z / mu, per-unit scales.
__global__ void __launch_bounds__(256)
prep_unit_kernel(UnitGeom q, const double *Z, const double *mu, int u0_global, int s_levels,
                 const double *G, double *A, long long mat_stride,
                 double *XP, double *MP, long long pan_stride,
                 double *tau_out, double *neg_tau_out, const int *perm = nullptr) {
    const int unit = blockIdx.x, ug = perm ? perm[u0_global + unit] : u0_global + unit;
    const int sys = ug / q.nk, kl = ug % q.nk, k = q.k_begin + kl;
    const double *zk = Z + (long long)k * q.z_dim;
    const double tau = ldexp(zk[q.dt_off], -s_levels);
    const long long dd = q.gsz;
    const double *Gu = G + (long long)unit * mat_stride;
    double *Au = A + (long long)unit * mat_stride;
    for (long long e = threadIdx.x; e < dd; e += blockDim.x) Au[e] = tau * Gu[e];
    const double *x = zk + q.x_off[sys];
    const double *mk = mu ? mu + ((long long)sys * q.K + k) * ((long long)q.d * q.c) : nullptr;
    double *xp = XP + (long long)unit * pan_stride, *mp = MP ? MP + (long long)unit * pan_stride : nullptr;
    for (int e = threadIdx.x; e < q.dp * q.cp; e += blockDim.x) {
        int r = e % q.dp, b = e / q.dp;
        bool in = (r < q.d) && (b < q.c);
        xp[e] = in ? x[r + b * q.d] : 0.0;
        if (mp) mp[e] = (in && mk) ? mk[r + b * q.d] : 0.0;
    }
    if (threadIdx.x == 0) { tau_out[unit] = tau; neg_tau_out[unit] = -tau; }
}

// Hessian scalar blocks of one unit: HUU = -tau^2/2 (H + H^T), HUT_j = -(c1_j + tau c2_j), HTT = -c3
__global__ void finalize_scalars_kernel(int m, const double *tau, const double *H, const double *c1,
                                        const double *c2, const double *c3, double *harena,
                                        long long harena_stride, long long off_huu, double fac) {
    // fac = 2 in iso half storage: <S, L>_full = 2 <S_L, L_L>
    const int unit = blockIdx.x;
    const double t = tau[unit];
    double *out = harena + (long long)unit * harena_stride + off_huu;
    const double *Hu = H + (long long)unit * m * m;
    for (int e = threadIdx.x; e < m * m; e += blockDim.x) {
        int i = e / m, j = e % m;
        out[e] = -fac * t * t * (Hu[i * m + j] + Hu[j * m + i]);   // H = C + C^T (node contractions)
    }
    for (int j = threadIdx.x; j < m; j += blockDim.x)
        out[m * m + j] = -fac * (c1[(long long)unit * m + j] + t * c2[(long long)unit * m + j]);
    if (threadIdx.x == 0) { out[m * m + m] = -fac * c3[unit]; out[m * m + m + 1] = 0.0; }
}

// constants 0.0 / 1.0 at the tail of the Jacobian arena
__global__ void arena_consts_kernel(double *jarena, long long stride, long long off, int units) {
    int u = blockIdx.x * blockDim.x + threadIdx.x;
    if (u < units) { jarena[u * stride + off] = 0.0; jarena[u * stride + off + 1] = 1.0; }
}

// ---------------------------------------------------------------------------------------
// Output assembly: COO values in the caller's order, coalesced stores, gathers from the arena
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
assemble_kernel(const double *arena, long long arena_stride, const uint32_t *code, long long nnz,
                double *out, int u0_global, int nk, int chunks_per_unit, int out_u0, const int *perm = nullptr) {
    const int unit = blockIdx.x / chunks_per_unit, chunk = blockIdx.x % chunks_per_unit;
    const int sys = (perm ? perm[u0_global + unit] : u0_global + unit) / nk;
    const uint32_t *cd = code + (long long)sys * nnz;
    const double *ar = arena + (long long)unit * arena_stride;
    // out_u0: unit index of `out`'s first row; with perm the row is the unit's own position
    double *o = out + (long long)(perm ? perm[out_u0 + unit] : out_u0 + unit) * nnz;
    long long per = (nnz + chunks_per_unit - 1) / chunks_per_unit;
    per += per & 1;                                   // even chunk starts keep 16-byte alignment
    const long long b0 = chunk * per, e1 = (b0 + per < nnz) ? b0 + per : nnz;
    if ((nnz & 1) == 0 && (reinterpret_cast<uintptr_t>(o) & 15) == 0) {
        // two entries per thread, one 16-byte store (rows of `out` are then all 16-byte aligned)
        for (long long e = b0 + 2 * threadIdx.x; e < e1; e += 2 * blockDim.x) {
            const uint2 cc = *reinterpret_cast<const uint2 *>(cd + e);
            double v0 = ar[cc.x & 0x7fffffffu], v1 = ar[cc.y & 0x7fffffffu];
            if (cc.x >> 31) v0 = -v0;
            if (cc.y >> 31) v1 = -v1;
            *reinterpret_cast<double2 *>(o + e) = make_double2(v0, v1);
        }
    } else {
        for (long long e = b0 + threadIdx.x; e < e1; e += blockDim.x) {
            uint32_t cc = cd[e];
            double v = ar[cc & 0x7fffffffu];
            o[e] = (cc >> 31) ? -v : v;
        }
    }
}

// Run-based assembly.  The COO value vector of a unit is mostly long runs of consecutive arena
// entries (block-diagonal Jacobian: c copies of -E; the panels), so the host cuts the slot-code table
// into pieces (qcb200.cu: build_pieces): COPY pieces stream a run with 16-byte loads and stores and
// never touch the code table, GATHER pieces cover the short runs in between through the codes.
// One CTA per (unit, piece).  Measured 7.5 TB/s for pure HBM writes on this GPU; the per-entry gather
// of assemble_kernel reaches 4.4.
struct Piece { long long dst; unsigned src; unsigned len; unsigned kind; unsigned pad_; };   // kind 0 gather, 1 copy, 2 copy negated
__global__ void __launch_bounds__(256)
assemble_runs_kernel(const double *arena, long long arena_stride, const uint32_t *code, long long nnz,
                     const Piece *pieces, int npieces, double *out, int u0_global, int nk, int out_u0,
                     const int *perm, unsigned split = 0, int parts = 3) {
    const int unit = blockIdx.x / npieces, pi = blockIdx.x - unit * npieces;
    const int sys = (perm ? perm[u0_global + unit] : u0_global + unit) / nk;
    const Piece pc = pieces[(long long)sys * npieces + pi];
    if (pc.len == 0) return;
    if (parts != 3) {       // two-pass assembly: COPY pieces wholly inside arena[0, split) first, the rest later
        const bool head = pc.kind != 0 && pc.src + pc.len <= split;
        if (head != (parts == 1)) return;
    }
    const double *ar = arena + (long long)unit * arena_stride;
    double *o = out + (long long)(perm ? perm[out_u0 + unit] : out_u0 + unit) * nnz + pc.dst;
    if (pc.kind == 0) {
        const uint32_t *cd = code + (long long)sys * nnz + pc.dst;
        for (unsigned e = threadIdx.x; e < pc.len; e += blockDim.x) {
            const uint32_t cc = cd[e];
            const double v = ar[cc & 0x7fffffffu];
            o[e] = (cc >> 31) ? -v : v;
        }
        return;
    }
    const double *s = ar + pc.src;
    const bool neg = pc.kind == 2;
    if (((reinterpret_cast<uintptr_t>(o) | reinterpret_cast<uintptr_t>(s)) & 15) == 0) {
        const unsigned n2 = pc.len >> 1;
        const double2 *s2 = reinterpret_cast<const double2 *>(s);
        double2 *o2 = reinterpret_cast<double2 *>(o);
        constexpr int U = 4;
        for (unsigned e0 = threadIdx.x; e0 < n2; e0 += U * blockDim.x) {
            double2 v[U];
#pragma unroll
            for (int u = 0; u < U; ++u) { const unsigned e = e0 + u * blockDim.x; if (e < n2) v[u] = __ldg(s2 + e); }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const unsigned e = e0 + u * blockDim.x;
                if (e < n2) __stcs(o2 + e, neg ? make_double2(-v[u].x, -v[u].y) : v[u]);   // written once, never re-read here
            }
        }
        if ((pc.len & 1) && threadIdx.x == 0) o[pc.len - 1] = neg ? -s[pc.len - 1] : s[pc.len - 1];
    } else {
        for (unsigned e = threadIdx.x; e < pc.len; e += blockDim.x) o[e] = neg ? -s[e] : s[e];
    }
}

// delta = x_{k+1} - (E X)  with EX in a padded panel
__global__ void __launch_bounds__(256)
delta_kernel(UnitGeom q, const double *Z, const double *EX, long long pan_stride, double *delta, int u0_global,
             const int *perm = nullptr) {
    const int unit = blockIdx.x, ug = perm ? perm[u0_global + unit] : u0_global + unit;
    const int sys = ug / q.nk, k = q.k_begin + ug % q.nk;
    const double *xn = Z + (long long)(k + 1) * q.z_dim + q.x_off[sys];
    const double *ex = EX + (long long)unit * pan_stride;
    double *o = delta + (long long)ug * q.d * q.c;
    for (int e = threadIdx.x; e < q.d * q.c; e += blockDim.x) {
        int r = e % q.d, b = e / q.d;
        o[e] = xn[e] - ex[r + b * q.dp];
    }
}


// out[u][r + q*d] = arena[u*stride + r + q*dp]: the unpadded propagator of every unit of a chunk
__global__ void __launch_bounds__(256)
extract_E_kernel(const double *arena, long long stride, int dp, int d, double *out) {
    const double *E = arena + (long long)blockIdx.x * stride;
    double *o = out + (long long)blockIdx.x * d * d;
    for (int e = threadIdx.x; e < d * d; e += blockDim.x) o[e] = E[(e % d) + (long long)(e / d) * dp];
}

// Rollout: x_{k+1} = E_k x_k along the trajectory, one CTA per system, the state panel (d x c) in
// shared memory, E_k streamed from global memory (computed for all k beforehand, in parallel).
// states: [S][N][x_dim].  Thread t owns row r = t % d for the columns b = t / d, t / d + nb, ...
__global__ void __launch_bounds__(256)
rollout_chain_kernel(const double *E, const double *x_init, const double *Z, int z_dim, const int *x_off,
                     int d, int c, int K, double *states) {
    extern __shared__ double xs[];          // two panels of d*c
    const int s = blockIdx.x, dc = d * c;
    double *cur = xs, *nxt = xs + dc;
    double *out = states + (long long)s * (K + 1) * dc;
    for (int e = threadIdx.x; e < dc; e += blockDim.x) {
        const double v = x_init ? x_init[(long long)s * dc + e] : Z[x_off[s] + e];
        cur[e] = v; out[e] = v;
    }
    __syncthreads();
    for (int k = 0; k < K; ++k) {
        const double *Ek = E + ((long long)s * K + k) * d * d;
        for (int e = threadIdx.x; e < dc; e += blockDim.x) {
            const int r = e % d, b = e / d;
            double acc = 0.0;
            for (int q = 0; q < d; ++q) acc = fma(Ek[r + (long long)q * d], cur[q + b * d], acc);
            nxt[e] = acc;
            out[(long long)(k + 1) * dc + e] = acc;
        }
        __syncthreads();
        double *t = cur; cur = nxt; nxt = t;
    }
}

// Rollout from the scanned propagators: states[s][k + 1] = P[s][k] x0[s] with P[k] = E_k ... E_0.
// One CTA per (system, interval); P is read once (L2 / HBM), x0 comes from the trajectory or the caller.
__global__ void __launch_bounds__(256)
rollout_apply_kernel(const double *P, const double *x_init, const double *Z, const int *x_off, int d, int c, int K,
                     double *states) {
    const int s = blockIdx.x / K, k = blockIdx.x % K, dc = d * c;
    const double *Pk = P + ((long long)s * K + k) * d * d;
    const double *x0 = x_init ? x_init + (long long)s * dc : Z + x_off[s];
    double *out = states + ((long long)s * (K + 1) + k + 1) * dc;
    if (k == 0) for (int e = threadIdx.x; e < dc; e += blockDim.x) out[e - dc] = x0[e];      // knot 1: the initial state
    for (int e = threadIdx.x; e < dc; e += blockDim.x) {
        const int r = e % d, b = e / d;
        double a0 = 0.0, a1 = 0.0;
        int q = 0;
        for (; q + 1 < d; q += 2) {
            a0 = fma(Pk[r + (long long)q * d], x0[q + b * d], a0);
            a1 = fma(Pk[r + (long long)(q + 1) * d], x0[q + 1 + b * d], a1);
        }
        if (q < d) a0 = fma(Pk[r + (long long)q * d], x0[q + b * d], a0);
        out[e] = a0 + a1;
    }
}

}  // namespace qcb
