// qcb200.cu -- C ABI + host executor of the B200-native collocation evaluator.
// See include/qcb200.h for the contract and DESIGN.md for the architecture.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <functional>
#include <map>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include <immintrin.h>
#include <pthread.h>
#include <sched.h>

#include <cuda_runtime.h>

#include "../../include/qcb200.h"
#include "qcb_program.h"
#include "qcb_host_tuner.h"
#include "qcb_large.cuh"
#include "qcb_small.cuh"

using namespace qcb;

namespace {

thread_local std::string g_create_error;

struct CudaFail { cudaError_t e; const char *what; int line; };
#define CK(call)                                                         \
    do {                                                                 \
        cudaError_t e_ = (call);                                         \
        if (e_ != cudaSuccess) throw CudaFail{e_, #call, __LINE__};      \
    } while (0)

// the fused small-d kernel cannot hold a unit with this many squaring levels in shared memory
struct SmallCapacity { int squarings; };

struct Entry { int64_t row, col; uint32_t code; };

// Run-length form of a unit's slot-code table, used by the host-side expansion of the compact
// transfer (qcb_eval_all on the large-d path): copy / negated copy of `len` consecutive arena
// entries, or `len` repeats of one entry.
struct Seg { int64_t dst; uint32_t src; uint32_t len; uint32_t kind; };   // kind: bit0 negate, bit1 fill

inline int round_up(int v, int q) { return (v + q - 1) / q * q; }

}  // namespace

struct qcb_handle_s {
    // geometry
    int d = 0, c = 0, m = 0, N = 0, S = 1, K = 0, z_dim = 0, u_off = 0, dt_off = 0, jac_mode = 0;
    int k0 = 0, k1 = 0, nk = 0, units = 0, xdim = 0;
    int dp = 0, cp = 0;            // padded dims of the device matrices / panels
    bool large = false;
    bool iso = false;              // generators have the real-iso block form: half storage + half flops
    int gcols = 0;                 // stored columns per matrix (dp, or dp/2 when iso)
    int64_t msz = 0;               // doubles per stored matrix
    int device = 0, n_sm = 148;
    std::vector<int> x_off;
    std::string err;
    cudaStream_t stream = nullptr, copy_stream = nullptr, stream2 = nullptr;   // stream2: second lane of eval_large
    cudaEvent_t lane_ev[2] = {nullptr, nullptr};
    cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    std::vector<cudaEvent_t> chunk_events;
    int64_t launches = 0;
    double ms[4] = {0, 0, 0, 0};
    bool xfer_compact = false;
    int64_t xfer[2] = {0, 0};      // bytes moved host->device / device->host by the last host-buffer call
    // optional per-launch profiling of the batched DMMA GEMM (qcb_profile_*)
    bool profiling = false;
    std::vector<cudaEvent_t> prof_ev;     // pairs
    std::vector<double> prof_flops;       // algorithmic flops of each recorded launch
    size_t prof_used = 0;
    double prof_acc[4] = {0, 0, 0, 0};    // launches, total ms, total flops, reserved

    // device constants
    double *dG0 = nullptr, *dGd = nullptr;
    int *d_xoff = nullptr;
    uint32_t *d_jcode = nullptr, *d_hcode = nullptr;   // [S][nnz]
    int64_t jnnz_unit = 0, hnnz_unit = 0;
    int64_t JA = 0, HA = 0;                            // arena strides (doubles)
    bool custom_j = false, custom_h = false;
    unsigned long long *d_norm = nullptr;
    int *d_bad = nullptr;          // [0] flags (bit0 non-finite / out of range, bit1 chain capacity), [1] max squarings seen
    int *h_flags = nullptr;        // pinned host mirror of d_bad[0..1]
    int s_hint = 2;                // small-d path: chain capacity tried first (squarings of the last call)

    // staging for the host-buffer API
    double *dZ = nullptr, *dmu = nullptr;
    double *d_delta = nullptr, *d_jac = nullptr, *d_hess = nullptr;

    // large-path scratch
    int chunk_units = 0, s_cap = -1;
    double *sc_fixed = nullptr, *sc_E = nullptr, *sc_W = nullptr, *sc_dslot = nullptr, *sc_L = nullptr;   // sc_L: [s_cap + 1] levels
    double *sc_XP = nullptr, *sc_MP = nullptr, *sc_EX = nullptr, *sc_ja = nullptr, *sc_ha = nullptr;
    double *sc_H = nullptr, *sc_c1 = nullptr, *sc_c2 = nullptr, *sc_c3 = nullptr, *sc_tau = nullptr, *sc_ntau = nullptr;
    int small_threads = 0;
    double *d_Eall = nullptr, *d_Escan = nullptr, *d_states = nullptr, *d_xinit = nullptr;   // rollout buffers
    // sparse drive generators (ELL): products with G_j become gathers
    bool run_assembly = true;                 // run-based device assembly (QCB_GATHER_ASSEMBLY=1: per-entry gather)
    bool sparse_drives = false;
    int ell_w = 0;
    int *d_ell_idx = nullptr, *d_ellT_idx = nullptr;
    double *d_ell_val = nullptr, *d_ellT_val = nullptr;
    double *d_unit_norm = nullptr;            // [units] ||dt G(u)||_1 of the last norm pre-pass
    std::vector<double> h_unit_norm;
    int *d_perm = nullptr;                    // units sorted by squaring count (device-resident calls, S = 1)
    std::vector<int> h_perm, h_class_end;
    double *d_ja_all = nullptr, *d_ha_all = nullptr, *hs_ja = nullptr, *hs_ha = nullptr;   // small-d compact host route
    // compact host transfer: arenas (not expanded COO vectors) cross PCIe, the host cores expand
    std::vector<uint32_t> h_jcode, h_hcode;           // host copies of the code tables [S][nnz]
    std::vector<std::vector<Seg>> jsegs, hsegs;       // per system
    // Direct panels (host-buffer calls, S = 1): long plain-copy runs of the COO vectors whose source is used
    // nowhere else (the u / dt blocks of the Jacobian, the (x,u) / (x,dt) blocks of the Hessian) are DMA'd from
    // the arenas straight to their place in the caller's pinned buffer -- they neither pass through the staging
    // buffers nor occupy the cores; the staging copy shrinks to the arena ranges the rest still needs.
    struct Direct { int64_t dst; uint32_t src, len; };
    std::vector<Direct> jdirect, hdirect;
    std::vector<std::vector<Seg>> jsegs_x, hsegs_x;   // expansion lists without the direct runs
    std::vector<std::pair<uint32_t, uint32_t>> jstage, hstage;   // arena ranges [lo, hi) still staged
    bool direct_built = false;
    Piece *d_jpieces = nullptr, *d_hpieces = nullptr; // run-based device assembly: [S][n*pieces]
    int n_jpieces = 0, n_hpieces = 0;
    double *sc_ja2 = nullptr, *sc_ha2 = nullptr;      // second arena set (double buffering)
    double *stage_j[2] = {nullptr, nullptr}, *stage_h[2] = {nullptr, nullptr};   // pinned
    int stage_units = 0;
    double *part_j[2] = {nullptr, nullptr}, *part_h[2] = {nullptr, nullptr};   // device COO rows of the units sent directly
    int part_units = 0;
    // share of each chunk assembled on the device and DMA'd as full COO rows (QCB_DIRECT_FRACTION).
    // Measured on the B200 box: 0 is best -- DMA writes and the cores' streaming stores share the same
    // host DRAM write bandwidth (~107 GB/s), so splitting the work between them gains nothing.
    double direct_fraction = 0.0;
    cudaEvent_t copied[2] = {nullptr, nullptr}, part_done[2] = {nullptr, nullptr};

    // What the scratch still holds of the last evaluation's Z (new_x = 0 calls reuse it): bit0 the
    // propagator chain (G, A, powers, E levels, E X), bit1 the forward tangents (per-drive registers,
    // L_j at every level, G E, the Jacobian arena).  Valid only for device-resident whole-slab calls.
    int cache_have = 0;
    int cache_s_slab = -1;
    int reuse = 0;                 // set by the *_x entry points for the call in flight
    bool dZ_valid = false;         // the staged device copy of Z is the one the last host-buffer call uploaded
    HostTuner tune[2];             // expansion threads of host-buffer calls that deliver the Jacobian: [0] without, [1] with the Hessian

    // arena delivery (qcb_eval_arenas_device): where the per-unit value arenas of the slab are written
    // instead of being assembled into COO vectors here; may be another GPU's memory
    double *arena_out_j = nullptr, *arena_out_h = nullptr;
    cudaEvent_t ship_ev[3] = {nullptr, nullptr, nullptr};   // last delivery of [0] the E part, [1] the rest of the Jacobian arena, [2] the Hessian arena
    int delivery_split = 1;        // QCB_DELIVERY_SPLIT: sequential pieces per slab in a multi-GPU delivery
    int arena_parts = 3;           // qcb_set_arena_parts: 1 = propagators only (E part), 2 = the rest of the Jacobian arena, 3 = all
    // Asynchronous mode (qcb_set_async): device-pointer calls return once their work is queued; completion
    // and the range check are collected by qcb_wait.  Lets a caller queue the Jacobian and the Hessian side
    // of one iterate back to back and overlap the delivery of the first with the computation of the second.
    bool async_mode = false;

    // State index map (variational integrators, src/quantum_integrators.jl:140-169: the state is a
    // concatenation of several named components).  When the caller gives one, the kernels work on a
    // packed copy of Z ([all panels | u | dt] per knot -- z_dim, x_off, u_off, dt_off above then
    // describe that packed knot) and `ext` keeps the caller's layout for the structure arrays, the
    // gather of Z / mu on the way in and the row order of delta on the way out.
    struct Ext {
        bool on = false, row_identity = true;
        int z_dim = 0, u_off = 0, dt_off = 0;
        std::vector<int> xmap, rmap;               // [S][xdim] knot offsets; [xdim] defect rows
        int *d_xmap = nullptr, *d_rmap = nullptr;
        double *dZ_ext = nullptr, *dZp = nullptr, *dmup = nullptr, *d_delta_p = nullptr;
    } ext;
    int64_t ez() const { return ext.on ? ext.z_dim : z_dim; }
    int64_t eu() const { return ext.on ? ext.u_off : u_off; }
    int64_t et() const { return ext.on ? ext.dt_off : dt_off; }
    int64_t ecol_x(int s, int q) const { return ext.on ? ext.xmap[(size_t)s * xdim + q] : x_off[s] + q; }
    int64_t erow(int r) const { return ext.on ? ext.rmap[r] : r; }

    UnitGeom geom() const {
        UnitGeom q;
        q.d = d; q.c = c; q.m = m; q.dp = dp; q.cp = cp; q.z_dim = z_dim; q.u_off = u_off; q.dt_off = dt_off;
        q.K = K; q.nk = nk; q.k_begin = k0; q.S = S; q.x_off = d_xoff; q.G0 = dG0; q.Gd = dGd;
        q.gcols = gcols; q.gsz = msz;
        return q;
    }
};

namespace {

// ---------------------------------------------------------------------------------------
// Native structure patterns and slot codes
// ---------------------------------------------------------------------------------------
struct Codes {
    qcb_handle_s *h;
    int64_t dd() const { return (int64_t)h->dp * h->dp; }      // E slot of the Jacobian arena (always full)
    int64_t pan() const { return (int64_t)h->dp * h->cp; }
    uint32_t jE(int rr, int qq) const { return (uint32_t)(rr + (int64_t)qq * h->dp) | QCB_CODE_NEG; }   // -E(rr, qq)
    uint32_t jU(int j, int i) const { return (uint32_t)(dd() + j * pan() + (i % h->d) + (int64_t)(i / h->d) * h->dp); }
    uint32_t jT(int i) const { return jU(h->m, i); }
    uint32_t jZero() const { return (uint32_t)(dd() + (h->m + 1) * pan()); }
    uint32_t jOne() const { return jZero() + 1; }
    uint32_t hXU(int j, int i) const { return (uint32_t)(j * pan() + (i % h->d) + (int64_t)(i / h->d) * h->dp); }
    uint32_t hXT(int i) const { return hXU(h->m, i); }
    uint32_t hUU(int i, int j) const { return (uint32_t)((h->m + 1) * pan() + i * h->m + j); }
    uint32_t hUT(int j) const { return (uint32_t)((h->m + 1) * pan() + h->m * h->m + j); }
    uint32_t hTT() const { return hUT(h->m); }
    uint32_t hZero() const { return hTT() + 1; }
};

// per-unit native Jacobian pattern of system s: local row, local col in [z_k; z_{k+1}], code;
// sorted by (col, row) -- the order Julia's findnz gives for a sparse block.
std::vector<Entry> jac_pattern(qcb_handle_s *h, int s) {
    Codes cd{h};
    const int d = h->d, xd = h->xdim, m = h->m;
    std::vector<Entry> v;
    v.reserve((size_t)h->jnnz_unit);
    for (int q = 0; q < xd; ++q) {
        const int b = q / d, qq = q % d;
        const int64_t col = h->ecol_x(s, q);
        if (h->jac_mode == QCB_JAC_DENSE_X) {
            for (int r = 0; r < xd; ++r)
                v.push_back({h->erow(r), col, (r / d == b) ? cd.jE(r % d, qq) : cd.jZero()});
        } else {
            for (int rr = 0; rr < d; ++rr) v.push_back({h->erow(b * d + rr), col, cd.jE(rr, qq)});
        }
    }
    for (int j = 0; j < m; ++j)
        for (int r = 0; r < xd; ++r) v.push_back({h->erow(r), h->eu() + j, cd.jU(j, r)});
    for (int r = 0; r < xd; ++r) v.push_back({h->erow(r), h->et(), cd.jT(r)});
    for (int r = 0; r < xd; ++r) v.push_back({h->erow(r), h->ez() + h->ecol_x(s, r), cd.jOne()});
    std::stable_sort(v.begin(), v.end(), [](const Entry &a, const Entry &b) {
        return a.col != b.col ? a.col < b.col : a.row < b.row;
    });
    return v;
}

std::vector<Entry> hess_pattern(qcb_handle_s *h, int s) {
    Codes cd{h};
    const int xd = h->xdim, m = h->m;
    std::vector<Entry> v;
    v.reserve((size_t)h->hnnz_unit);
    auto add = [&](int64_t a, int64_t b, uint32_t code) { v.push_back({std::min(a, b), std::max(a, b), code}); };
    for (int j = 0; j < m; ++j)
        for (int i = 0; i < xd; ++i) add(h->ecol_x(s, i), h->eu() + j, cd.hXU(j, i));
    for (int i = 0; i < xd; ++i) add(h->ecol_x(s, i), h->et(), cd.hXT(i));
    for (int i = 0; i < m; ++i)
        for (int j = i; j < m; ++j) add(h->eu() + i, h->eu() + j, cd.hUU(i, j));
    for (int j = 0; j < m; ++j) add(h->eu() + j, h->et(), cd.hUT(j));
    add(h->et(), h->et(), cd.hTT());
    std::stable_sort(v.begin(), v.end(), [](const Entry &a, const Entry &b) {
        return a.col != b.col ? a.col < b.col : a.row < b.row;
    });
    return v;
}

std::vector<Seg> build_segments(const uint32_t *code, int64_t nnz) {
    std::vector<Seg> v;
    int64_t e = 0;
    while (e < nnz) {
        const uint32_t c0 = code[e], off0 = c0 & 0x7fffffffu, neg = c0 >> 31;
        int64_t len = 1;
        uint32_t kind = neg;
        if (e + 1 < nnz && code[e + 1] == c0) {                       // repeats of one entry
            kind |= 2u;
            while (e + len < nnz && code[e + len] == c0 && len < 0x7fffffff) ++len;
        } else {
            while (e + len < nnz && (code[e + len] >> 31) == neg &&
                   (code[e + len] & 0x7fffffffu) == off0 + (uint32_t)len && len < 0x7fffffff) ++len;
        }
        v.push_back({e, off0, (uint32_t)len, kind});
        e += len;
    }
    return v;
}

// Pieces of at most PIECE entries (QCB_PIECE, default 8192; flat between 4096 and 32768) for assemble_runs_kernel: runs of >= 256 consecutive arena entries
// become COPY pieces, everything between them is covered by GATHER pieces (through the code table).
std::vector<Piece> build_pieces(const uint32_t *code, int64_t nnz) {
    int64_t PIECE = 8192;
    const int64_t MIN_RUN = 256;
    if (const char *e = getenv("QCB_PIECE")) PIECE = std::max(512L, atol(e));
    std::vector<Piece> out;
    int64_t gather_from = -1;
    auto flush_gather = [&](int64_t upto) {
        for (int64_t e = gather_from; gather_from >= 0 && e < upto; e += PIECE)
            out.push_back({e, 0u, (unsigned)std::min(PIECE, upto - e), 0u, 0u});
        gather_from = -1;
    };
    for (const Seg &sg : build_segments(code, nnz)) {
        if (!(sg.kind & 2u) && sg.len >= MIN_RUN) {
            flush_gather(sg.dst);
            for (int64_t q = 0; q < sg.len; q += PIECE)
                out.push_back({sg.dst + q, sg.src + (unsigned)q, (unsigned)std::min<int64_t>(PIECE, sg.len - q), (sg.kind & 1u) ? 2u : 1u, 0u});
        } else if (gather_from < 0) {
            gather_from = sg.dst;
        }
    }
    flush_gather(nnz);
    return out;
}

void upload_pieces(qcb_handle_s *h, const std::vector<uint32_t> &codes, int64_t nnz, Piece **dptr, int *count) {
    std::vector<std::vector<Piece>> per(h->S);
    size_t mx = 0;
    for (int s = 0; s < h->S; ++s) { per[s] = build_pieces(codes.data() + (size_t)s * nnz, nnz); mx = std::max(mx, per[s].size()); }
    std::vector<Piece> flat((size_t)h->S * mx, Piece{0, 0u, 0u, 0u, 0u});
    for (int s = 0; s < h->S; ++s) std::copy(per[s].begin(), per[s].end(), flat.begin() + (size_t)s * mx);
    if (*dptr) { CK(cudaFree(*dptr)); *dptr = nullptr; }
    if (mx) {
        CK(cudaMalloc(dptr, flat.size() * sizeof(Piece)));
        CK(cudaMemcpy(*dptr, flat.data(), flat.size() * sizeof(Piece), cudaMemcpyHostToDevice));
    }
    *count = (int)mx;
}

// dst[0..len) = +-src[0..len) with non-temporal 64-byte stores (the destination is written once
// and never read by the library; streaming stores avoid the read-for-ownership traffic)
__attribute__((target("avx512f"))) void copy_run_avx512(double *dst, const double *src, size_t len, bool neg) {
    const __m512i sgn = neg ? _mm512_set1_epi64((long long)0x8000000000000000ull) : _mm512_setzero_si512();
    size_t i = 0;
    while (i < len && (reinterpret_cast<uintptr_t>(dst + i) & 63)) { dst[i] = neg ? -src[i] : src[i]; ++i; }
    for (; i + 8 <= len; i += 8) _mm512_stream_pd(dst + i, _mm512_castsi512_pd(_mm512_xor_si512(_mm512_castpd_si512(_mm512_loadu_pd(src + i)), sgn)));
    for (; i < len; ++i) dst[i] = neg ? -src[i] : src[i];
}
void copy_run_plain(double *dst, const double *src, size_t len, bool neg) {
    if (neg) for (size_t i = 0; i < len; ++i) dst[i] = -src[i];
    else memcpy(dst, src, len * sizeof(double));
}
__attribute__((target("avx512f"))) void fill_run_avx512(double *dst, double v, size_t len) {
    const __m512d vv = _mm512_set1_pd(v);
    size_t i = 0;
    while (i < len && (reinterpret_cast<uintptr_t>(dst + i) & 63)) dst[i++] = v;
    for (; i + 8 <= len; i += 8) _mm512_stream_pd(dst + i, vv);
    for (; i < len; ++i) dst[i] = v;
}

// Persistent worker threads for the host-side expansion (created once per process, shared by all
// handles): run(fn) executes fn() on every worker and on the caller, returns when all are done.
// Size: QCB_HOST_THREADS, else OMP_NUM_THREADS (unless it is the launcher's default of 1), else the cores this process may run on divided by
// LOCAL_WORLD_SIZE (one process per GPU on one host: N ranks must not each start a thread per core).
// Workers are pinned one per core, rank r taking the r-th block of the allowed cores, so that the
// ranks' streaming stores do not migrate between cores (QCB_HOST_PIN=0 disables pinning).
// run() is serialised by a mutex: handles driven from different host threads take turns.
// Index of the calling thread inside the pool (0: the caller of run()/start(), i.e. any non-pool thread) and the
// number of pool threads an expansion job may use (0: all).  The limit exists because on some hosts the PCIe
// writes that run beside the expansion get starved by sixteen cores of streaming stores, and the whole call is
// faster with fewer store threads; it is chosen per handle by measurement (HostTuner).
thread_local int tl_pool_id = 0;
std::atomic<int> g_expand_limit{0};

class HostPool {
public:
    static HostPool &get() { static HostPool p; return p; }
    int size() const { return (int)workers_.size() + 1; }
    void run(const std::function<void()> &fn) { start(fn); finish(); }
    // start(): the workers begin fn() and the caller goes on (fn must stay alive until finish());
    // finish(): the caller joins the work, then waits for the workers.  One job at a time per process.
    void start(const std::function<void()> &fn) {
        run_m_.lock();
        {
            std::lock_guard<std::mutex> lk(m_);
            fn_ = &fn; pending_ = (int)workers_.size(); ++epoch_;
        }
        cv_.notify_all();
    }
    void finish() {
        (*fn_)();
        {
            std::unique_lock<std::mutex> lk(m_);
            done_.wait(lk, [&] { return pending_ == 0; });
            fn_ = nullptr;
        }
        run_m_.unlock();
    }
private:
    static int env_int(const char *name) { const char *e = getenv(name); return e ? atoi(e) : 0; }
    HostPool() {
        std::vector<int> cores;
        cpu_set_t set;
        CPU_ZERO(&set);
        if (sched_getaffinity(0, sizeof(set), &set) == 0)
            for (int c = 0; c < CPU_SETSIZE; ++c) if (CPU_ISSET(c, &set)) cores.push_back(c);
        if (cores.empty()) for (unsigned c = 0; c < std::max(1u, std::thread::hardware_concurrency()); ++c) cores.push_back((int)c);
        const int world = std::max(1, env_int("LOCAL_WORLD_SIZE")), rank = std::max(0, env_int("LOCAL_RANK")) % world;
        // OMP_NUM_THREADS = 1 next to LOCAL_WORLD_SIZE is what torch.distributed.run exports by default for
        // every rank, not a request for a serial expansion (measured: 4.4 k instead of 11 k knot-evals/s on
        // two ranks): only an explicit QCB_HOST_THREADS, or an OMP_NUM_THREADS > 1, overrides the share of cores.
        int n = env_int("QCB_HOST_THREADS");
        if (n <= 0) { n = env_int("OMP_NUM_THREADS"); if (n == 1 && getenv("LOCAL_WORLD_SIZE")) n = 0; }
        if (n <= 0) n = std::max(1, (int)cores.size() / world);
        n = std::max(1, std::min(n, (int)cores.size()));
        const char *pin_env = getenv("QCB_HOST_PIN");
        const bool pin = !(pin_env && atoi(pin_env) == 0) && (int)cores.size() >= n;
        const int first = (int)(((long long)rank * n) % (long long)cores.size());
        for (int i = 1; i < n; ++i) {
            workers_.emplace_back([this, i] { tl_pool_id = i; loop(); });
            if (pin) {
                cpu_set_t one;
                CPU_ZERO(&one);
                CPU_SET(cores[(size_t)(first + i) % cores.size()], &one);
                pthread_setaffinity_np(workers_.back().native_handle(), sizeof(one), &one);
            }
        }
    }
    ~HostPool() {
        { std::lock_guard<std::mutex> lk(m_); stop_ = true; ++epoch_; }
        cv_.notify_all();
        for (auto &w : workers_) w.join();
    }
    void loop() {
        unsigned long long seen = 0;
        for (;;) {
            const std::function<void()> *fn;
            {
                std::unique_lock<std::mutex> lk(m_);
                cv_.wait(lk, [&] { return epoch_ != seen; });
                seen = epoch_;
                if (stop_) return;
                fn = fn_;
            }
            (*fn)();
            { std::lock_guard<std::mutex> lk(m_); if (--pending_ == 0) done_.notify_one(); }
        }
    }
    std::vector<std::thread> workers_;
    std::mutex m_, run_m_;
    std::condition_variable cv_, done_;
    const std::function<void()> *fn_ = nullptr;
    int pending_ = 0;
    unsigned long long epoch_ = 0;
    bool stop_ = false;
};

// One expansion job: `nu` units of one COO vector from arena rows staged in host memory.
struct ExpandJob {
    const std::vector<std::vector<Seg>> *segs;
    const double *stage;
    int64_t arena_stride;
    double *out;
    int64_t nnz;
};

// Expand units [u0, u0 + nu) of every job: out[(u0 + u) * nnz + seg.dst ...] = +-arena[seg.src ...].
// A task can run in the background: start() hands it to the pool's workers and returns, so that the calling
// thread queues the next chunk's kernels meanwhile; finish() lets the caller help and waits for the rest.
struct ExpandTask {
    ExpandJob jobs[2];
    int njobs = 0, nk = 1, u0 = 0, nu = 0;
    std::atomic<int> next{0};
    std::function<void()> work;
    bool running = false;
    ExpandTask() {
        work = [this]() {
            static const bool has512 = __builtin_cpu_supports("avx512f");
            const int limit = g_expand_limit.load(std::memory_order_relaxed);
            if (limit > 0 && tl_pool_id >= limit) return;
            const int total = nu * njobs;
            for (;;) {
                const int w = next.fetch_add(1);
                if (w >= total) break;
                const ExpandJob &jb = jobs[w / nu];
                const int u = w % nu;
                const std::vector<Seg> &sg = (*jb.segs)[(size_t)((u0 + u) / nk)];
                const double *ar = jb.stage + (size_t)u * jb.arena_stride;
                double *o = jb.out + (size_t)(u0 + u) * jb.nnz;
                for (const Seg &q : sg) {
                    const bool neg = q.kind & 1u;
                    if (q.kind & 2u) {
                        const double v = neg ? -ar[q.src] : ar[q.src];
                        if (has512) fill_run_avx512(o + q.dst, v, q.len);
                        else for (uint32_t i = 0; i < q.len; ++i) o[q.dst + i] = v;
                    } else if (has512) copy_run_avx512(o + q.dst, ar + q.src, q.len, neg);
                    else copy_run_plain(o + q.dst, ar + q.src, q.len, neg);
                }
            }
            if (has512) _mm_sfence();
        };
    }
    void start(const ExpandJob *jb, int nj, int nk_, int u0_, int nu_) {
        for (int i = 0; i < nj; ++i) jobs[i] = jb[i];
        njobs = nj; nk = nk_; u0 = u0_; nu = nu_;
        next.store(0);
        running = true;
        HostPool::get().start(work);
    }
    void finish() {
        if (!running) return;
        HostPool::get().finish();
        running = false;
    }
    ~ExpandTask() { finish(); }     // an exception on the way out must not leave the workers with a dead job
};
void expand_units(const ExpandJob *jobs, int njobs, int nk, int u0, int nu) {
    ExpandTask t;
    t.start(jobs, njobs, nk, u0, nu);
    t.finish();
}

void upload_codes(qcb_handle_s *h, uint32_t **dptr, const std::vector<uint32_t> &codes) {
    if (*dptr) { CK(cudaFree(*dptr)); *dptr = nullptr; }
    CK(cudaMalloc(dptr, codes.size() * sizeof(uint32_t)));
    CK(cudaMemcpy(*dptr, codes.data(), codes.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
}

void build_native_codes(qcb_handle_s *h, bool jac, bool hess) {
    if (jac) {
        std::vector<uint32_t> codes;
        for (int s = 0; s < h->S; ++s)
            for (auto &e : jac_pattern(h, s)) codes.push_back(e.code);
        upload_codes(h, &h->d_jcode, codes);
        h->custom_j = false;
        h->h_jcode = codes;
    }
    if (hess) {
        std::vector<uint32_t> codes;
        for (int s = 0; s < h->S; ++s)
            for (auto &e : hess_pattern(h, s)) codes.push_back(e.code);
        upload_codes(h, &h->d_hcode, codes);
        h->custom_h = false;
        h->h_hcode = codes;
    }
}

// ---------------------------------------------------------------------------------------
// Launch helpers
// ---------------------------------------------------------------------------------------
template <int BM, int BN, int WM, int WN, int STAGES>
void launch_gemm_cfg(qcb_handle_s *h, const GemmArgs &g, int batch, cudaStream_t st) {
    static bool attr_set[64] = {false};
    auto kern = bgemm_kernel<BM, BN, WM, WN, STAGES>;
    const size_t smem = GemmSmem<BM, BN>::bytes(STAGES);
    if (!attr_set[h->device]) {
        CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_set[h->device] = true;
    }
    const int tiles = (g.M / BM) * (g.N / BN);
    const int threads = (BM / WM) * (BN / WN) * 32;
    const bool prof = h->profiling;
    if (prof) {
        if (h->prof_used + 2 > h->prof_ev.size()) {
            for (int i = 0; i < 2; ++i) { cudaEvent_t e; CK(cudaEventCreate(&e)); h->prof_ev.push_back(e); }
        }
        CK(cudaEventRecord(h->prof_ev[h->prof_used], st));
    }
    kern<<<(unsigned)(batch * tiles), threads, smem, st>>>(g);
    if (prof) {
        CK(cudaEventRecord(h->prof_ev[h->prof_used + 1], st));
        h->prof_flops.push_back(2.0 * g.M * g.N * g.K * g.nterm * batch);
        h->prof_used += 2;
    }
    ++h->launches;
}

void launch_gemm_iso(qcb_handle_s *h, const GemmArgs &g, int units, cudaStream_t st) {
    static bool attr_set[64] = {false};
#ifndef QCB_ISO_BK
#define QCB_ISO_BK 16
#endif
    constexpr int IBK = QCB_ISO_BK, ST = IBK == 16 ? 3 : 2;      // 16-deep slabs x 3 stages (92 KB), or 32-deep x 2 (98 KB)
    auto kern = bgemm_iso_kernel<ST, IBK>;
    const size_t smem = GemmSmem<128, 64, IBK>::bytes(ST);
    if (!attr_set[h->device]) {
        CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_set[h->device] = true;
    }
    const int batch = units * g.ndir;
    const int tiles = (g.M / 128) * (g.N / 64);
    const bool prof = h->profiling;
    if (prof) {
        if (h->prof_used + 2 > h->prof_ev.size())
            for (int i = 0; i < 2; ++i) { cudaEvent_t e; CK(cudaEventCreate(&e)); h->prof_ev.push_back(e); }
        CK(cudaEventRecord(h->prof_ev[h->prof_used], st));
    }
    kern<<<(unsigned)(batch * tiles), 256, smem, st>>>(g);
    if (prof) {
        CK(cudaEventRecord(h->prof_ev[h->prof_used + 1], st));
        h->prof_flops.push_back(2.0 * g.M * g.N * g.K * g.nterm * batch);
        h->prof_used += 2;
    }
    ++h->launches;
    CK(cudaGetLastError());
}

void launch_gemm(qcb_handle_s *h, GemmArgs &g, int units, cudaStream_t st) {
    const int batch = units * g.ndir;
    const bool big_ok = (g.M % 128 == 0) && (g.N % 128 == 0);
    if (big_ok && (long long)batch * (g.M / 128) * (g.N / 128) >= 2LL * h->n_sm)
        launch_gemm_cfg<128, 128, 32, 32, 4>(h, g, batch, st);
    else
        launch_gemm_cfg<64, 64, 32, 32, 4>(h, g, batch, st);
    CK(cudaGetLastError());
}

struct LargeExec {
    qcb_handle_s *h;
    int nu;        // units in this chunk
    int u0;        // slab-local index of the chunk's first unit
    int s;         // squarings (uniform)
    int level = 0;
    long long dd, pan;
    double *ja = nullptr, *ha = nullptr;   // value arenas (BASE pointers of the scratch arrays: see ubase)
    // Lane support: squaring classes run concurrently on two streams, each on its own slice of the
    // scratch arrays.  Every array has a per-unit stride, so a lane is just an offset of `ubase` units:
    // mat() adds it for everything addressed through operands, raw kernel arguments use at().
    int ubase = 0;
    cudaStream_t st = nullptr;

    Opnd mat(double *base, long long su, long long sd = 0, long long ss = 0) const {
        Opnd o; o.p = base + (long long)ubase * su; o.su = su; o.sd = sd; o.ss = ss; o.ld = h->dp; o.pad_ = 0; return o;
    }
    double *at(double *base, long long su) const { return base + (long long)ubase * su; }
    Opnd E(int lvl) const {
        if (lvl == s) return mat(ja, h->JA);
        return mat(h->sc_E + (long long)lvl * h->chunk_units * dd, dd);
    }
    Opnd W(int lvl) const { return mat(h->sc_W + (long long)lvl * h->chunk_units * dd, dd); }
    Opnd fixed(int r) const { return mat(h->sc_fixed + (long long)r * h->chunk_units * dd, dd); }
    // L_j of every squaring level stays (a later Hessian call on the same Z contracts against each of them)
    Opnd Llev(int lvl) const { return mat(h->sc_L + (long long)lvl * h->chunk_units * h->m * dd, (long long)h->m * dd, dd); }
    Opnd reg(int r) const {
        if (r < QR_NUM_FIXED) return fixed(r);
        switch (r) {
        case QR_EL: return E(level);
        case QR_EL1: return E(level + 1);
        case QR_WL: return W(level);
        case QR_WL1: return W(level + 1);
        case QR_E0: return E(0);
        case QR_W0: return W(0);
        case QR_GJ: return mat(h->dGd, 0, dd, (long long)h->m * dd);
        case QR_L: return Llev(level);
        case QR_LN: return Llev(level + 1);
        default: return mat(h->sc_dslot + (long long)qcb_dslot(r) * h->chunk_units * h->m * dd, (long long)h->m * dd, dd);
        }
    }
    GemmArgs base_args(int ndir) const {
        GemmArgs g;
        memset(&g, 0, sizeof(g));
        g.M = g.K = h->dp;
        g.N = h->iso ? h->dp / 2 : h->dp;
        g.d_ident = h->iso ? h->d / 2 : h->d;   // iso half storage: I_L = [I_n; 0]
        g.alpha = 1.0;
        g.ndir = ndir;
        g.unit0_global = u0;
        g.nk = h->nk;
        return g;
    }
    void run(const qcb_op &op, int ndir) {
        GemmArgs g = base_args(ndir);
        g.nterm = op.nterm;
        g.nlin = op.nlin;
        for (int t = 0; t < op.nterm; ++t) {
            g.term[t].a = reg(op.term[t].a);
            g.term[t].b = reg(op.term[t].b);
            g.term[t].ta = op.term[t].ta;
            g.term[t].tb = op.term[t].tb;
        }
        for (int l = 0; l < op.nlin; ++l) {
            g.lin[l] = reg(op.lin[l].reg);
            g.lin_coef[l] = QCB_TAYLOR[op.lin[l].coef];
        }
        g.ident = op.ident >= 0 ? QCB_TAYLOR[op.ident] : 0.0;
        g.dst = reg(op.dst);
        if (h->sparse_drives && op.nterm > 0) {
            // terms with the drive generator as a factor are gathers (sparse_drive_kernel -> dst);
            // the remaining dense terms are added by the GEMM through a linear term reading dst back
            SparseArgs sa;
            memset(&sa, 0, sizeof(sa));
            int nd = 0, nleft = 0, nright = 0;
            for (int t = 0; t < op.nterm; ++t) {
                const bool ga = op.term[t].a == QR_GJ, gb = op.term[t].b == QR_GJ;
                if (ga && !op.term[t].tb && nleft == 0) {                    // G_j X  or  G_j^T X
                    sa.xl = g.term[t].b;
                    sa.l_idx = op.term[t].ta ? h->d_ellT_idx : h->d_ell_idx;
                    sa.l_val = op.term[t].ta ? h->d_ellT_val : h->d_ell_val;
                    ++nleft;
                } else if (gb && !op.term[t].ta && !op.term[t].tb && nright == 0) {   // X G_j
                    sa.xr = g.term[t].a;
                    sa.r_idx = h->d_ellT_idx; sa.r_val = h->d_ellT_val;
                    ++nright;
                } else {
                    g.term[nd++] = g.term[t];
                }
            }
            // forms of the program tables: one left term, optionally with the right term (dA2 = G A + A G)
            if (nleft == 1 && op.nlin + (nd > 0 ? 1 : 0) <= QCB_MAX_LIN) {
                sa.dst = g.dst; sa.d = h->dp; sa.ncols = g.N; sa.iso = h->iso ? 1 : 0; sa.m = h->m;
                sa.ndir = ndir; sa.unit0_global = u0; sa.nk = h->nk;
                const dim3 sgrid((unsigned)(nu * ndir)), sblock(64, 4);
#define QCB_SPARSE_LAUNCH(W_)                                                                      \
    if (nright) sparse_drive_kernel<W_, true><<<sgrid, sblock, 0, st>>>(sa);              \
    else sparse_drive_kernel<W_, false><<<sgrid, sblock, 0, st>>>(sa)
                switch (h->ell_w) {
                case 1: QCB_SPARSE_LAUNCH(1); break;
                case 2: QCB_SPARSE_LAUNCH(2); break;
                case 4: QCB_SPARSE_LAUNCH(4); break;
                default: QCB_SPARSE_LAUNCH(8); break;
                }
#undef QCB_SPARSE_LAUNCH
                ++h->launches;
                CK(cudaGetLastError());
                if (nd == 0 && op.nlin == 0 && op.ident < 0) return;     // the op was sparse only
                g.lin[g.nlin] = g.dst; g.lin_coef[g.nlin] = 1.0; ++g.nlin;
                g.nterm = nd;
            } else {
                for (int t = 0; t < op.nterm; ++t) {      // not applicable: restore the dense form
                    g.term[t].a = reg(op.term[t].a); g.term[t].b = reg(op.term[t].b);
                    g.term[t].ta = op.term[t].ta; g.term[t].tb = op.term[t].tb;
                }
            }
        }
        if (g.nterm == 0) {        // pure linear combination: no tensor work
            lincomb_kernel<<<(unsigned)(nu * ndir * (g.N / 8)), 256, 0, st>>>(g);
            ++h->launches;
            CK(cudaGetLastError());
            return;
        }
        if (h->iso) launch_gemm_iso(h, g, nu, st); else launch_gemm(h, g, nu, st);
    }
    // dst = alpha * [scale] * op(a) * op(b), general shapes.  In iso mode the left operand is a
    // half-stored matrix (expanded on the fly) unless `plain` says both operands are ordinary.
    void product(Opnd dst, Opnd a, int ta, Opnd b, int tb, int M, int N, int Kd, int ndir, double alpha,
                 const double *unit_scale, bool plain = false) {
        GemmArgs g = base_args(ndir);
        g.M = M; g.N = N; g.K = Kd;
        g.d_ident = 0;
        g.nterm = 1;
        g.term[0].a = a; g.term[0].b = b; g.term[0].ta = ta; g.term[0].tb = tb;
        g.alpha = alpha;
        g.unit_scale = unit_scale;
        g.dst = dst;
        if (h->iso && !plain) launch_gemm_iso(h, g, nu, st); else launch_gemm(h, g, nu, st);
    }
    void contract(Opnd P, int nj, Opnd Q, int ni, double *out, int accumulate) {
        ContractArgs a;
        a.P = P; a.Q = Q; a.nj = nj; a.ni = ni; a.nelem = dd; a.out = out; a.accumulate = accumulate;
        a.unit0_global = u0; a.nk = h->nk;
        if (nj <= 16 && ni <= 16 && (dd % 128) == 0)
            contract_mma_kernel<<<(unsigned)nu, 256, 0, st>>>(a);
        else
            contract_kernel<<<(unsigned)(nu * nj), 256, 0, st>>>(a);
        ++h->launches;
        CK(cudaGetLastError());
    }
};

void free_scratch(qcb_handle_s *h) {
    h->cache_have = 0;
    double **ptrs[] = {&h->sc_fixed, &h->sc_E, &h->sc_W, &h->sc_dslot, &h->sc_L, &h->sc_XP, &h->sc_MP,
                       &h->sc_EX, &h->sc_ja, &h->sc_ha, &h->sc_H, &h->sc_c1, &h->sc_c2, &h->sc_c3, &h->sc_tau, &h->sc_ntau};
    for (auto p : ptrs)
        if (*p) { cudaFree(*p); *p = nullptr; }
    if (h->sc_ja2) { cudaFree(h->sc_ja2); h->sc_ja2 = nullptr; }
    if (h->sc_ha2) { cudaFree(h->sc_ha2); h->sc_ha2 = nullptr; }
    for (int b = 0; b < 2; ++b) {
        if (h->stage_j[b]) { cudaFreeHost(h->stage_j[b]); h->stage_j[b] = nullptr; }
        if (h->stage_h[b]) { cudaFreeHost(h->stage_h[b]); h->stage_h[b] = nullptr; }
    }
    for (int b = 0; b < 2; ++b) {
        if (h->part_j[b]) { cudaFree(h->part_j[b]); h->part_j[b] = nullptr; }
        if (h->part_h[b]) { cudaFree(h->part_h[b]); h->part_h[b] = nullptr; }
    }
    h->part_units = 0;
    h->stage_units = 0;
    h->s_cap = -1;
}

// buffers of the compact host transfer (second arena set, pinned staging, run-length tables)
// Units per chunk of a host-complete call.  Those are bound by the host cores' expansion, not by the
// GPU: small chunks let the first arenas reach the host early, so the expansion pipeline starts after a
// millisecond instead of a third of the evaluation, and its last chunk is short.
int compact_step(const qcb_handle_s *h, bool with_jacobian = true) {
    if (getenv("QCB_E2E_BIG_CHUNKS")) return h->chunk_units;
    if (const char *e = getenv("QCB_E2E_UNITS")) return (int)std::min<long long>(h->chunk_units, std::max(1L, atol(e)));
    const long long slots = (long long)h->n_sm * (h->iso ? 2 : 1);
    const long long tiles = h->iso ? (h->dp / 128) : ((h->dp % 128 == 0) ? (long long)(h->dp / 128) * (h->dp / 128)
                                                                          : (long long)(h->dp / 64) * (h->dp / 64));
    // one and a half waves of per-drive CTAs.  Measured on cfg5 (e2e, knot-evals/s) with the background expansion:
    // 16 units 6.4 k, 24: 11.9 k (probably a slower box), 32: 16.2 k, 40: 16.2 k, 48: 15.6 k, 80: 14.7 k,
    // 148 (six waves, the round-1 choice): 14.5 k -- short chunks shorten the pipeline's fill and drain
    // A Hessian-only callback has a tenth of the host-side work (1.7 GB of values, most of them DMA'd straight into
    // place): it is bound by the GPU, so its chunks are six waves long (eval_h through host buffers with new_x = 0:
    // 62 ms with the short chunks, 48 ms with these).
    const long long waves2 = with_jacobian ? 3 : 12;      // chunk length in half waves
    const long long small = std::max(1LL, waves2 * slots / std::max<long long>(1, 2LL * std::max(1, h->m) * std::max(1LL, tiles)));
    return (int)std::min<long long>(h->chunk_units, small);
}

void ensure_compact(qcb_handle_s *h) {
    const int step = std::max(compact_step(h, true), compact_step(h, false));     // staging sized for the longer chunks
    if (h->stage_units != step) {
        const size_t n = (size_t)step;
        CK(cudaMalloc(&h->sc_ja2, n * (size_t)h->JA * 8));
        CK(cudaMalloc(&h->sc_ha2, n * (size_t)h->HA * 8));
        for (int b = 0; b < 2; ++b) {
            CK(cudaMallocHost(&h->stage_j[b], n * (size_t)h->JA * 8));
            CK(cudaMallocHost(&h->stage_h[b], n * (size_t)h->HA * 8));
            if (!h->copied[b]) CK(cudaEventCreateWithFlags(&h->copied[b], cudaEventDisableTiming));
            if (!h->part_done[b]) CK(cudaEventCreateWithFlags(&h->part_done[b], cudaEventDisableTiming));
        }
        if (const char *e = getenv("QCB_DIRECT_FRACTION")) h->direct_fraction = std::min(1.0, std::max(0.0, atof(e)));
        h->part_units = (int)(h->direct_fraction * step);
        for (int b = 0; b < 2 && h->part_units > 0; ++b) {
            CK(cudaMalloc(&h->part_j[b], (size_t)h->part_units * h->jnnz_unit * 8));
            CK(cudaMalloc(&h->part_h[b], (size_t)h->part_units * h->hnnz_unit * 8));
        }
        h->stage_units = step;
    }
    if (h->jsegs.empty()) {
        h->jsegs.resize(h->S);
        for (int s = 0; s < h->S; ++s) h->jsegs[s] = build_segments(h->h_jcode.data() + (size_t)s * h->jnnz_unit, h->jnnz_unit);
    }
    if (h->hsegs.empty()) {
        h->hsegs.resize(h->S);
        for (int s = 0; s < h->S; ++s) h->hsegs[s] = build_segments(h->h_hcode.data() + (size_t)s * h->hnnz_unit, h->hnnz_unit);
    }
}

// Split a unit's run-length table into direct runs and the rest (see qcb_handle_s::Direct).
void split_direct(const std::vector<Seg> &segs, int64_t arena_len, std::vector<qcb_handle_s::Direct> &direct,
                  std::vector<Seg> &rest, std::vector<std::pair<uint32_t, uint32_t>> &stage) {
    constexpr uint32_t MIN_DIRECT = 4096;       // 32 KB per unit and run: worth a DMA descriptor row
    std::vector<uint8_t> uses((size_t)arena_len, 0);
    for (const Seg &q : segs) {
        const uint32_t n = (q.kind & 2u) ? 1u : q.len;
        for (uint32_t i = 0; i < n; ++i) { uint8_t &c = uses[q.src + i]; if (c < 255) ++c; }
    }
    direct.clear(); rest.clear(); stage.clear();
    std::vector<uint8_t> needed((size_t)arena_len, 0);
    for (const Seg &q : segs) {
        bool ok = q.kind == 0 && q.len >= MIN_DIRECT;
        for (uint32_t i = 0; ok && i < q.len; ++i) ok = uses[q.src + i] == 1;
        if (ok) direct.push_back({q.dst, q.src, q.len});
        else {
            rest.push_back(q);
            const uint32_t n = (q.kind & 2u) ? 1u : q.len;
            for (uint32_t i = 0; i < n; ++i) needed[q.src + i] = 1;
        }
    }
    for (int64_t e = 0; e < arena_len;) {
        if (!needed[e]) { ++e; continue; }
        int64_t b = e;
        while (e < arena_len && needed[e]) ++e;
        stage.push_back({(uint32_t)b, (uint32_t)e});
    }
}

void ensure_direct(qcb_handle_s *h) {
    if (h->direct_built) return;
    h->direct_built = true;
    h->jdirect.clear(); h->hdirect.clear();
    if (h->S != 1 || getenv("QCB_E2E_NO_DIRECT_PANELS")) return;
    h->jsegs_x.assign(1, {}); h->hsegs_x.assign(1, {});
    split_direct(h->jsegs[0], h->JA, h->jdirect, h->jsegs_x[0], h->jstage);
    split_direct(h->hsegs[0], h->HA, h->hdirect, h->hsegs_x[0], h->hstage);
    // a handful of long runs, or nothing: many short ones would turn into as many DMA calls per chunk
    if (h->jdirect.size() > 8 || h->jstage.size() > 8) h->jdirect.clear();
    if (h->hdirect.size() > 8 || h->hstage.size() > 8) h->hdirect.clear();
}

bool host_pinned(const void *p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeHost;
}

size_t scratch_bytes_per_unit(const qcb_handle_s *h, int s_cap) {
    const size_t dd = (size_t)h->msz, pan = (size_t)h->dp * h->cp;
    size_t n = (QR_NUM_FIXED + 1 + 2 * (size_t)(s_cap + 1) + (size_t)(QCB_NUM_DSLOTS + s_cap + 1) * h->m) * dd;
    n += 3 * pan + h->JA + h->HA + (size_t)h->m * h->m + 2 * h->m + 3;
    return n * sizeof(double);
}

void ensure_scratch(qcb_handle_s *h, int s_need) {
    if (h->s_cap >= s_need && h->sc_fixed) return;
    free_scratch(h);
    const int s_cap = std::max(s_need, 2);
    // Scratch budget: a quarter of the free HBM, at most 24 GB (a cfg5 slab of 2000 knots then runs as
    // one chunk per squaring class: fewer, longer launches -- 37.9 -> 36.2 ms against the former 8 GB).
    size_t budget_mb = 8192;
    {
        size_t free_b = 0, total_b = 0;
        if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess)
            budget_mb = std::min<size_t>(40960, std::max<size_t>(512, (free_b >> 20) / 4));
        else
            cudaGetLastError();
    }
    if (const char *e = getenv("QCB_SCRATCH_MB")) budget_mb = (size_t)std::max(64L, atol(e));
    const size_t per = scratch_bytes_per_unit(h, s_cap);
    long long cu = (long long)((budget_mb << 20) / per);
    cu = std::max(1LL, std::min<long long>(cu, h->units));
    if (cu < h->units && h->m > 0) {
        // choose the chunk size (<= the budget) that wastes the fewest CTA waves over the whole slab:
        // the per-direction launches dominate, each has chunk*m CTAs on n_sm*ctas_per_sm slots
        const long long slots = (long long)h->n_sm * (h->iso ? 2 : 1);
        const long long tiles = h->iso ? (h->dp / 128) : ((h->dp % 128 == 0) ? (long long)(h->dp / 128) * (h->dp / 128)
                                                                              : (long long)(h->dp / 64) * (h->dp / 64));
        long long best = cu, best_waves = -1;
        for (long long t = cu; t >= std::max(1LL, cu * 2 / 3); --t) {
            // ~13 per-drive launches and ~20 per-unit launches per chunk, all of similar tile cost
            long long waves = 0;
            for (long long done = 0; done < h->units; done += t) {
                const long long n_ = std::min<long long>(t, h->units - done);
                waves += 13 * ((n_ * h->m * tiles + slots - 1) / slots) + 20 * ((n_ * tiles + slots - 1) / slots);
            }
            if (best_waves < 0 || waves < best_waves) { best_waves = waves; best = t; }
        }
        cu = best;
    }
    h->chunk_units = (int)cu;
    const size_t dd = (size_t)h->msz, pan = (size_t)h->dp * h->cp, n = (size_t)cu;
    const size_t mm = std::max(1, h->m);
    CK(cudaMalloc(&h->sc_fixed, (QR_NUM_FIXED + 1) * n * dd * 8));     // + one register for G E
    CK(cudaMalloc(&h->sc_E, (size_t)(s_cap + 1) * n * dd * 8));
    CK(cudaMalloc(&h->sc_W, (size_t)(s_cap + 1) * n * dd * 8));
    CK(cudaMalloc(&h->sc_dslot, (size_t)QCB_NUM_DSLOTS * n * mm * dd * 8));
    CK(cudaMalloc(&h->sc_L, (size_t)(s_cap + 1) * n * mm * dd * 8));
    CK(cudaMalloc(&h->sc_XP, n * pan * 8));
    CK(cudaMalloc(&h->sc_MP, n * pan * 8));
    CK(cudaMalloc(&h->sc_EX, n * pan * 8));
    CK(cudaMalloc(&h->sc_ja, n * (size_t)h->JA * 8));
    CK(cudaMalloc(&h->sc_ha, n * (size_t)h->HA * 8));
    CK(cudaMalloc(&h->sc_H, n * mm * mm * 8));
    CK(cudaMalloc(&h->sc_c1, n * mm * 8));
    CK(cudaMalloc(&h->sc_c2, n * mm * 8));
    CK(cudaMalloc(&h->sc_c3, n * 8));
    CK(cudaMalloc(&h->sc_tau, n * 8));
    CK(cudaMalloc(&h->sc_ntau, n * 8));
    h->s_cap = s_cap;
}

struct HostOut { double *delta, *jac, *hess; bool compact; };

// dst[unit] = src[unit]: one arena (stride doubles) per unit, 16-byte accesses; dst may be peer memory.
constexpr long long SHIP_PIECE = 8192;
__global__ void __launch_bounds__(256)
arena_ship_kernel(const double *src, long long stride, double *dst, int u0, const int *perm, long long lo, long long hi) {
    const int unit = blockIdx.y;
    const long long ug = perm ? perm[u0 + unit] : u0 + unit;
    const long long b = lo + (long long)blockIdx.x * SHIP_PIECE, len = min(SHIP_PIECE, hi - b);     // entries [lo, hi) of the arena
    const double *s = src + (long long)unit * stride + b;
    double *o = dst + ug * stride + b;
    if (((reinterpret_cast<uintptr_t>(o) | reinterpret_cast<uintptr_t>(s)) & 15) == 0) {
        const double2 *s2 = reinterpret_cast<const double2 *>(s);
        double2 *o2 = reinterpret_cast<double2 *>(o);
        const long long n2 = len >> 1;
        for (long long e = threadIdx.x; e < n2; e += blockDim.x) o2[e] = __ldg(s2 + e);
        if ((len & 1) && threadIdx.x == 0) o[len - 1] = s[len - 1];
    } else {
        for (long long e = threadIdx.x; e < len; e += blockDim.x) o[e] = s[e];
    }
}
void ensure_out(qcb_handle_s *h, double **p, size_t n);

// norm pre-pass over the whole slab -> uniform number of squarings (or -1)
int slab_squarings(qcb_handle_s *h, const double *dZ) {
    CK(cudaMemsetAsync(h->d_norm, 0, sizeof(unsigned long long), h->stream));
    CK(cudaMemsetAsync(h->d_bad, 0, 2 * sizeof(int), h->stream));
    if (!h->d_unit_norm) {
        CK(cudaMalloc(&h->d_unit_norm, (size_t)h->units * 8));
        h->h_unit_norm.resize(h->units);
    }
    build_G_kernel<<<h->units, 256, 0, h->stream>>>(h->geom(), dZ, 0, nullptr, 0, h->d_norm, h->d_bad, h->d_unit_norm);
    ++h->launches;
    CK(cudaGetLastError());
    unsigned long long bits = 0;
    int bad = 0;
    // per-unit norms: every chunk of units takes the squaring count its own members need
    CK(cudaMemcpyAsync(h->h_unit_norm.data(), h->d_unit_norm, (size_t)h->units * 8, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaMemcpyAsync(&bits, h->d_norm, sizeof(bits), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaMemcpyAsync(&bad, h->d_bad, sizeof(bad), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    if (bad) return -1;
    double nrm;
    memcpy(&nrm, &bits, sizeof(nrm));
    return qcb_squarings(nrm);
}

void eval_large(qcb_handle_s *h, const double *dZ, const double *dmu, double *d_delta, double *d_jac,
                double *d_hess, int want, const HostOut *host, int s_slab, double *E_out = nullptr) {
    ensure_scratch(h, s_slab);
    const bool wj = want & 1, wh = (want & 2) && dmu;
    const long long dd = h->msz, pan = (long long)h->dp * h->cp;
    const int m = h->m;
    const int ncol = h->gcols;                       // columns of a stored matrix (dp, or dp/2 in iso mode)
    UnitGeom q = h->geom();
    const bool compact = host && host->compact && (wj || wh);
    if (compact) { ensure_compact(h); ensure_direct(h); }
    // direct panels need a pinned destination (the DMA engine writes it); pageable buffers keep the staged route
    const bool jdir = compact && wj && host->jac && !h->jdirect.empty() && h->direct_fraction == 0.0 && host_pinned(host->jac);
    const bool hdir = compact && wh && host->hess && !h->hdirect.empty() && h->direct_fraction == 0.0 && host_pinned(host->hess);
    // chunk whose arenas are staged on the host, not yet expanded; its first prev_nd units went
    // the direct route (assembled on the device, full COO rows DMA'd) and need no expansion
    int prev_u0 = -1, prev_nu = 0, prev_nd = 0;
    // The expansion of a staged chunk runs on the pool's workers in the background while this thread queues
    // the next chunk's kernels; it is joined (the caller helps) before the staging buffer it reads is reused.
    ExpandTask task;
    auto expand_prev = [&](size_t idx_prev) {
        if (prev_u0 < 0) return;
        const int b = (int)(idx_prev & 1);
        CK(cudaEventSynchronize(h->copied[b]));
        ExpandJob jobs[2];
        int nj = 0;
        if (wj && host->jac) jobs[nj++] = {jdir ? &h->jsegs_x : &h->jsegs, h->stage_j[b] + (size_t)prev_nd * h->JA, h->JA, host->jac, h->jnnz_unit};
        if (wh && host->hess) jobs[nj++] = {hdir ? &h->hsegs_x : &h->hsegs, h->stage_h[b] + (size_t)prev_nd * h->HA, h->HA, host->hess, h->hnnz_unit};
        if (nj > 0 && prev_nu > prev_nd) task.start(jobs, nj, h->nk, prev_u0 + prev_nd, prev_nu - prev_nd);
        prev_u0 = -1;
    };
    // delivery_split > 1 (multi-GPU delivery): the slab is cut into that many sequential pieces even though the
    // scratch holds all of it, so that the first piece's arenas cross NVLink while the next is computed
    const int chunk_step = compact ? compact_step(h, wj)
                                   : ((h->delivery_split > 1 && (h->arena_out_j || h->arena_out_h) && h->chunk_units >= h->units)
                                          ? std::max(1, (h->units + h->delivery_split - 1) / h->delivery_split)
                                          : h->chunk_units);
    // Device-resident calls on one system: visit the units in the order of their squaring count, so
    // that a chunk never runs levels only a few of its members need (random pulses: most units sit
    // one level below the largest).  perm maps position -> unit; the kernels that touch Z, mu and
    // the outputs read it, everything in between works on chunk-local scratch.  Host-complete calls
    // are bound by the host cores and keep the natural order (contiguous D2H ranges).
    const int *perm = nullptr;
    h->h_class_end.clear();
    // switches are read per call (tests flip them between calls), never inside the chunk loop
    const bool per_unit_s = (int)h->h_unit_norm.size() == h->units && !getenv("QCB_UNIFORM_SQUARINGS");
    if (!host && !E_out && h->S == 1 && per_unit_s && !getenv("QCB_NO_UNIT_SORT")) {
        std::vector<int> cls(h->units);
        int lo = QCB_MAX_SQUARINGS, hi = 0;
        for (int u = 0; u < h->units; ++u) {
            cls[u] = std::min(s_slab, std::max(0, qcb_squarings(h->h_unit_norm[u])));
            lo = std::min(lo, cls[u]); hi = std::max(hi, cls[u]);
        }
        if (hi > lo) {
            h->h_perm.resize(h->units);
            int p = 0;
            for (int c = lo; c <= hi; ++c) {
                for (int u = 0; u < h->units; ++u) if (cls[u] == c) h->h_perm[p++] = u;
                if (h->h_class_end.empty() || h->h_class_end.back() != p) h->h_class_end.push_back(p);
            }
            if (!h->d_perm) CK(cudaMalloc(&h->d_perm, (size_t)h->units * sizeof(int)));
            CK(cudaMemcpyAsync(h->d_perm, h->h_perm.data(), (size_t)h->units * sizeof(int), cudaMemcpyHostToDevice, h->stream));
            perm = h->d_perm;
        }
    }
    // Lanes: when the scratch holds the whole slab, every squaring class is one chunk on its own
    // slice of the scratch arrays (offset = its first position), and the classes alternate between two
    // streams.  A small class (a few dozen units: launches of one or two CTA waves) then fills the
    // SM slots the big class leaves idle in its wave tails instead of serialising ~50 under-filled
    // launches behind it -- what strong scaling over 8 GPUs (250 units per GPU) otherwise pays for.
    // When the scratch holds the whole slab every chunk works on its own slice (offset = its first
    // position): nothing is overwritten by a later chunk, which is what lets a new_x = 0 call find the
    // previous call's intermediates in place.
    // (host-buffer calls included: their small chunks then work on consecutive slices, which is what lets
    // qcb_eval_jacobian_x / qcb_eval_hessian_x with new_x = 0 reuse what an earlier callback computed)
    const bool slices = !E_out && h->chunk_units >= h->units && (long long)h->m * dd >= (long long)h->dp * h->dp &&
                        !(host && h->direct_fraction > 0.0);
    const bool lanes = perm && slices && !h->profiling && !getenv("QCB_NO_LANES");   // per-launch timing needs kernels that run alone
    const int reuse = slices ? (h->reuse & h->cache_have) : 0;
    h->cache_have = 0;              // re-established below once this call's launches are queued
    if (lanes) {
        if (!h->stream2) CK(cudaStreamCreateWithFlags(&h->stream2, cudaStreamNonBlocking));
        if (!h->lane_ev[0]) for (int i = 0; i < 2; ++i) CK(cudaEventCreateWithFlags(&h->lane_ev[i], cudaEventDisableTiming));
        CK(cudaEventRecord(h->lane_ev[0], h->stream));          // inputs, norms and perm are ready
        CK(cudaStreamWaitEvent(h->stream2, h->lane_ev[0], 0));
    }
    size_t chunk_idx = 0, class_idx = 0;
    int nu_next = 0;
    for (int u0 = 0; u0 < h->units; u0 += nu_next, ++chunk_idx) {
        int nu = std::min(chunk_step, h->units - u0);
        if (perm) {     // chunks do not straddle a class boundary
            while (class_idx < h->h_class_end.size() && h->h_class_end[class_idx] <= u0) ++class_idx;
            if (class_idx < h->h_class_end.size()) nu = std::min(nu, h->h_class_end[class_idx] - u0);
        }
        nu_next = nu;
        const long long ub = slices ? u0 : 0;                          // scratch slice of this chunk
        // lanes alternate by squaring class: the pieces of one class follow each other on their stream
        cudaStream_t st = (lanes && (class_idx & 1)) ? h->stream2 : h->stream;
        // squarings of this chunk: what the largest ||dt G(u)||_1 among its units needs (smooth pulses
        // have slowly varying norms, so quiet stretches of the trajectory skip levels the peaks need)
        int s = s_slab;
        if (per_unit_s) {
            double mx = 0.0;
            for (int u = u0; u < u0 + nu; ++u) mx = std::max(mx, h->h_unit_norm[perm ? h->h_perm[u] : u]);
            s = std::min(s_slab, std::max(0, qcb_squarings(mx)));
        }
        LargeExec X{h, nu, u0, s};
        X.dd = dd; X.pan = pan; X.ubase = (int)ub; X.st = st;
        const int buf = compact ? (int)(chunk_idx & 1) : 0;
        // sliced scratch: every chunk has arena rows of its own (nothing to double-buffer)
        X.ja = (buf && !slices) ? h->sc_ja2 : h->sc_ja;
        X.ha = (buf && !slices) ? h->sc_ha2 : h->sc_ha;
        // the D2H of chunk idx-2 read this arena set: let it finish before overwriting
        if (compact && chunk_idx >= 2) {
            CK(cudaStreamWaitEvent(st, h->copied[buf], 0));
            CK(cudaStreamWaitEvent(st, h->part_done[buf], 0));
        }
        // hybrid: the PCIe link and the host cores both write COO rows -- the first nd units of the
        // chunk are assembled on the device and DMA'd in full, the rest travel as compact arenas
        const int nd = compact ? std::min(h->part_units, (int)(h->direct_fraction * nu)) : 0;
        Opnd G = X.fixed(QR_G), A = X.fixed(QR_A);
        if (!(reuse & 1)) {
            build_G_kernel<<<nu, 256, 0, st>>>(q, dZ, u0, const_cast<double *>(G.p), dd, h->d_norm, h->d_bad, nullptr, perm);
            ++h->launches;
        }
        // panels of X and M, tau, A = tau G (cheap; M changes with the multipliers even when Z does not)
        prep_unit_kernel<<<nu, 256, 0, st>>>(q, dZ, wh ? dmu : nullptr, u0, s, G.p, const_cast<double *>(A.p), dd,
                                                     X.at(h->sc_XP, pan), wh ? X.at(h->sc_MP, pan) : nullptr, pan, X.at(h->sc_tau, 1),
                                                     X.at(h->sc_ntau, 1), perm);
        ++h->launches;
        CK(cudaGetLastError());
        Opnd XP = X.mat(h->sc_XP, pan), MP = X.mat(h->sc_MP, pan), EXp = X.mat(h->sc_EX, pan);
        Opnd Ef = X.E(s);
        if (!(reuse & 1)) {
            // polynomial + squarings
            X.level = 0;
            for (int i = 0; i < QCB_NOPS(QCB_P1); ++i) X.run(QCB_P1[i], 1);
            for (int l = 0; l < s; ++l) { X.level = l; X.run(QCB_P2[0], 1); }
            // defect
            X.product(EXp, Ef, 0, XP, 0, h->dp, h->cp, h->dp, 1, 1.0, nullptr);
        }
        delta_kernel<<<nu, 256, 0, st>>>(q, dZ, X.at(h->sc_EX, pan), pan, d_delta, u0, perm);
        ++h->launches;
        const long long e_part = (long long)h->dp * h->dp;        // the arena starts with the full E
        if (h->arena_out_j && (h->arena_parts & 1) && !(wj && h->arena_parts == 3)) {
            // propagators first: the x-block of the Jacobian (c copies of -E, 90 % of its entries on the unitary
            // path) depends on nothing else, so the Ipopt rank can assemble it while the tangents are computed
            if (h->iso) { iso_expand_kernel<<<nu, 256, 0, st>>>(X.at(X.ja, h->JA), h->JA, h->dp); ++h->launches; }
            if (h->chunk_events.size() <= chunk_idx) {
                cudaEvent_t e;
                CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
                h->chunk_events.push_back(e);
            }
            CK(cudaEventRecord(h->chunk_events[chunk_idx], st));
            CK(cudaStreamWaitEvent(h->copy_stream, h->chunk_events[chunk_idx], 0));
            arena_ship_kernel<<<dim3((unsigned)((e_part + SHIP_PIECE - 1) / SHIP_PIECE), (unsigned)nu), 256, 0, h->copy_stream>>>(
                X.at(X.ja, h->JA), h->JA, h->arena_out_j, u0, perm, 0, e_part);
            ++h->launches;
        }
        if (E_out) {      // rollouts: keep the (unpadded, full) propagator of every unit
            if (h->iso) { iso_expand_kernel<<<nu, 256, 0, st>>>(X.at(X.ja, h->JA), h->JA, h->dp); ++h->launches; }
            extract_E_kernel<<<nu, 256, 0, st>>>(X.at(X.ja, h->JA), h->JA, h->dp, h->d, E_out + (size_t)u0 * h->d * h->d);
            ++h->launches;
        }
        if (wj || wh) {
            const bool tangents = !(reuse & 2);      // forward tangents, L_j levels, G E and the Jacobian arena
            // Hessian side of an iterate whose tangents are already there (new_x = 0): the (x,u) and (x,dt)
            // panels -- all but a few hundred numbers of the Hessian arena -- need only L_j, G E and the
            // multipliers, so they are formed first and, in a multi-GPU delivery, leave while the second-order
            // sweeps run
            const bool hx_early = wh && !tangents;
            const long long h_panels = (long long)(m + 1) * pan;
            auto hx_panels = [&]() {
                X.level = s;
                Opnd GE_ = X.fixed(QR_NUM_FIXED), L_ = X.reg(QR_L);
                Opnd HXU = X.mat(X.ha, h->HA, pan), HXT = X.mat(X.ha + m * pan, h->HA);
                if (m > 0) X.product(HXU, L_, 1, MP, 0, h->dp, h->cp, h->dp, m, 1.0, X.at(h->sc_ntau, 1));
                X.product(HXT, GE_, 1, MP, 0, h->dp, h->cp, h->dp, 1, -1.0, nullptr);
            };
            auto ship_h = [&](long long lo, long long hi) {
                if (h->chunk_events.size() <= chunk_idx) {
                    cudaEvent_t e;
                    CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
                    h->chunk_events.push_back(e);
                }
                CK(cudaEventRecord(h->chunk_events[chunk_idx], st));
                CK(cudaStreamWaitEvent(h->copy_stream, h->chunk_events[chunk_idx], 0));
                arena_ship_kernel<<<dim3((unsigned)((hi - lo + SHIP_PIECE - 1) / SHIP_PIECE), (unsigned)nu), 256, 0, h->copy_stream>>>(
                    X.at(X.ha, h->HA), h->HA, h->arena_out_h, u0, perm, lo, hi);
                ++h->launches;
            };
            if (hx_early) {
                hx_panels();
                if (h->arena_out_h) ship_h(0, h_panels);
            }
            if (wh) {
                if (h->iso) {   // W = M X^T is a general matrix: project it onto the iso algebra (half storage)
                    // W (dp x dp per unit) borrows the head of this slice of the per-drive temporary U (dead
                    // until the first second-order product; the persistent per-drive registers stay intact)
                    // (with a single drive in half storage U is smaller than W: the head of the whole per-drive
                    // region is used then, and nothing is kept across calls -- see `slices`)
                    double *wfull = ((long long)h->m * dd >= (long long)h->dp * h->dp)
                                        ? h->sc_dslot + ((long long)qcb_dslot(QR_U) * h->chunk_units + ub) * h->m * dd
                                        : h->sc_dslot + ub * (long long)h->m * dd;
                    Opnd Wfull = X.mat(wfull, (long long)h->dp * h->dp);
                    Wfull.p = wfull;
                    X.product(Wfull, MP, 0, XP, 1, h->dp, h->dp, h->cp, 1, 1.0, nullptr, true);
                    iso_project_kernel<<<nu, 256, 0, st>>>(wfull, (long long)h->dp * h->dp,
                                                                  const_cast<double *>(X.W(s).p), dd, h->dp);
                    ++h->launches;
                } else {
                    X.product(X.W(s), MP, 0, XP, 1, h->dp, h->dp, h->cp, 1, 1.0, nullptr);   // W = M X^T
                }
                for (int l = s - 1; l >= 0; --l) { X.level = l; X.run(QCB_P3[0], 1); }
                X.level = 0;
                for (int i = 0; i < QCB_NOPS(QCB_P4); ++i) X.run(QCB_P4[i], 1);
            }
            X.level = 0;
            if (m > 0) {
                if (tangents) for (int i = 0; i < QCB_NOPS(QCB_P5); ++i) X.run(QCB_P5[i], m);
                if (wh) {   // second order: one product per node and direction, then m x m contractions
                    for (int i = 0; i < QCB_NOPS(QCB_P6); ++i) {
                        X.run(QCB_P6[i], m);
                        X.contract(X.reg(QR_U), m, X.reg(QCB_P6_PARTNER[i]), m, X.at(h->sc_H, (long long)std::max(1, m) * std::max(1, m)), i > 0);
                    }
                }
                for (int l = 0; l < s; ++l) {
                    X.level = l;
                    if (wh) {
                        X.run(QCB_P7U[0], m);
                        X.contract(X.reg(QR_U), m, X.reg(QR_L), m, X.at(h->sc_H, (long long)std::max(1, m) * std::max(1, m)), 1);
                    }
                    if (tangents) X.run(QCB_P7L[0], m);
                }
            }
            // final blocks.  G E has a register of its own (the powers A2..A4 must survive for a later
            // Hessian call on the same Z); W E^T and G^T W reuse reverse-sweep registers that are dead by now
            X.level = s;
            Opnd GE = X.fixed(QR_NUM_FIXED), WE = X.fixed(QR_ZB1), GW = X.fixed(QR_ZB2), L = X.reg(QR_L);
            const long long esz = (long long)h->dp * h->dp;   // the arena always holds the full E
            if (tangents) {
                X.product(GE, G, 0, Ef, 0, h->dp, ncol, h->dp, 1, 1.0, nullptr);
                Opnd JU = X.mat(X.ja + esz, h->JA, pan), JT = X.mat(X.ja + esz + m * pan, h->JA);
                if (m > 0) X.product(JU, L, 0, XP, 0, h->dp, h->cp, h->dp, m, 1.0, X.at(h->sc_ntau, 1));
                X.product(JT, GE, 0, XP, 0, h->dp, h->cp, h->dp, 1, -1.0, nullptr);
                arena_consts_kernel<<<(nu + 127) / 128, 128, 0, st>>>(X.at(X.ja, h->JA), h->JA, esz + (m + 1) * pan, nu);
                ++h->launches;
                if (h->iso) {   // right block column of E from the computed left one: the COO x-block is then c
                                // contiguous copies of -E (long runs for the device gather and the host expansion)
                    iso_expand_kernel<<<nu, 256, 0, st>>>(X.at(X.ja, h->JA), h->JA, h->dp);
                    ++h->launches;
                }
            }
            if (wh) {
                Opnd Wt = X.W(s);
                if (!hx_early) hx_panels();
                X.product(WE, Wt, 0, Ef, 1, h->dp, ncol, h->dp, 1, 1.0, nullptr);
                X.product(GW, G, 1, Wt, 0, h->dp, ncol, h->dp, 1, 1.0, nullptr);
                if (m > 0) {
                    X.contract(WE, 1, X.reg(QR_GJ), m, X.at(h->sc_c1, std::max(1, m)), 0);
                    X.contract(GW, 1, L, m, X.at(h->sc_c2, std::max(1, m)), 0);
                }
                X.contract(GW, 1, GE, 1, X.at(h->sc_c3, 1), 0);
                finalize_scalars_kernel<<<nu, 64, 0, st>>>(m, X.at(h->sc_tau, 1), X.at(h->sc_H, (long long)std::max(1, m) * std::max(1, m)),
                                                          X.at(h->sc_c1, std::max(1, m)), X.at(h->sc_c2, std::max(1, m)), X.at(h->sc_c3, 1),
                                                          X.at(X.ha, h->HA), h->HA, (m + 1) * pan, h->iso ? 2.0 : 1.0);
                ++h->launches;
            }
            if ((wj && h->arena_out_j) || (wh && h->arena_out_h)) {
                // multi-GPU delivery of the compact form: every distinct number of the unit once, stored
                // straight into the Ipopt rank's HBM (NVLink peer stores); that rank assembles the COO vectors.
                // The copy runs on the side stream behind this chunk's kernels, so it overlaps whatever the
                // compute streams do next (the next chunk, or the Hessian side of the same iterate).
                if (h->chunk_events.size() <= chunk_idx) {
                    cudaEvent_t e;
                    CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
                    h->chunk_events.push_back(e);
                }
                CK(cudaEventRecord(h->chunk_events[chunk_idx], st));
                CK(cudaStreamWaitEvent(h->copy_stream, h->chunk_events[chunk_idx], 0));
                if (wj && h->arena_out_j && (h->arena_parts & 2)) {
                    const long long lo = h->arena_parts == 3 ? 0 : e_part;       // the E part may have gone ahead
                    arena_ship_kernel<<<dim3((unsigned)((h->JA - lo + SHIP_PIECE - 1) / SHIP_PIECE), (unsigned)nu), 256, 0, h->copy_stream>>>(
                        X.at(X.ja, h->JA), h->JA, h->arena_out_j, u0, perm, lo, h->JA);
                    ++h->launches;
                }
                if (wh && h->arena_out_h) {     // the panels may have gone ahead: then only the scalars are left
                    const long long lo = hx_early ? h_panels : 0;
                    arena_ship_kernel<<<dim3((unsigned)((h->HA - lo + SHIP_PIECE - 1) / SHIP_PIECE), (unsigned)nu), 256, 0, h->copy_stream>>>(
                        X.at(X.ha, h->HA), h->HA, h->arena_out_h, u0, perm, lo, h->HA);
                    ++h->launches;
                }
            }
            if (wj && d_jac && !compact) {
                if (h->run_assembly) {
                    if (!h->d_jpieces) upload_pieces(h, h->h_jcode, h->jnnz_unit, &h->d_jpieces, &h->n_jpieces);
                    assemble_runs_kernel<<<(unsigned)((size_t)nu * h->n_jpieces), 256, 0, st>>>(
                        X.at(X.ja, h->JA), h->JA, h->d_jcode, h->jnnz_unit, h->d_jpieces, h->n_jpieces, d_jac, u0, h->nk, u0, perm);
                } else {
                    int cpu = (int)std::min<int64_t>(64, std::max<int64_t>(1, h->jnnz_unit / 4096));
                    assemble_kernel<<<nu * cpu, 256, 0, st>>>(X.at(X.ja, h->JA), h->JA, h->d_jcode, h->jnnz_unit, d_jac, u0, h->nk, cpu, u0, perm);
                }
                ++h->launches;
            }
            if (wh && d_hess && !compact) {
                if (h->run_assembly) {
                    if (!h->d_hpieces) upload_pieces(h, h->h_hcode, h->hnnz_unit, &h->d_hpieces, &h->n_hpieces);
                    assemble_runs_kernel<<<(unsigned)((size_t)nu * h->n_hpieces), 256, 0, st>>>(
                        X.at(X.ha, h->HA), h->HA, h->d_hcode, h->hnnz_unit, h->d_hpieces, h->n_hpieces, d_hess, u0, h->nk, u0, perm);
                } else {
                    int cpu = (int)std::min<int64_t>(64, std::max<int64_t>(1, h->hnnz_unit / 4096));
                    assemble_kernel<<<nu * cpu, 256, 0, st>>>(X.at(X.ha, h->HA), h->HA, h->d_hcode, h->hnnz_unit, d_hess, u0, h->nk, cpu, u0, perm);
                }
                ++h->launches;
            }
            if (compact && nd > 0) {
                if (wj && host->jac) {
                    int cpu = (int)std::min<int64_t>(64, std::max<int64_t>(1, h->jnnz_unit / 4096));
                    assemble_kernel<<<nd * cpu, 256, 0, st>>>(X.ja, h->JA, h->d_jcode, h->jnnz_unit, h->part_j[buf], u0, h->nk, cpu, 0);
                    ++h->launches;
                }
                if (wh && host->hess) {
                    int cpu = (int)std::min<int64_t>(64, std::max<int64_t>(1, h->hnnz_unit / 4096));
                    assemble_kernel<<<nd * cpu, 256, 0, st>>>(X.ha, h->HA, h->d_hcode, h->hnnz_unit, h->part_h[buf], u0, h->nk, cpu, 0);
                    ++h->launches;
                }
            }
        }
        CK(cudaGetLastError());
        if (host) {   // stream this chunk's results to the host while the next chunk computes
            if (h->chunk_events.size() <= chunk_idx) {
                cudaEvent_t e;
                CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
                h->chunk_events.push_back(e);
            }
            CK(cudaEventRecord(h->chunk_events[chunk_idx], h->stream));
            CK(cudaStreamWaitEvent(h->copy_stream, h->chunk_events[chunk_idx], 0));
            if (host->delta) {
                CK(cudaMemcpyAsync(host->delta + (size_t)u0 * h->xdim, d_delta + (size_t)u0 * h->xdim,
                                   (size_t)nu * h->xdim * 8, cudaMemcpyDeviceToHost, h->copy_stream));
                h->xfer[1] += (int64_t)((size_t)nu * h->xdim * 8);
            }
            if (compact) {
                // compact transfer: the value arenas (every distinct number once) cross PCIe; the
                // host cores write the COO vectors from them while the GPU works on the next chunk
                // one arena set -> host: in one piece, or (direct panels) the long plain runs straight to their
                // place in the caller's vector and only the ranges the cores still read to the staging buffer
                auto to_host = [&](bool direct, const double *arena, int64_t stride, double *stage, double *out, int64_t nnz,
                                   const std::vector<qcb_handle_s::Direct> &runs, const std::vector<std::pair<uint32_t, uint32_t>> &ranges) {
                    const size_t rows = (size_t)(nu - nd);
                    if (!direct) {
                        CK(cudaMemcpyAsync(stage + (size_t)nd * stride, arena + (size_t)nd * stride, rows * stride * 8,
                                           cudaMemcpyDeviceToHost, h->copy_stream));
                        h->xfer[1] += (int64_t)(rows * stride * 8);
                        return;
                    }
                    for (const auto &r : ranges) {
                        CK(cudaMemcpy2DAsync(stage + (size_t)nd * stride + r.first, (size_t)stride * 8, arena + (size_t)nd * stride + r.first,
                                             (size_t)stride * 8, (size_t)(r.second - r.first) * 8, rows, cudaMemcpyDeviceToHost, h->copy_stream));
                        h->xfer[1] += (int64_t)(rows * (r.second - r.first) * 8);
                    }
                    for (const auto &d : runs) {
                        CK(cudaMemcpy2DAsync(out + (size_t)(u0 + nd) * nnz + d.dst, (size_t)nnz * 8, arena + (size_t)nd * stride + d.src,
                                             (size_t)stride * 8, (size_t)d.len * 8, rows, cudaMemcpyDeviceToHost, h->copy_stream));
                        h->xfer[1] += (int64_t)(rows * d.len * 8);
                    }
                };
                task.finish();      // chunk idx - 2 read the staging buffers the copies below overwrite
                if (wj && host->jac && nu > nd) to_host(jdir, X.at(X.ja, h->JA), h->JA, h->stage_j[buf], host->jac, h->jnnz_unit, h->jdirect, h->jstage);
                if (wh && host->hess && nu > nd) to_host(hdir, X.at(X.ha, h->HA), h->HA, h->stage_h[buf], host->hess, h->hnnz_unit, h->hdirect, h->hstage);
                CK(cudaEventRecord(h->copied[buf], h->copy_stream));
                if (nd > 0) {
                    if (wj && host->jac) {
                        CK(cudaMemcpyAsync(host->jac + (size_t)u0 * h->jnnz_unit, h->part_j[buf],
                                           (size_t)nd * h->jnnz_unit * 8, cudaMemcpyDeviceToHost, h->copy_stream));
                        h->xfer[1] += (int64_t)((size_t)nd * h->jnnz_unit * 8);
                    }
                    if (wh && host->hess) {
                        CK(cudaMemcpyAsync(host->hess + (size_t)u0 * h->hnnz_unit, h->part_h[buf],
                                           (size_t)nd * h->hnnz_unit * 8, cudaMemcpyDeviceToHost, h->copy_stream));
                        h->xfer[1] += (int64_t)((size_t)nd * h->hnnz_unit * 8);
                    }
                }
                CK(cudaEventRecord(h->part_done[buf], h->copy_stream));
                expand_prev(chunk_idx - 1);
                prev_u0 = u0; prev_nu = nu; prev_nd = nd;
            } else {
                if (wj && host->jac) {
                    CK(cudaMemcpyAsync(host->jac + (size_t)u0 * h->jnnz_unit, d_jac + (size_t)u0 * h->jnnz_unit,
                                       (size_t)nu * h->jnnz_unit * 8, cudaMemcpyDeviceToHost, h->copy_stream));
                    h->xfer[1] += (int64_t)((size_t)nu * h->jnnz_unit * 8);
                }
                if (wh && host->hess) {
                    CK(cudaMemcpyAsync(host->hess + (size_t)u0 * h->hnnz_unit, d_hess + (size_t)u0 * h->hnnz_unit,
                                       (size_t)nu * h->hnnz_unit * 8, cudaMemcpyDeviceToHost, h->copy_stream));
                    h->xfer[1] += (int64_t)((size_t)nu * h->hnnz_unit * 8);
                }
            }
        }
    }
    if (compact) { task.finish(); expand_prev(chunk_idx - 1); task.finish(); }
    if (h->arena_out_j || h->arena_out_h) {
        for (int i = 0; i < 3; ++i) if (!h->ship_ev[i]) CK(cudaEventCreateWithFlags(&h->ship_ev[i], cudaEventDisableTiming));
        if (h->arena_out_j && (h->arena_parts & 1)) CK(cudaEventRecord(h->ship_ev[0], h->copy_stream));
        if (wj && h->arena_out_j && (h->arena_parts & 2)) CK(cudaEventRecord(h->ship_ev[1], h->copy_stream));
        if (wh && h->arena_out_h) CK(cudaEventRecord(h->ship_ev[2], h->copy_stream));
    }
    if (slices) {       // every slice is intact: a new_x = 0 call may pick up from here
        h->cache_have = 1 | (((wj || wh) || (reuse & 2)) ? 2 : 0);
        h->cache_s_slab = s_slab;
    }
    if (lanes) {        // join: everything the caller orders after this call waits for both lanes
        CK(cudaEventRecord(h->lane_ev[1], h->stream2));
        CK(cudaStreamWaitEvent(h->stream, h->lane_ev[1], 0));
    }
}

template <int D, int C = 0, int M = 0, int TPU = 0>
void launch_small(qcb_handle_s *h, const SmallParams &p, size_t smem) {
    auto kern = small_unit_kernel<D, C, M, TPU>;
    CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int ctas = (p.units + p.upc - 1) / p.upc;
    kern<<<ctas, p.upc * p.tpu, smem, h->stream>>>(p);
}

void eval_small(qcb_handle_s *h, const double *dZ, const double *dmu, double *d_delta, double *d_jac,
                double *d_hess, int want, int s, double *E_out = nullptr, bool arenas = false) {
    SmallParams p;
    p.E_out = E_out;
    p.ja_out = arenas ? h->d_ja_all : nullptr; p.ha_out = arenas ? h->d_ha_all : nullptr;
    p.ja_stride = h->JA; p.ha_stride = h->HA;
    p.d = h->d; p.c = h->c; p.m = h->m; p.z_dim = h->z_dim; p.u_off = h->u_off; p.dt_off = h->dt_off;
    p.K = h->K; p.nk = h->nk; p.k_begin = h->k0; p.S = h->S; p.units = h->units;
    p.x_off = h->d_xoff; p.G0 = h->dG0; p.Gd = h->dGd; p.Z = dZ; p.mu = dmu;
    p.delta = d_delta; p.jac = d_jac; p.hess = d_hess;
    p.jcode = h->d_jcode; p.hcode = h->d_hcode; p.jnnz = h->jnnz_unit; p.hnnz = h->hnnz_unit;
    p.want = want; p.smax = s; p.bad_flag = h->d_bad;
    // threads per unit: one per matrix element, as a power of two <= 32 or a multiple of 32 (<= 512)
    // run-time d: one thread per matrix element; compile-time d: one thread per register tile of RB rows
    const int dd = (h->d / small_tile_rows(h->d)) * h->d;
    int tpu = 1;
    if (dd <= 32) { while (tpu < dd) tpu <<= 1; } else tpu = std::min(512, round_up(dd, 32));
    const size_t unit_d = small_unit_doubles(h->d, h->c, h->m, s);
    if (small_smem_doubles(h->d, h->c, h->m, s, 1) * 8 > 227 * 1024) throw SmallCapacity{s};
    // units per CTA: the kernel is latency bound (dependent tiny products, sub-warp barriers), so what
    // counts is the number of units resident per SM = CTAs that fit (shared memory, threads) x upc.
    // Small problems keep upc low enough to put a CTA on every SM twice over.
    int upc = 1;
    long best_resident = 0;
    for (int cand = 1; cand <= 16; cand *= 2) {
        if (cand * tpu > (small_tiled(h->d) ? 256 : 512)) break;
        if (cand > 1 && (long long)cand * h->n_sm * 2 > h->units) break;
        const size_t bytes = small_smem_doubles(h->d, h->c, h->m, s, cand) * 8 + 1024;   // + per-CTA reservation
        if (bytes > 227 * 1024) break;
        const long ctas = std::min<long>(std::min<long>(32, (long)(228 * 1024 / bytes)), 2048 / (cand * tpu));
        const long resident = ctas * cand;
        if (resident >= best_resident) { best_resident = resident; upc = cand; }
    }
    p.tpu = tpu; p.upc = upc; p.unit_doubles = (int)unit_d;
    const size_t smem = small_smem_doubles(h->d, h->c, h->m, s, upc) * sizeof(double);
    // fully specialised instances of the BASELINE shapes (state columns, drives, threads per unit)
    const int dcm = h->d * 10000 + h->c * 100 + h->m;
    if (dcm == 40202 && tpu == 16 / QCB_TILE4) { launch_small<4, 2, 2, 16 / QCB_TILE4>(h, p, smem); ++h->launches; CK(cudaGetLastError()); return; }
    if (dcm == 60302 && tpu == 64) { launch_small<6, 3, 2, 64>(h, p, smem); ++h->launches; CK(cudaGetLastError()); return; }
    if (dcm == 180904 && tpu == 352) { launch_small<18, 9, 4, 352>(h, p, smem); ++h->launches; CK(cudaGetLastError()); return; }
    if (dcm == 180104 && tpu == 352) { launch_small<18, 1, 4, 352>(h, p, smem); ++h->launches; CK(cudaGetLastError()); return; }
    switch (h->d) {
    case 4: launch_small<4>(h, p, smem); break;
    case 6: launch_small<6>(h, p, smem); break;
    case 8: launch_small<8>(h, p, smem); break;
    case 10: launch_small<10>(h, p, smem); break;
    case 12: launch_small<12>(h, p, smem); break;
    case 16: launch_small<16>(h, p, smem); break;
    case 18: launch_small<18>(h, p, smem); break;
    default: launch_small<0>(h, p, smem); break;
    }
    ++h->launches;
    CK(cudaGetLastError());
}

// State index map: the caller's Z / mu (device copies) -> the packed layout the kernels read.
// One CTA per knot of the slab (plus the halo knot); mu rows are reordered from the caller's defect
// rows to panel order.
__global__ void pack_knots_kernel(const double *Zext, int ez, double *Zp, int zp, const int *xmap, int nx, int eu, int m,
                                  int et, int k_lo) {
    const long long k = k_lo + blockIdx.x;
    const double *src = Zext + k * ez;
    double *dst = Zp + k * zp;
    for (int i = threadIdx.x; i < nx; i += blockDim.x) dst[i] = src[xmap[i]];
    for (int j = threadIdx.x; j < m; j += blockDim.x) dst[nx + j] = src[eu + j];
    if (threadIdx.x == 0) dst[nx + m] = src[et];
}
__global__ void permute_rows_kernel(const double *src, double *dst, const int *rmap, int xdim, long long rows, int scatter) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= rows * xdim) return;
    const long long r = t / xdim;
    const int i = (int)(t % xdim);
    if (scatter) dst[r * xdim + rmap[i]] = src[t];      // panel order -> caller's rows
    else dst[t] = src[r * xdim + rmap[i]];              // caller's rows -> panel order
}

qcb_status eval_device_inner(qcb_handle_s *h, const double *dZ, const double *dmu, double *d_delta, double *d_jac,
                             double *d_hess, int want, const HostOut *host, double *E_out);

// common driver: device inputs -> device outputs (+ optional streamed host outputs)
qcb_status eval_device(qcb_handle_s *h, const double *dZ, const double *dmu, double *d_delta, double *d_jac,
                       double *d_hess, int want, const HostOut *host, double *E_out = nullptr) {
    if (!h->ext.on) return eval_device_inner(h, dZ, dmu, d_delta, d_jac, d_hess, want, host, E_out);
    qcb_handle_s::Ext &x = h->ext;
    ensure_out(h, &x.dZp, (size_t)h->N * h->z_dim);
    pack_knots_kernel<<<h->nk + 1, 128, 0, h->stream>>>(dZ, x.z_dim, x.dZp, h->z_dim, x.d_xmap, h->S * h->xdim, x.u_off,
                                                         h->m, x.dt_off, h->k0);
    ++h->launches;
    const double *mu_in = dmu;
    if (dmu && !x.row_identity) {
        const long long rows = (long long)h->S * h->K;
        ensure_out(h, &x.dmup, (size_t)rows * h->xdim);
        permute_rows_kernel<<<(unsigned)((rows * h->xdim + 255) / 256), 256, 0, h->stream>>>(dmu, x.dmup, x.d_rmap, h->xdim, rows, 0);
        ++h->launches;
        mu_in = x.dmup;
    }
    CK(cudaGetLastError());
    if (x.row_identity) return eval_device_inner(h, x.dZp, mu_in, d_delta, d_jac, d_hess, want, host, E_out);
    // delta leaves the kernels in panel order: evaluate into a scratch vector, reorder, then deliver
    ensure_out(h, &x.d_delta_p, (size_t)h->units * h->xdim);
    HostOut inner;
    if (host) { inner = *host; inner.delta = nullptr; }
    qcb_status st = eval_device_inner(h, x.dZp, mu_in, x.d_delta_p, d_jac, d_hess, want, host ? &inner : nullptr, E_out);
    if (st != QCB_OK) return st;
    permute_rows_kernel<<<(unsigned)(((long long)h->units * h->xdim + 255) / 256), 256, 0, h->stream>>>(
        x.d_delta_p, d_delta, x.d_rmap, h->xdim, h->units, 1);
    ++h->launches;
    CK(cudaGetLastError());
    if (host && host->delta)
        CK(cudaMemcpyAsync(host->delta, d_delta, (size_t)h->units * h->xdim * 8, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return QCB_OK;
}

qcb_status eval_device_inner(qcb_handle_s *h, const double *dZ, const double *dmu, double *d_delta, double *d_jac,
                             double *d_hess, int want, const HostOut *host, double *E_out) {
    const char *range_msg = "non-finite input or ||dt*G(u)||_1 beyond the supported range";
    if (h->large) {
        // new_x = 0: the norms, the squaring classes and the unit order are those of the previous call
        const bool same_x = (h->reuse & h->cache_have & 1) && !E_out && h->cache_s_slab >= 0;
        const int s = same_x ? h->cache_s_slab : slab_squarings(h, dZ);
        if (!same_x) h->cache_have = 0;
        if (s < 0) { h->err = range_msg; return QCB_ERR_RANGE; }
        eval_large(h, dZ, dmu, d_delta, d_jac, d_hess, want, host, s, E_out);
    } else {
        // No norm pre-pass on the small-d path: the kernel sizes its squaring chains for the count
        // seen by the previous call, reports the count it needs, and is re-run once if that was
        // not enough (rare: the count only changes when ||dt*G(u)||_1 crosses a power of two).
        // compact host route when the COO vectors are much larger than the distinct values
        // (dense-x Jacobians): the arenas cross PCIe, the host cores expand them
        const bool compact = host && host->compact;
        if (compact) {
            ensure_out(h, &h->d_ja_all, (size_t)h->units * h->JA);
            ensure_out(h, &h->d_ha_all, (size_t)h->units * h->HA);
            if (!h->hs_ja) CK(cudaMallocHost(&h->hs_ja, (size_t)h->units * h->JA * 8));
            if (!h->hs_ha) CK(cudaMallocHost(&h->hs_ha, (size_t)h->units * h->HA * 8));
            if (h->jsegs.empty()) {
                h->jsegs.resize(h->S);
                for (int s2 = 0; s2 < h->S; ++s2) h->jsegs[s2] = build_segments(h->h_jcode.data() + (size_t)s2 * h->jnnz_unit, h->jnnz_unit);
            }
            if (h->hsegs.empty()) {
                h->hsegs.resize(h->S);
                for (int s2 = 0; s2 < h->S; ++s2) h->hsegs[s2] = build_segments(h->h_hcode.data() + (size_t)s2 * h->hnnz_unit, h->hnnz_unit);
            }
        }
        int cap = std::min(h->s_hint, QCB_MAX_SQUARINGS);
        for (int attempt = 0;; ++attempt) {
            CK(cudaMemsetAsync(h->d_bad, 0, 2 * sizeof(int), h->stream));
            try {
                eval_small(h, dZ, dmu, d_delta, compact ? nullptr : d_jac, compact ? nullptr : d_hess, want, cap, E_out, compact);
            } catch (const SmallCapacity &sc) {
                // a pulse whose ||dt*G(u)||_1 needs more squaring levels than a unit of this size can keep
                // in shared memory: reported like any other out-of-range input (Ipopt: evaluation failed)
                char buf[200];
                snprintf(buf, sizeof buf, "||dt*G(u)||_1 needs %d squarings; the fused small-d kernel holds fewer for d = %d "
                                          "(create the handle with force_path = 2 for such pulses)", sc.squarings, h->d);
                h->err = buf;
                return QCB_ERR_RANGE;
            }
            CK(cudaMemcpyAsync(h->h_flags, h->d_bad, 2 * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
            CK(cudaStreamSynchronize(h->stream));
            if (h->h_flags[0] & 1) { h->err = range_msg; return QCB_ERR_RANGE; }
            if (!(h->h_flags[0] & 2)) break;
            cap = h->h_flags[1];
            if (attempt > 1 || cap > QCB_MAX_SQUARINGS) { h->err = range_msg; return QCB_ERR_RANGE; }
        }
        h->s_hint = std::max(h->h_flags[1], 0);
        if (host) {
            const bool wj = (want & 1) && host->jac, wh = (want & 2) && host->hess;
            if (host->delta) {
                CK(cudaMemcpyAsync(host->delta, d_delta, (size_t)h->units * h->xdim * 8, cudaMemcpyDeviceToHost, h->copy_stream));
                h->xfer[1] += (int64_t)((size_t)h->units * h->xdim * 8);
            }
            if (compact) {
                if (wj) { CK(cudaMemcpyAsync(h->hs_ja, h->d_ja_all, (size_t)h->units * h->JA * 8, cudaMemcpyDeviceToHost, h->copy_stream)); h->xfer[1] += (int64_t)h->units * h->JA * 8; }
                if (wh) { CK(cudaMemcpyAsync(h->hs_ha, h->d_ha_all, (size_t)h->units * h->HA * 8, cudaMemcpyDeviceToHost, h->copy_stream)); h->xfer[1] += (int64_t)h->units * h->HA * 8; }
                CK(cudaStreamSynchronize(h->copy_stream));
                ExpandJob jobs[2];
                int nj = 0;
                if (wj) jobs[nj++] = {&h->jsegs, h->hs_ja, h->JA, host->jac, h->jnnz_unit};
                if (wh) jobs[nj++] = {&h->hsegs, h->hs_ha, h->HA, host->hess, h->hnnz_unit};
                if (nj > 0) expand_units(jobs, nj, h->nk, 0, h->units);
            } else {
                if (wj) {
                    CK(cudaMemcpyAsync(host->jac, d_jac, (size_t)h->units * h->jnnz_unit * 8, cudaMemcpyDeviceToHost, h->copy_stream));
                    h->xfer[1] += (int64_t)((size_t)h->units * h->jnnz_unit * 8);
                }
                if (wh) {
                    CK(cudaMemcpyAsync(host->hess, d_hess, (size_t)h->units * h->hnnz_unit * 8, cudaMemcpyDeviceToHost, h->copy_stream));
                    h->xfer[1] += (int64_t)((size_t)h->units * h->hnnz_unit * 8);
                }
                CK(cudaStreamSynchronize(h->copy_stream));
            }
        }
        return QCB_OK;
    }
    if (h->async_mode && !host && !h->profiling) return QCB_OK;      // qcb_wait collects completion and the range flag
    CK(cudaStreamSynchronize(h->stream));
    if (host || h->arena_out_j || h->arena_out_h) CK(cudaStreamSynchronize(h->copy_stream));
    if (h->profiling) {
        for (size_t i = 0; i + 1 < h->prof_used; i += 2) {
            float t = 0;
            cudaEventElapsedTime(&t, h->prof_ev[i], h->prof_ev[i + 1]);
            h->prof_acc[0] += 1; h->prof_acc[1] += t; h->prof_acc[2] += h->prof_flops[i / 2];
        }
        h->prof_used = 0;
        h->prof_flops.clear();
    }
    int bad = 0;
    CK(cudaMemcpy(&bad, h->d_bad, sizeof(int), cudaMemcpyDeviceToHost));
    if (bad) { h->err = range_msg; return QCB_ERR_RANGE; }
    return QCB_OK;
}

qcb_status fail(qcb_handle_s *h, qcb_status st, const std::string &msg) {
    if (h) h->err = msg; else g_create_error = msg;
    return st;
}
qcb_status fail_cuda(qcb_handle_s *h, const CudaFail &f) {
    char buf[512];
    snprintf(buf, sizeof buf, "CUDA error '%s' in %s (qcb200.cu:%d)", cudaGetErrorString(f.e), f.what, f.line);
    cudaGetLastError();
    return fail(h, f.e == cudaErrorMemoryAllocation ? QCB_ERR_NOMEM : QCB_ERR_CUDA, buf);
}

void init_constants() {
    CK(cudaMemcpyToSymbol(c_P1, QCB_P1, sizeof(QCB_P1)));
    CK(cudaMemcpyToSymbol(c_P2, QCB_P2, sizeof(QCB_P2)));
    CK(cudaMemcpyToSymbol(c_P3, QCB_P3, sizeof(QCB_P3)));
    CK(cudaMemcpyToSymbol(c_P4, QCB_P4, sizeof(QCB_P4)));
    CK(cudaMemcpyToSymbol(c_P5, QCB_P5, sizeof(QCB_P5)));
    CK(cudaMemcpyToSymbol(c_P6, QCB_P6, sizeof(QCB_P6)));
    CK(cudaMemcpyToSymbol(c_P6_partner, QCB_P6_PARTNER, sizeof(QCB_P6_PARTNER)));
    CK(cudaMemcpyToSymbol(c_P7U, QCB_P7U, sizeof(QCB_P7U)));
    CK(cudaMemcpyToSymbol(c_P7L, QCB_P7L, sizeof(QCB_P7L)));
    CK(cudaMemcpyToSymbol(c_taylor, QCB_TAYLOR, sizeof(QCB_TAYLOR)));
}

void ensure_out(qcb_handle_s *h, double **p, size_t n) {
    if (!*p) CK(cudaMalloc(p, std::max<size_t>(n, 1) * sizeof(double)));
}

}  // namespace

// =========================================================================================
// C ABI
// =========================================================================================
extern "C" {

const char *qcb_version(void) { return "qcb200 0.1.0 (sm_100a)"; }

const char *qcb_last_error(qcb_handle h) { return h ? h->err.c_str() : g_create_error.c_str(); }

qcb_status qcb_create(qcb_handle *out, const qcb_desc *desc) {
    if (!out || !desc) return fail(nullptr, QCB_ERR_ARG, "null argument");
    *out = nullptr;
    const bool has_maps = desc->struct_bytes == (int32_t)sizeof(qcb_desc);     // older callers stop before x_index
    if (!has_maps && desc->struct_bytes != (int32_t)offsetof(qcb_desc, x_index))
        return fail(nullptr, QCB_ERR_ARG, "qcb_desc.struct_bytes mismatch");
    const int32_t *x_index = has_maps ? desc->x_index : nullptr, *row_index = has_maps ? desc->row_index : nullptr;
    if (row_index && !x_index) return fail(nullptr, QCB_ERR_ARG, "row_index needs x_index");
    if (desc->d < 1 || desc->c < 1 || desc->m < 0 || desc->N < 2 || desc->S < 1 || desc->z_dim < 1)
        return fail(nullptr, QCB_ERR_ARG, "bad dimensions in qcb_desc");
    if (desc->m > CONTRACT_MAXI) return fail(nullptr, QCB_ERR_ARG, "more than 32 drives are not supported");
    if (!desc->x_off || !desc->G_drift || (desc->m > 0 && !desc->G_drives))
        return fail(nullptr, QCB_ERR_ARG, "null array in qcb_desc");
    if (desc->jac_mode != QCB_JAC_DENSE_X && desc->jac_mode != QCB_JAC_BLOCK_DIAG)
        return fail(nullptr, QCB_ERR_ARG, "bad jac_mode");
    const int xdim = desc->d * desc->c;
    if (x_index) {
        // every state entry names a distinct place of the knot, none of them a control or the timestep
        for (int s = 0; s < desc->S; ++s) {
            std::vector<char> used((size_t)desc->z_dim, 0);
            for (int q = 0; q < xdim; ++q) {
                const int o = x_index[(size_t)s * xdim + q];
                if (o < 0 || o >= desc->z_dim || used[(size_t)o] || o == desc->dt_off || (o >= desc->u_off && o < desc->u_off + desc->m))
                    return fail(nullptr, QCB_ERR_ARG, "x_index: entries must be distinct state positions inside the knot");
                used[(size_t)o] = 1;
            }
        }
        if (row_index) {
            std::vector<char> used((size_t)xdim, 0);
            for (int q = 0; q < xdim; ++q) {
                if (row_index[q] < 0 || row_index[q] >= xdim || used[(size_t)row_index[q]]) return fail(nullptr, QCB_ERR_ARG, "row_index is not a permutation of the unit's rows");
                used[(size_t)row_index[q]] = 1;
            }
        }
    } else {
        for (int s = 0; s < desc->S; ++s)
            if (desc->x_off[s] < 0 || desc->x_off[s] + xdim > desc->z_dim) return fail(nullptr, QCB_ERR_ARG, "x_off out of the knot");
    }
    if (desc->u_off < 0 || desc->u_off + desc->m > desc->z_dim || desc->dt_off < 0 || desc->dt_off >= desc->z_dim)
        return fail(nullptr, QCB_ERR_ARG, "u_off / dt_off out of the knot");
    const int K = desc->N - 1;
    const int k0 = desc->k_begin, k1 = desc->k_end == 0 ? K : desc->k_end;
    if (k0 < 0 || k1 > K || k0 >= k1) return fail(nullptr, QCB_ERR_ARG, "bad interval slab [k_begin, k_end)");

    qcb_handle_s *h = new qcb_handle_s();
    h->run_assembly = !getenv("QCB_GATHER_ASSEMBLY");
    if (const char *e = getenv("QCB_DELIVERY_SPLIT")) h->delivery_split = std::max(1, atoi(e));
    try {
        int ndev = 0;
        if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
            cudaGetLastError();
            delete h;
            return fail(nullptr, QCB_ERR_CUDA, "no usable CUDA device: this library has no CPU fallback");
        }
        if (desc->device < 0 || desc->device >= ndev) { delete h; return fail(nullptr, QCB_ERR_ARG, "bad device ordinal"); }
        h->device = desc->device;
        CK(cudaSetDevice(h->device));
        cudaDeviceProp prop;
        CK(cudaGetDeviceProperties(&prop, h->device));
        h->n_sm = prop.multiProcessorCount;
        h->d = desc->d; h->c = desc->c; h->m = desc->m; h->N = desc->N; h->S = desc->S; h->K = K;
        h->z_dim = desc->z_dim; h->u_off = desc->u_off; h->dt_off = desc->dt_off; h->jac_mode = desc->jac_mode;
        h->k0 = k0; h->k1 = k1; h->nk = k1 - k0; h->units = h->S * h->nk; h->xdim = xdim;
        h->x_off.assign(desc->x_off, desc->x_off + desc->S);
        if (x_index) {
            // kernels see a packed knot [panel of system 0 | ... | u | dt]; the caller's layout stays in ext
            h->ext.on = true;
            h->ext.z_dim = desc->z_dim; h->ext.u_off = desc->u_off; h->ext.dt_off = desc->dt_off;
            h->ext.xmap.assign(x_index, x_index + (size_t)desc->S * xdim);
            h->ext.rmap.resize(xdim);
            for (int q = 0; q < xdim; ++q) {
                h->ext.rmap[q] = row_index ? row_index[q] : q;
                if (h->ext.rmap[q] != q) h->ext.row_identity = false;
            }
            for (int s = 0; s < desc->S; ++s) h->x_off[s] = s * xdim;
            h->u_off = desc->S * xdim; h->dt_off = h->u_off + desc->m; h->z_dim = h->dt_off + 1;
        }
        // path selection: the fused kernel needs every register of a unit in shared memory
        const size_t small_need = small_smem_doubles(h->d, h->c, h->m, 3) * sizeof(double);
        h->large = small_need > 200 * 1024;
        if (desc->force_path == 1) h->large = false;
        if (desc->force_path == 2) h->large = true;
        if (h->large) { h->dp = round_up(h->d, 64); h->cp = round_up(h->c, 64); }
        else { h->dp = h->d; h->cp = h->c; }
        // real-iso block structure [[P,-Q],[Q,P]] of every generator -> half storage / half flops
        h->iso = false;
        if (h->large && h->d % 128 == 0 && !getenv("QCB_NO_ISO")) {
            const int n = h->d / 2;
            auto is_iso = [&](const double *M) {
                for (int j = 0; j < n; ++j)
                    for (int i = 0; i < n; ++i) {
                        if (M[(i + n) + (size_t)(j + n) * h->d] != M[i + (size_t)j * h->d]) return false;
                        if (M[i + (size_t)(j + n) * h->d] != -M[(i + n) + (size_t)j * h->d]) return false;
                    }
                return true;
            };
            bool ok = true;
            for (int s = 0; s < h->S && ok; ++s) {
                ok = is_iso(desc->G_drift + (size_t)s * h->d * h->d);
                for (int j = 0; j < h->m && ok; ++j) ok = is_iso(desc->G_drives + ((size_t)s * h->m + j) * h->d * h->d);
            }
            h->iso = ok;
        }
        h->gcols = h->iso ? h->dp / 2 : h->dp;
        h->msz = (int64_t)h->dp * h->gcols;
        h->small_threads = std::min(512, std::max(32, round_up(h->d * h->d, 32)));
        const int64_t xb = h->jac_mode == QCB_JAC_DENSE_X ? (int64_t)xdim * xdim : (int64_t)h->c * h->d * h->d;
        h->jnnz_unit = xb + (int64_t)xdim * (h->m + 2);
        h->hnnz_unit = (int64_t)xdim * (h->m + 1) + (int64_t)h->m * (h->m + 1) / 2 + h->m + 1;
        const int64_t dd = h->msz, pan = (int64_t)h->dp * h->cp;
        h->JA = (int64_t)h->dp * h->dp + (h->m + 1) * pan + 2;
        h->HA = (h->m + 1) * pan + (int64_t)h->m * h->m + h->m + 2;
        h->HA += h->HA & 1;
        if (h->JA >= (1LL << 31) || h->HA >= (1LL << 31)) { delete h; return fail(nullptr, QCB_ERR_ARG, "unit too large for 31-bit slot codes"); }

        CK(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
        {   // side stream (deliveries, foreign-slab assembly, D2H): highest priority, so that its small CTAs take
            // the SM slots that free up instead of queueing behind a compute kernel's pending waves
            int lo_p = 0, hi_p = 0;
            CK(cudaDeviceGetStreamPriorityRange(&lo_p, &hi_p));
            CK(cudaStreamCreateWithPriority(&h->copy_stream, cudaStreamNonBlocking, hi_p));
        }
        for (auto &e : h->ev) CK(cudaEventCreate(&e));
        init_constants();
        // generators, padded to dp x dp
        {
            std::vector<double> g0((size_t)h->S * dd, 0.0), gd((size_t)h->S * std::max(1, h->m) * dd, 0.0);
            for (int s = 0; s < h->S; ++s) {
                const int qmax = h->iso ? h->d / 2 : h->d;    // iso: only the left block column is kept
                for (int q = 0; q < qmax; ++q)
                    for (int r = 0; r < h->d; ++r)
                        g0[(size_t)s * dd + r + (size_t)q * h->dp] = desc->G_drift[(size_t)s * h->d * h->d + r + (size_t)q * h->d];
                for (int j = 0; j < h->m; ++j)
                    for (int q = 0; q < qmax; ++q)
                        for (int r = 0; r < h->d; ++r)
                            gd[((size_t)s * h->m + j) * dd + r + (size_t)q * h->dp] =
                                desc->G_drives[((size_t)s * h->m + j) * h->d * h->d + r + (size_t)q * h->d];
            }
            CK(cudaMalloc(&h->dG0, g0.size() * 8));
            CK(cudaMalloc(&h->dGd, gd.size() * 8));
            CK(cudaMemcpy(h->dG0, g0.data(), g0.size() * 8, cudaMemcpyHostToDevice));
            CK(cudaMemcpy(h->dGd, gd.data(), gd.size() * 8, cudaMemcpyHostToDevice));
        }
        // sparse drive generators -> ELL (large-d path only; unpadded systems)
        if (h->large && h->dp == h->d && h->m > 0 && !getenv("QCB_DENSE_DRIVES")) {
            const int d = h->d;
            int W = 0;
            for (int s = 0; s < h->S; ++s)
                for (int j = 0; j < h->m; ++j) {
                    const double *Gj = desc->G_drives + ((size_t)s * h->m + j) * d * d;
                    std::vector<int> rown(d, 0), coln(d, 0);
                    for (int q = 0; q < d; ++q)
                        for (int r = 0; r < d; ++r)
                            if (Gj[r + (size_t)q * d] != 0.0) { ++rown[r]; ++coln[q]; }
                    for (int i = 0; i < d; ++i) W = std::max(W, std::max(rown[i], coln[i]));
                }
            if (W >= 1 && W <= 8) {
                W = W <= 2 ? W : (W <= 4 ? 4 : 8);     // kernel instantiations: 1, 2, 4, 8
                const size_t tot = (size_t)h->S * h->m * d * W;
                std::vector<int> idx(tot, 0), idxT(tot, 0);
                std::vector<double> val(tot, 0.0), valT(tot, 0.0);
                for (int s = 0; s < h->S; ++s)
                    for (int j = 0; j < h->m; ++j) {
                        const double *Gj = desc->G_drives + ((size_t)s * h->m + j) * d * d;
                        const size_t base = ((size_t)s * h->m + j) * d * W;
                        std::vector<int> rown(d, 0), coln(d, 0);
                        for (int q = 0; q < d; ++q)
                            for (int r = 0; r < d; ++r) {
                                const double v = Gj[r + (size_t)q * d];
                                if (v == 0.0) continue;
                                idx[base + (size_t)r * W + rown[r]] = q; val[base + (size_t)r * W + rown[r]] = v; ++rown[r];
                                idxT[base + (size_t)q * W + coln[q]] = r; valT[base + (size_t)q * W + coln[q]] = v; ++coln[q];
                            }
                    }
                CK(cudaMalloc(&h->d_ell_idx, tot * sizeof(int)));
                CK(cudaMalloc(&h->d_ellT_idx, tot * sizeof(int)));
                CK(cudaMalloc(&h->d_ell_val, tot * 8));
                CK(cudaMalloc(&h->d_ellT_val, tot * 8));
                CK(cudaMemcpy(h->d_ell_idx, idx.data(), tot * sizeof(int), cudaMemcpyHostToDevice));
                CK(cudaMemcpy(h->d_ellT_idx, idxT.data(), tot * sizeof(int), cudaMemcpyHostToDevice));
                CK(cudaMemcpy(h->d_ell_val, val.data(), tot * 8, cudaMemcpyHostToDevice));
                CK(cudaMemcpy(h->d_ellT_val, valT.data(), tot * 8, cudaMemcpyHostToDevice));
                h->ell_w = W;
                h->sparse_drives = true;
            }
        }
        if (h->ext.on) {
            CK(cudaMalloc(&h->ext.d_xmap, h->ext.xmap.size() * sizeof(int)));
            CK(cudaMemcpy(h->ext.d_xmap, h->ext.xmap.data(), h->ext.xmap.size() * sizeof(int), cudaMemcpyHostToDevice));
            CK(cudaMalloc(&h->ext.d_rmap, h->ext.rmap.size() * sizeof(int)));
            CK(cudaMemcpy(h->ext.d_rmap, h->ext.rmap.data(), h->ext.rmap.size() * sizeof(int), cudaMemcpyHostToDevice));
        }
        CK(cudaMalloc(&h->d_xoff, h->S * sizeof(int)));
        CK(cudaMemcpy(h->d_xoff, h->x_off.data(), h->S * sizeof(int), cudaMemcpyHostToDevice));
        CK(cudaMalloc(&h->d_norm, sizeof(unsigned long long)));
        CK(cudaMalloc(&h->d_bad, 2 * sizeof(int)));
        CK(cudaMemset(h->d_bad, 0, 2 * sizeof(int)));
        CK(cudaMallocHost(&h->h_flags, 2 * sizeof(int)));
        build_native_codes(h, true, true);
    } catch (const CudaFail &f) {
        qcb_status st = fail_cuda(nullptr, f);
        qcb_destroy(h);
        return st;
    }
    *out = h;
    return QCB_OK;
}

void qcb_destroy(qcb_handle h) {
    if (!h) return;
    cudaSetDevice(h->device);
    free_scratch(h);
    void *ptrs[] = {h->dG0, h->dGd, h->d_xoff, h->d_jcode, h->d_hcode, h->d_norm, h->d_bad, h->dZ, h->dmu, h->d_delta, h->d_jac, h->d_hess,
                    h->d_Eall, h->d_Escan, h->d_states, h->d_xinit, h->d_ja_all, h->d_ha_all,
                    h->d_ell_idx, h->d_ellT_idx, h->d_ell_val, h->d_ellT_val, h->d_unit_norm, h->d_perm, h->d_jpieces, h->d_hpieces,
                    h->ext.d_xmap, h->ext.d_rmap, h->ext.dZ_ext, h->ext.dZp, h->ext.dmup, h->ext.d_delta_p};
    for (void *p : ptrs)
        if (p) cudaFree(p);
    if (h->h_flags) cudaFreeHost(h->h_flags);
    if (h->hs_ja) cudaFreeHost(h->hs_ja);
    if (h->hs_ha) cudaFreeHost(h->hs_ha);
    for (auto &e : h->ev)
        if (e) cudaEventDestroy(e);
    for (auto &e : h->chunk_events) cudaEventDestroy(e);
    for (auto &e : h->copied) if (e) cudaEventDestroy(e);
    for (auto &e : h->part_done) if (e) cudaEventDestroy(e);
    for (auto &e : h->prof_ev) cudaEventDestroy(e);
    if (h->stream) cudaStreamDestroy(h->stream);
    if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
    if (h->stream2) cudaStreamDestroy(h->stream2);
    for (int i = 0; i < 2; ++i) if (h->lane_ev[i]) cudaEventDestroy(h->lane_ev[i]);
    for (int i = 0; i < 3; ++i) if (h->ship_ev[i]) cudaEventDestroy(h->ship_ev[i]);
    delete h;
}

int64_t qcb_defect_dim(qcb_handle h) { return h ? (int64_t)h->units * h->xdim : 0; }
int64_t qcb_jacobian_nnz(qcb_handle h) { return h ? (int64_t)h->units * h->jnnz_unit : 0; }
int64_t qcb_hessian_nnz(qcb_handle h) { return h ? (int64_t)h->units * h->hnnz_unit : 0; }
int64_t qcb_kernel_launches(qcb_handle h) { return h ? h->launches : 0; }
void *qcb_stream(qcb_handle h) { return h ? (void *)h->stream : nullptr; }

qcb_status qcb_get_transfer_bytes(qcb_handle h, int64_t bytes[2]) {
    if (!h || !bytes) return QCB_ERR_ARG;
    bytes[0] = h->xfer[0]; bytes[1] = h->xfer[1];
    return QCB_OK;
}

qcb_status qcb_profile_enable(qcb_handle h, int32_t on) {
    if (!h) return QCB_ERR_ARG;
    h->profiling = on != 0;
    for (double &v : h->prof_acc) v = 0;
    return QCB_OK;
}
qcb_status qcb_profile_read(qcb_handle h, double out[4]) {
    if (!h || !out) return QCB_ERR_ARG;
    for (int i = 0; i < 4; ++i) out[i] = h->prof_acc[i];
    return QCB_OK;
}

qcb_status qcb_get_timings(qcb_handle h, double ms[4]) {
    if (!h || !ms) return QCB_ERR_ARG;
    for (int i = 0; i < 4; ++i) ms[i] = h->ms[i];
    return QCB_OK;
}

qcb_status qcb_jacobian_structure(qcb_handle h, int64_t row_base, int64_t *rows, int64_t *cols) {
    if (!h || !rows || !cols) return fail(h, QCB_ERR_ARG, "null argument");
    if (h->custom_j) return fail(h, QCB_ERR_STRUCTURE, "a caller-supplied Jacobian structure is in force");
    int64_t p = 0;
    for (int s = 0; s < h->S; ++s) {
        std::vector<Entry> pat = jac_pattern(h, s);
        for (int kl = 0; kl < h->nk; ++kl) {
            const int64_t rb = row_base + ((int64_t)s * h->nk + kl) * h->xdim, cb = (int64_t)(h->k0 + kl) * h->ez();
            for (auto &e : pat) { rows[p] = rb + e.row + 1; cols[p] = cb + e.col + 1; ++p; }
        }
    }
    return QCB_OK;
}

qcb_status qcb_hessian_structure(qcb_handle h, int64_t *rows, int64_t *cols) {
    if (!h || !rows || !cols) return fail(h, QCB_ERR_ARG, "null argument");
    if (h->custom_h) return fail(h, QCB_ERR_STRUCTURE, "a caller-supplied Hessian structure is in force");
    int64_t p = 0;
    for (int s = 0; s < h->S; ++s) {
        std::vector<Entry> pat = hess_pattern(h, s);
        for (int kl = 0; kl < h->nk; ++kl) {
            const int64_t cb = (int64_t)(h->k0 + kl) * h->ez();
            for (auto &e : pat) { rows[p] = cb + e.row + 1; cols[p] = cb + e.col + 1; ++p; }
        }
    }
    return QCB_OK;
}

qcb_status qcb_set_structure(qcb_handle h, int64_t row_base, const int64_t *jr, const int64_t *jc, int64_t jn,
                             const int64_t *hr, const int64_t *hc, int64_t hn) {
    if (!h) return QCB_ERR_ARG;
    Codes cd{h};
    const int d = h->d, xd = h->xdim, m = h->m;
    const int64_t EZ = h->ez(), EU = h->eu(), ET = h->et();
    // caller-side knot offset -> panel entry of system s (-1: not a state entry); defect row -> panel entry
    std::vector<std::vector<int>> x_of(h->S, std::vector<int>((size_t)EZ, -1));
    for (int s = 0; s < h->S; ++s)
        for (int q = 0; q < xd; ++q) x_of[s][(size_t)h->ecol_x(s, q)] = q;
    std::vector<int> panel_of_row(xd);
    for (int r = 0; r < xd; ++r) panel_of_row[(size_t)h->erow(r)] = r;
    try {
        CK(cudaSetDevice(h->device));
        if (jr && jc && jn > 0) {
            if (jn % h->units) return fail(h, QCB_ERR_STRUCTURE, "Jacobian nnz is not a multiple of the number of units");
            const int64_t per = jn / h->units;
            std::vector<uint32_t> codes((size_t)h->S * per);
            for (int64_t p = 0; p < jn; ++p) {
                const int64_t un = p / per, e = p % per;
                const int s = (int)(un / h->nk), kl = (int)(un % h->nk), k = h->k0 + kl;
                const int64_t r_ext = jr[p] - 1 - row_base - un * xd;
                const int64_t col = jc[p] - 1 - (int64_t)k * EZ;
                if (r_ext < 0 || r_ext >= xd) return fail(h, QCB_ERR_STRUCTURE, "Jacobian row outside its unit: entries must be grouped [system][interval]");
                if (col < 0 || col >= 2 * EZ) return fail(h, QCB_ERR_STRUCTURE, "Jacobian column outside knots k, k+1 of its unit");
                const int r = panel_of_row[(size_t)r_ext];
                uint32_t code = cd.jZero();
                if (col < EZ && x_of[s][(size_t)col] >= 0) {
                    const int qx = x_of[s][(size_t)col];
                    if (qx / d == r / d) code = cd.jE(r % d, qx % d);
                } else if (col >= EU && col < EU + m) code = cd.jU((int)(col - EU), r);
                else if (col == ET) code = cd.jT(r);
                else if (col >= EZ && x_of[s][(size_t)(col - EZ)] == r) code = cd.jOne();
                uint32_t &slot = codes[(size_t)s * per + e];
                if (kl == 0) slot = code;
                else if (slot != code) return fail(h, QCB_ERR_STRUCTURE, "Jacobian pattern differs between intervals");
            }
            upload_codes(h, &h->d_jcode, codes);
            if (per != h->jnnz_unit) free_scratch(h);     // staging buffers were sized with the old nnz
            h->jnnz_unit = per;
            h->custom_j = true;
            h->h_jcode = codes;
            h->jsegs.clear(); h->direct_built = false;
            if (h->d_jpieces) { CK(cudaFree(h->d_jpieces)); h->d_jpieces = nullptr; }
            if (h->d_jac) { CK(cudaFree(h->d_jac)); h->d_jac = nullptr; }
        }
        if (hr && hc && hn > 0) {
            if (hn % h->units) return fail(h, QCB_ERR_STRUCTURE, "Hessian nnz is not a multiple of the number of units");
            const int64_t per = hn / h->units;
            std::vector<uint32_t> codes((size_t)h->S * per);
            auto classify = [&](int s, int64_t v, int &kind, int &idx) {   // 0 x, 1 u, 2 dt, 3 other
                if (x_of[s][(size_t)v] >= 0) { kind = 0; idx = x_of[s][(size_t)v]; }
                else if (v >= EU && v < EU + m) { kind = 1; idx = (int)(v - EU); }
                else if (v == ET) { kind = 2; idx = 0; }
                else { kind = 3; idx = 0; }
            };
            std::vector<std::map<std::pair<int64_t, int64_t>, int64_t>> seen_pairs(h->S);   // pair -> first entry
            for (int64_t p = 0; p < hn; ++p) {
                const int64_t un = p / per, e = p % per;
                const int s = (int)(un / h->nk), kl = (int)(un % h->nk), k = h->k0 + kl;
                const int64_t a = hr[p] - 1 - (int64_t)k * EZ, b = hc[p] - 1 - (int64_t)k * EZ;
                if (a < 0 || a >= EZ || b < 0 || b >= EZ)
                    return fail(h, QCB_ERR_STRUCTURE, "Hessian entry outside the knot of its unit: entries must be grouped [system][interval]");
                int ka, ia, kb, ib;
                classify(s, a, ka, ia);
                classify(s, b, kb, ib);
                if (ka > kb) { std::swap(ka, kb); std::swap(ia, ib); }
                uint32_t code = cd.hZero();
                if (ka == 0 && kb == 1) code = cd.hXU(ib, ia);
                else if (ka == 0 && kb == 2) code = cd.hXT(ia);
                else if (ka == 1 && kb == 1) code = cd.hUU(std::min(ia, ib), std::max(ia, ib));
                else if (ka == 1 && kb == 2) code = cd.hUT(ia);
                else if (ka == 2 && kb == 2) code = cd.hTT();
                // MOI sums entries that name the same variable pair, in either order: when the adopted
                // structure lists a pair twice (both triangles, or a plain repeat) only its first
                // occurrence carries the value, later ones are filled with 0
                if (code != cd.hZero()) {
                    const std::pair<int64_t, int64_t> key{std::min(a, b), std::max(a, b)};
                    if (kl == 0) {
                        if (!seen_pairs[s].insert({key, e}).second) code = cd.hZero();
                    } else {
                        auto it = seen_pairs[s].find(key);
                        if (it != seen_pairs[s].end() && it->second != e) code = cd.hZero();
                    }
                }
                uint32_t &slot = codes[(size_t)s * per + e];
                if (kl == 0) slot = code;
                else if (slot != code) return fail(h, QCB_ERR_STRUCTURE, "Hessian pattern differs between intervals");
            }
            upload_codes(h, &h->d_hcode, codes);
            if (per != h->hnnz_unit) free_scratch(h);
            h->hnnz_unit = per;
            h->custom_h = true;
            h->h_hcode = codes;
            h->hsegs.clear(); h->direct_built = false;
            if (h->d_hpieces) { CK(cudaFree(h->d_hpieces)); h->d_hpieces = nullptr; }
            if (h->d_hess) { CK(cudaFree(h->d_hess)); h->d_hess = nullptr; }
        }
    } catch (const CudaFail &f) {
        return fail_cuda(h, f);
    }
    return QCB_OK;
}

qcb_status qcb_eval_all_device(qcb_handle h, const double *dZ, const double *dmu, double *d_delta,
                               double *d_jac, double *d_hess) {
    if (!h || !dZ) return fail(h, QCB_ERR_ARG, "null argument");
    h->dZ_valid = false;        // what the scratch holds now belongs to the caller's dZ, not to the staged copy
    try {
        CK(cudaSetDevice(h->device));
        int want = (d_jac ? 1 : 0) | ((d_hess && dmu) ? 2 : 0);
        if (!d_delta) { ensure_out(h, &h->d_delta, (size_t)h->units * h->xdim); d_delta = h->d_delta; }
        return eval_device(h, dZ, dmu, d_delta, d_jac, d_hess, want, nullptr);
    } catch (const CudaFail &f) {
        return fail_cuda(h, f);
    }
}

int64_t qcb_arena_doubles(qcb_handle h, int32_t which) { return h ? (which == 0 ? h->JA : h->HA) : 0; }

qcb_status qcb_eval_arenas_device(qcb_handle h, const double *dZ, const double *dmu, double *d_delta,
                                  double *d_jac_arenas, double *d_hess_arenas) {
    if (!h || !dZ) return fail(h, QCB_ERR_ARG, "null argument");
    if (!h->large) return fail(h, QCB_ERR_ARG, "arena delivery exists on the large-d path only (small units are cheaper to deliver assembled)");
    h->dZ_valid = false;
    try {
        CK(cudaSetDevice(h->device));
        const int want = ((d_jac_arenas && (h->arena_parts & 2)) ? 1 : 0) | ((d_hess_arenas && dmu) ? 2 : 0);
        if (!d_delta) { ensure_out(h, &h->d_delta, (size_t)h->units * h->xdim); d_delta = h->d_delta; }
        h->arena_out_j = d_jac_arenas; h->arena_out_h = d_hess_arenas;
        qcb_status st;
        try { st = eval_device(h, dZ, dmu, d_delta, nullptr, nullptr, want, nullptr); }
        catch (...) { h->arena_out_j = h->arena_out_h = nullptr; throw; }
        h->arena_out_j = h->arena_out_h = nullptr;
        return st;
    } catch (const CudaFail &f) {
        return fail_cuda(h, f);
    }
}

qcb_status qcb_eval_arenas_device_x(qcb_handle h, const double *dZ, const double *dmu, int32_t new_x, double *d_delta,
                                    double *d_jac_arenas, double *d_hess_arenas) {
    if (!h) return QCB_ERR_ARG;
    h->reuse = new_x ? 0 : 3;
    qcb_status st = qcb_eval_arenas_device(h, dZ, dmu, d_delta, d_jac_arenas, d_hess_arenas);
    h->reuse = 0;
    return st;
}

qcb_status qcb_set_arena_parts(qcb_handle h, int32_t parts) {
    if (!h || parts < 1 || parts > 3) return QCB_ERR_ARG;
    h->arena_parts = parts;
    return QCB_OK;
}

qcb_status qcb_set_async(qcb_handle h, int32_t on) {
    if (!h) return QCB_ERR_ARG;
    if (on && h->ext.on) return fail(h, QCB_ERR_ARG, "asynchronous mode is not available on a handle created with a state index map");
    h->async_mode = on != 0;
    return QCB_OK;
}

qcb_status qcb_wait(qcb_handle h, int32_t what) {
    if (!h) return QCB_ERR_ARG;
    try {
        CK(cudaSetDevice(h->device));
        if (what >= 1 && what <= 3) {          // the last delivery of the E part (1), of the rest of the Jacobian arena (2),
            if (h->ship_ev[what - 1]) CK(cudaEventSynchronize(h->ship_ev[what - 1]));   // of the Hessian arena (3) has completed
            return QCB_OK;
        }
        CK(cudaStreamSynchronize(h->stream));
        if (h->stream2) CK(cudaStreamSynchronize(h->stream2));
        CK(cudaStreamSynchronize(h->copy_stream));
        int bad = 0;
        CK(cudaMemcpy(&bad, h->d_bad, sizeof(int), cudaMemcpyDeviceToHost));
        if (bad & 1) { h->err = "non-finite input or ||dt*G(u)||_1 beyond the supported range"; return QCB_ERR_RANGE; }
        return QCB_OK;
    } catch (const CudaFail &f) {
        return fail_cuda(h, f);
    }
}

qcb_status qcb_assemble_device(qcb_handle h, const double *d_jac_arenas, const double *d_hess_arenas, int64_t n_units,
                               int32_t nk_src, double *d_jac_vals, double *d_hess_vals, int32_t parts) {
    if (parts < 1 || parts > 3) return fail(h, QCB_ERR_ARG, "parts must be 1, 2 or 3");
    if (!h || n_units < 0 || nk_src < 1) return fail(h, QCB_ERR_ARG, "bad argument");
    if ((d_jac_arenas && !d_jac_vals) || (d_hess_arenas && !d_hess_vals)) return fail(h, QCB_ERR_ARG, "null output");
    if (n_units == 0) return QCB_OK;
    try {
        CK(cudaSetDevice(h->device));
        // on the side stream: the Ipopt rank assembles foreign slabs while its own kernels keep running
        cudaStream_t st = h->copy_stream;
        if (d_jac_arenas) {
            if (!h->d_jpieces) upload_pieces(h, h->h_jcode, h->jnnz_unit, &h->d_jpieces, &h->n_jpieces);
            // parts: 1 = entries copied from the E part of the arena (the x-block), 2 = everything else
            const unsigned e_part = (unsigned)((long long)h->dp * h->dp);
            assemble_runs_kernel<<<(unsigned)((size_t)n_units * h->n_jpieces), 256, 0, st>>>(
                d_jac_arenas, h->JA, h->d_jcode, h->jnnz_unit, h->d_jpieces, h->n_jpieces, d_jac_vals, 0, nk_src, 0, nullptr,
                parts == 3 ? 0u : e_part, parts);
            ++h->launches;
        }
        if (d_hess_arenas) {
            if (!h->d_hpieces) upload_pieces(h, h->h_hcode, h->hnnz_unit, &h->d_hpieces, &h->n_hpieces);
            assemble_runs_kernel<<<(unsigned)((size_t)n_units * h->n_hpieces), 256, 0, st>>>(
                d_hess_arenas, h->HA, h->d_hcode, h->hnnz_unit, h->d_hpieces, h->n_hpieces, d_hess_vals, 0, nk_src, 0, nullptr);
            ++h->launches;
        }
        CK(cudaGetLastError());
        if (!h->async_mode) CK(cudaStreamSynchronize(st));
        return QCB_OK;
    } catch (const CudaFail &f) {
        return fail_cuda(h, f);
    }
}

qcb_status qcb_eval_all_device_x(qcb_handle h, const double *dZ, const double *dmu, int32_t new_x, double *d_delta,
                                 double *d_jac, double *d_hess) {
    if (!h) return QCB_ERR_ARG;
    h->reuse = new_x ? 0 : 3;
    qcb_status st = qcb_eval_all_device(h, dZ, dmu, d_delta, d_jac, d_hess);
    h->reuse = 0;
    return st;
}

qcb_status qcb_eval_all_x(qcb_handle h, const double *Z, const double *mu, int32_t new_x, double *delta, double *jac,
                          double *hess) {
    if (!h || !Z) return fail(h, QCB_ERR_ARG, "null argument");
    if (hess && !mu) return fail(h, QCB_ERR_ARG, "Hessian values need the multipliers");
    try {
        // new_x = 0 with the previous iterate's intermediates still in place: Z is not even uploaded again
        const bool keep_z = !new_x && h->large && (h->cache_have & 1) && h->dZ_valid;
        h->reuse = keep_z ? 3 : 0;
        CK(cudaSetDevice(h->device));
        const size_t ezd = (size_t)h->ez();          // knot width of the caller's Z
        const size_t nz = (size_t)h->N * ezd, nmu = (size_t)h->S * h->K * h->xdim;
        double **dZ_in = h->ext.on ? &h->ext.dZ_ext : &h->dZ;
        // large-d path: send the compact value arenas over PCIe and expand on the host cores
        // (QCB_E2E_DIRECT=1 restores device-side assembly + full-size D2H)
        // (small-d path: only when the COO vectors are >= 4x larger than the arenas, e.g. dense-x Jacobians)
        const bool compact = (jac || hess) && !getenv("QCB_E2E_DIRECT") &&
                             (h->large || (h->jnnz_unit + h->hnnz_unit) >= 4 * (h->JA + h->HA));
        ensure_out(h, dZ_in, nz);
        ensure_out(h, &h->d_delta, (size_t)h->units * h->xdim);
        if (jac && !compact) ensure_out(h, &h->d_jac, (size_t)h->units * h->jnnz_unit);
        if (hess) ensure_out(h, &h->dmu, nmu);
        if (hess && !compact) ensure_out(h, &h->d_hess, (size_t)h->units * h->hnnz_unit);
        CK(cudaEventRecord(h->ev[0], h->stream));
        // only the knots of this slab (plus the halo knot k_end) are needed on the device
        if (!keep_z)
            CK(cudaMemcpyAsync(*dZ_in + (size_t)h->k0 * ezd, Z + (size_t)h->k0 * ezd,
                               (size_t)(h->nk + 1) * ezd * 8, cudaMemcpyHostToDevice, h->stream));
        h->dZ_valid = false;
        if (hess) {
            for (int s = 0; s < h->S; ++s) {
                const size_t off = ((size_t)s * h->K + h->k0) * h->xdim;
                CK(cudaMemcpyAsync(h->dmu + off, mu + off, (size_t)h->nk * h->xdim * 8, cudaMemcpyHostToDevice, h->stream));
            }
        }
        h->xfer[0] = (keep_z ? 0 : (int64_t)(h->nk + 1) * (int64_t)ezd * 8) + (hess ? (int64_t)h->S * h->nk * h->xdim * 8 : 0);
        h->xfer[1] = 0;   // counted where the device->host copies are issued
        h->xfer_compact = compact;
        CK(cudaEventRecord(h->ev[1], h->stream));
        HostOut ho{delta, jac, hess, compact};
        int want = (jac ? 1 : 0) | (hess ? 2 : 0);
        // calls that deliver the Jacobian are bound by the host side: pick their store-thread count by measurement
        static const bool tune_on = !getenv("QCB_HOST_THREADS") && !(getenv("QCB_HOST_TUNE") && atoi(getenv("QCB_HOST_TUNE")) == 0);
        HostTuner *tn = (tune_on && h->large && compact && jac) ? &h->tune[hess ? 1 : 0] : nullptr;
        const int pool = tn ? HostPool::get().size() : 0;
        if (tn) { const int t = tn->next(pool); g_expand_limit.store(t >= pool ? 0 : t); }
        const auto wall0 = std::chrono::steady_clock::now();
        // in compact mode no device COO buffer exists (the arenas travel instead): `want` alone says
        // what is evaluated
        qcb_status st = eval_device(h, *dZ_in, hess ? h->dmu : nullptr, h->d_delta,
                                    (jac && !compact) ? h->d_jac : nullptr, (hess && !compact) ? h->d_hess : nullptr,
                                    want, &ho);
        CK(cudaEventRecord(h->ev[2], h->stream));
        CK(cudaEventSynchronize(h->ev[2]));
        if (tn) {
            g_expand_limit.store(0);
            if (st == QCB_OK) tn->record(pool, std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - wall0).count());
        }
        float a = 0, b = 0;
        cudaEventElapsedTime(&a, h->ev[0], h->ev[1]);
        cudaEventElapsedTime(&b, h->ev[1], h->ev[2]);
        h->ms[0] = a; h->ms[1] = b; h->ms[2] = 0; h->ms[3] = a + b;
        h->reuse = 0;
        h->dZ_valid = st == QCB_OK;
        return st;
    } catch (const CudaFail &f) {
        h->reuse = 0;
        g_expand_limit.store(0);
        return fail_cuda(h, f);
    }
}

qcb_status qcb_eval_all(qcb_handle h, const double *Z, const double *mu, double *delta, double *jac, double *hess) {
    return qcb_eval_all_x(h, Z, mu, 1, delta, jac, hess);
}
qcb_status qcb_eval_defect_x(qcb_handle h, const double *Z, int32_t new_x, double *delta) {
    if (!delta) return fail(h, QCB_ERR_ARG, "null output");
    return qcb_eval_all_x(h, Z, nullptr, new_x, delta, nullptr, nullptr);
}
qcb_status qcb_eval_jacobian_x(qcb_handle h, const double *Z, int32_t new_x, double *jac) {
    if (!jac) return fail(h, QCB_ERR_ARG, "null output");
    return qcb_eval_all_x(h, Z, nullptr, new_x, nullptr, jac, nullptr);
}
qcb_status qcb_eval_hessian_x(qcb_handle h, const double *Z, int32_t new_x, const double *mu, double *hess) {
    if (!hess) return fail(h, QCB_ERR_ARG, "null output");
    return qcb_eval_all_x(h, Z, mu, new_x, nullptr, nullptr, hess);
}

qcb_status qcb_eval_defect(qcb_handle h, const double *Z, double *delta) {
    if (!delta) return fail(h, QCB_ERR_ARG, "null output");
    return qcb_eval_all(h, Z, nullptr, delta, nullptr, nullptr);
}
qcb_status qcb_eval_jacobian(qcb_handle h, const double *Z, double *jac) {
    if (!jac) return fail(h, QCB_ERR_ARG, "null output");
    return qcb_eval_all(h, Z, nullptr, nullptr, jac, nullptr);
}
qcb_status qcb_eval_hessian(qcb_handle h, const double *Z, const double *mu, double *hess) {
    if (!hess) return fail(h, QCB_ERR_ARG, "null output");
    return qcb_eval_all(h, Z, mu, nullptr, nullptr, hess);
}

qcb_status qcb_rollout(qcb_handle h, const double *Z, const double *x_init, double *states) {
    if (!h || !Z || !states) return fail(h, QCB_ERR_ARG, "null argument");
    if (h->k0 != 0 || h->k1 != h->K) return fail(h, QCB_ERR_ARG, "a rollout needs a handle that owns the whole trajectory");
    if (h->ext.on) return fail(h, QCB_ERR_ARG, "rollouts are not available on a handle created with a state index map");
    try {
        CK(cudaSetDevice(h->device));
        const size_t nz = (size_t)h->N * h->z_dim, dd = (size_t)h->d * h->d, ns = (size_t)h->S * h->N * h->xdim;
        ensure_out(h, &h->dZ, nz);
        ensure_out(h, &h->d_delta, (size_t)h->units * h->xdim);
        ensure_out(h, &h->d_Eall, (size_t)h->units * dd);
        ensure_out(h, &h->d_states, ns);
        CK(cudaMemcpyAsync(h->dZ, Z, nz * 8, cudaMemcpyHostToDevice, h->stream));
        if (x_init) {
            ensure_out(h, &h->d_xinit, (size_t)h->S * h->xdim);
            CK(cudaMemcpyAsync(h->d_xinit, x_init, (size_t)h->S * h->xdim * 8, cudaMemcpyHostToDevice, h->stream));
        }
        // every propagator E_k in parallel (polynomial + squarings only), then the sequential chain
        qcb_status st = eval_device(h, h->dZ, nullptr, h->d_delta, nullptr, nullptr, 0, nullptr, h->d_Eall);
        if (st != QCB_OK) return st;
        // Large matrices: the chain x_{k+1} = E_k x_k is a scan.  P_k = E_k ... E_0 for every k by a
        // Hillis-Steele prefix over the knots -- ceil(log2 K) rounds of batched d x d x d DMMA products on the
        // whole GPU instead of K dependent panel products on one SM -- then every state is P_k x_0 in parallel.
        // (Left products: round r combines P_k <- P_k * P_{k - 2^r}.)
        if (h->large && h->dp == h->d && h->K >= 8 && !getenv("QCB_ROLLOUT_CHAIN")) {
            ensure_out(h, &h->d_Escan, (size_t)h->units * dd);
            double *A = h->d_Eall, *B = h->d_Escan;
            const int K = h->K;
            for (int step = 1; step < K; step <<= 1) {
                for (int sys = 0; sys < h->S; ++sys) {
                    double *As = A + (size_t)sys * K * dd, *Bs = B + (size_t)sys * K * dd;
                    CK(cudaMemcpyAsync(Bs, As, (size_t)step * dd * 8, cudaMemcpyDeviceToDevice, h->stream));
                    GemmArgs g;
                    memset(&g, 0, sizeof(g));
                    g.M = g.N = g.K = h->dp;
                    g.alpha = 1.0; g.ndir = 1; g.nterm = 1; g.unit0_global = 0; g.nk = K;
                    auto opnd = [&](const double *p) { Opnd o; o.p = p; o.su = (long long)dd; o.sd = 0; o.ss = 0; o.ld = h->dp; o.pad_ = 0; return o; };
                    g.term[0].a = opnd(As + (size_t)step * dd);     // P_k,  k >= step
                    g.term[0].b = opnd(As);                         // P_{k - step}
                    g.dst = opnd(Bs + (size_t)step * dd);
                    launch_gemm(h, g, K - step, h->stream);
                }
                std::swap(A, B);
            }
            rollout_apply_kernel<<<(unsigned)(h->S * K), 256, 0, h->stream>>>(A, x_init ? h->d_xinit : nullptr, h->dZ, h->d_xoff,
                                                                             h->d, h->c, K, h->d_states);
            ++h->launches;
            CK(cudaGetLastError());
            CK(cudaMemcpyAsync(states, h->d_states, ns * 8, cudaMemcpyDeviceToHost, h->stream));
            CK(cudaStreamSynchronize(h->stream));
            return QCB_OK;
        }
        const size_t smem = 2 * (size_t)h->xdim * sizeof(double);
        if (smem > 200 * 1024) return fail(h, QCB_ERR_ARG, "state panel too large for the rollout kernel");
        CK(cudaFuncSetAttribute(rollout_chain_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        rollout_chain_kernel<<<h->S, 256, smem, h->stream>>>(h->d_Eall, x_init ? h->d_xinit : nullptr, h->dZ, h->z_dim,
                                                             h->d_xoff, h->d, h->c, h->K, h->d_states);
        ++h->launches;
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(states, h->d_states, ns * 8, cudaMemcpyDeviceToHost, h->stream));
        CK(cudaStreamSynchronize(h->stream));
        return QCB_OK;
    } catch (const CudaFail &f) {
        return fail_cuda(h, f);
    }
}

int32_t qcb_host_threads(void) { return HostPool::get().size(); }
int32_t qcb_host_threads_in_use(qcb_handle h, int32_t with_hessian) {
    if (!h) return 0;
    const HostTuner &t = h->tune[with_hessian ? 1 : 0];
    return t.chosen ? t.chosen : HostPool::get().size();
}

// out[0..n) = v by the host pool with streaming stores (2 MiB pieces, handed out dynamically)
static void host_fill_pool(double *out, size_t n, double v) {
    static const bool has512 = __builtin_cpu_supports("avx512f");
    const size_t piece = 1u << 18;
    std::atomic<size_t> next{0};
    std::function<void()> work = [&]() {
        for (;;) {
            const size_t b = next.fetch_add(piece);
            if (b >= n) break;
            const size_t len = std::min(piece, n - b);
            if (has512) fill_run_avx512(out + b, v, len);
            else for (size_t i = 0; i < len; ++i) out[b + i] = v;
        }
        if (has512) _mm_sfence();
    };
    HostPool::get().run(work);
}

qcb_status qcb_host_stream_fill(void *dst, size_t bytes, double *seconds) {
    if (!dst || bytes < 8) return QCB_ERR_ARG;
    const auto t0 = std::chrono::steady_clock::now();
    host_fill_pool(static_cast<double *>(dst), bytes / 8, 1.0);
    if (seconds) *seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    return QCB_OK;
}

qcb_status qcb_host_register(void *ptr, size_t bytes) {
    if (!ptr || !bytes) return QCB_ERR_ARG;
    cudaError_t e = cudaHostRegister(ptr, bytes, cudaHostRegisterDefault);
    if (e == cudaErrorHostMemoryAlreadyRegistered) { cudaGetLastError(); return QCB_OK; }
    if (e != cudaSuccess) { cudaGetLastError(); return QCB_ERR_CUDA; }
    return QCB_OK;
}
void qcb_host_unregister(void *ptr) { if (ptr && cudaHostUnregister(ptr) != cudaSuccess) cudaGetLastError(); }

qcb_status qcb_host_alloc(void **ptr, size_t bytes) {
    if (!ptr) return QCB_ERR_ARG;
    cudaError_t e = cudaMallocHost(ptr, bytes);
    if (e != cudaSuccess) { cudaGetLastError(); *ptr = nullptr; return QCB_ERR_NOMEM; }
    return QCB_OK;
}
void qcb_host_free(void *ptr) { if (ptr) cudaFreeHost(ptr); }

}  // extern "C"

#include "qcb_fidelity.cuh"
#include "qcb_aux.cuh"
#include "qcb_global.cuh"

// =========================================================================================
// Allocation-free batched fidelity (declared in include/qcb200.h next to qcb_fidelity_eval)
// =========================================================================================
// Scratch for the projections p[b][r] = g_r . x_b lives in this fixed device array instead of in
// memory obtained from cudaMalloc.  Every GPU that loads the library carries these 512 KiB until
// the process exits.  In host code `g_noalloc_proj` is only a stand-in for the symbol, not an
// address a kernel can use: the usable device address is looked up once per device with
// cudaGetSymbolAddress (below) and that pointer is what the kernels receive.
#define QCB_NOALLOC_MAX_PROJ 65536
static __device__ __align__(256) double g_noalloc_proj[QCB_NOALLOC_MAX_PROJ];

namespace qcb {
__global__ void __launch_bounds__(256)
fid_project_batch_kernel(const double *G, const double *X, double *p, int64_t n, int64_t R) {
    const int64_t r = blockIdx.x, b = blockIdx.y;
    const double *g = G + r * n, *x = X + b * n;
    double v = 0.0;
    for (int64_t i = threadIdx.x; i < n; i += blockDim.x) v = fma(g[i], x[i], v);
    __shared__ double red[8];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0;
        for (int w = 0; w < 8; ++w) s += red[w];
        p[b * R + r] = s;
    }
}
__global__ void fid_grad_batch_kernel(const double *G, const double *p, double *grad, double *F,
                                      int64_t n, int64_t R, double w, int linear) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, b = blockIdx.y;
    const double *pb = p + b * R;
    if (linear) {
        if (i < n && grad) grad[b * n + i] = w * G[i];
        if (i == 0 && F) F[b] = w * pb[0];
        return;
    }
    if (i < n && grad) {
        double v = 0.0;
        for (int64_t r = 0; r < R; ++r) v = fma(G[r * n + i], pb[r], v);
        grad[b * n + i] = 2.0 * w * v;
    }
    if (i == 0 && F) {
        double f = 0.0;
        for (int64_t r = 0; r < R; ++r) f = fma(pb[r], pb[r], f);
        F[b] = w * f;
    }
}
}  // namespace qcb

namespace {
constexpr int kMaxDevices = 64;
// Device address of g_noalloc_proj, per device.  Atomic because two host threads may make their
// first call on a device at the same time: both look the address up and store the same value.
std::atomic<double *> g_noalloc_ptr[kMaxDevices];
}  // namespace

extern "C" qcb_status qcb_fidelity_eval_batch_device_noalloc(qcb_fidelity f, const double *d_X, int64_t count,
                                                             double *d_F, double *d_grad) {
    if (!f || !d_X || count < 1) return QCB_ERR_ARG;
    if (count * f->R > QCB_NOALLOC_MAX_PROJ || count > 65535) {
        char buf[200];
        snprintf(buf, sizeof buf, "batch too large for the fixed scratch: count * %lld projections per state must be <= %d",
                 (long long)f->R, QCB_NOALLOC_MAX_PROJ);
        f->err = buf;
        return QCB_ERR_ARG;
    }
    if (f->device < 0 || f->device >= kMaxDevices) return QCB_ERR_ARG;
    cudaError_t e = cudaSetDevice(f->device);
    if (e == cudaSuccess && !g_noalloc_ptr[f->device].load(std::memory_order_acquire)) {
        // one-time set-up on this device (this device is current): resolve the symbol's address
        double *ptr = nullptr;
        e = cudaGetSymbolAddress((void **)&ptr, g_noalloc_proj);
        if (e == cudaSuccess) g_noalloc_ptr[f->device].store(ptr, std::memory_order_release);
    }
    if (e == cudaSuccess) {
        double *p = g_noalloc_ptr[f->device].load(std::memory_order_acquire);
        const int64_t n = f->x_dim;
        dim3 g1((unsigned)f->R, (unsigned)count), g2((unsigned)((n + 255) / 256), (unsigned)count);
        // every projection slot read below is written first in this same call: nothing has to start at zero
        qcb::fid_project_batch_kernel<<<g1, 256, 0, f->stream>>>(f->dG, d_X, p, n, f->R);
        qcb::fid_grad_batch_kernel<<<g2, 256, 0, f->stream>>>(f->dG, p, d_grad, d_F, n, f->R, f->w, f->linear);
        e = cudaGetLastError();
        if (e == cudaSuccess) e = cudaStreamSynchronize(f->stream);
    }
    if (e != cudaSuccess) {
        char buf[200];
        snprintf(buf, sizeof buf, "CUDA error '%s' in qcb_fidelity_eval_batch_device_noalloc", cudaGetErrorString(e));
        f->err = buf;
        cudaGetLastError();
        return QCB_ERR_CUDA;
    }
    return QCB_OK;
}

// =========================================================================================
// Device memory + CUDA IPC helpers for the multi-GPU time-slab driver
// =========================================================================================
extern "C" {

qcb_status qcb_device_alloc(int32_t device, void **ptr, size_t bytes) {
    if (!ptr) return QCB_ERR_ARG;
    *ptr = nullptr;
    if (cudaSetDevice(device) != cudaSuccess || cudaMalloc(ptr, bytes) != cudaSuccess) { cudaGetLastError(); return QCB_ERR_NOMEM; }
    return QCB_OK;
}
void qcb_device_free(int32_t device, void *ptr) {
    if (ptr && cudaSetDevice(device) == cudaSuccess) cudaFree(ptr);
}
qcb_status qcb_ipc_export(int32_t device, void *ptr, unsigned char handle[64]) {
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handle size");
    if (!ptr || !handle) return QCB_ERR_ARG;
    cudaIpcMemHandle_t hdl;
    if (cudaSetDevice(device) != cudaSuccess || cudaIpcGetMemHandle(&hdl, ptr) != cudaSuccess) { cudaGetLastError(); return QCB_ERR_CUDA; }
    memcpy(handle, &hdl, 64);
    return QCB_OK;
}
qcb_status qcb_ipc_open(int32_t device, const unsigned char handle[64], void **ptr) {
    if (!ptr || !handle) return QCB_ERR_ARG;
    cudaIpcMemHandle_t hdl;
    memcpy(&hdl, handle, 64);
    if (cudaSetDevice(device) != cudaSuccess ||
        cudaIpcOpenMemHandle(ptr, hdl, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); *ptr = nullptr; return QCB_ERR_CUDA; }
    return QCB_OK;
}
void qcb_ipc_close(int32_t device, void *ptr) {
    if (ptr && cudaSetDevice(device) == cudaSuccess) cudaIpcCloseMemHandle(ptr);
}
qcb_status qcb_device_to_host(int32_t device, void *dst_host, const void *src_device, size_t bytes) {
    if (cudaSetDevice(device) != cudaSuccess || cudaMemcpy(dst_host, src_device, bytes, cudaMemcpyDeviceToHost) != cudaSuccess) {
        cudaGetLastError();
        return QCB_ERR_CUDA;
    }
    return QCB_OK;
}

}  // extern "C"

// one C call per Ipopt callback over the pieces above (uses the batched fidelity entry point)
#include "qcb_problem.cuh"
