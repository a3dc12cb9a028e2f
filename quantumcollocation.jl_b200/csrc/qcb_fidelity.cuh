// qcb_fidelity.cuh -- terminal fidelities of src/quantum_objectives.jl as quadratic forms
//   F(x) = w * sum_r (g_r . x)^2,  grad F = 2w G^T (G x),  Hess F = 2w G^T G  (constant).
// Included at the end of qcb200.cu (shares its error helpers).
#pragma once

struct qcb_fidelity_s {
    int device = 0;
    int64_t x_dim = 0, R = 0;
    double w = 1.0;
    int linear = 0;           // density kind: F = w * (g_0 . x), not a sum of squares
    double *dG = nullptr;     // [R][x_dim] row-major
    double *dx = nullptr, *dp = nullptr, *dgrad = nullptr, *dF = nullptr, *dH = nullptr;
    cudaStream_t stream = nullptr;
    std::string err;
};

namespace qcb {

__global__ void __launch_bounds__(256) fid_project_kernel(const double *G, const double *x, double *p, int64_t n) {
    const double *g = G + (int64_t)blockIdx.x * n;
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
        p[blockIdx.x] = s;
    }
}
__global__ void fid_grad_kernel(const double *G, const double *p, double *grad, double *F, int64_t n, int64_t R, double w,
                                int linear) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (linear) {       // F = w g.x: constant gradient, no curvature
        if (i < n) grad[i] = w * G[i];
        if (i == 0) *F = w * p[0];
        return;
    }
    if (i < n) {
        double v = 0.0;
        for (int64_t r = 0; r < R; ++r) v = fma(G[r * n + i], p[r], v);
        grad[i] = 2.0 * w * v;
    }
    if (i == 0) {
        double f = 0.0;
        for (int64_t r = 0; r < R; ++r) f = fma(p[r], p[r], f);
        *F = w * f;
    }
}
__global__ void fid_hess_kernel(const double *G, double *H, int64_t n, int64_t R, double w, int linear) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, j = blockIdx.y;
    if (i >= n) return;
    double v = 0.0;
    for (int64_t r = 0; r < R && !linear; ++r) v = fma(G[r * n + i], G[r * n + j], v);
    H[i + j * n] = 2.0 * w * v;
}

}  // namespace qcb

namespace {

// isovec of a complex n x n operator given as interleaved (re,im) column-major, times phase
// (pr + i pi), restricted/embedded through `rows`/`cols` index maps into an n_full operator.
void add_isovec(std::vector<double> &g, int n_full, const std::vector<int> &rows, const std::vector<int> &cols,
                const double *M_re_im, int ldm, double pr, double pi) {
    for (size_t b = 0; b < cols.size(); ++b)
        for (size_t a = 0; a < rows.size(); ++a) {
            const double re = M_re_im[2 * (a + b * ldm)], im = M_re_im[2 * (a + b * ldm) + 1];
            const double vr = pr * re - pi * im, vi = pr * im + pi * re;
            g[(size_t)cols[b] * 2 * n_full + rows[a]] += vr;
            g[(size_t)cols[b] * 2 * n_full + n_full + rows[a]] += vi;
        }
}

}  // namespace

extern "C" {

qcb_status qcb_fidelity_create(qcb_fidelity *out, int32_t kind, int32_t n, int32_t count, const double *goal,
                               const int32_t *subspace, int32_t n_sub, int32_t device) {
    if (!out || !goal || n < 1) return fail(nullptr, QCB_ERR_ARG, "null/bad argument");
    *out = nullptr;
    std::vector<double> G;
    int64_t xd = 0, R = 2;
    double w = 1.0;
    std::vector<int> all(n);
    for (int i = 0; i < n; ++i) all[i] = i;
    if (kind == QCB_FID_UNITARY) {
        xd = 2LL * n * n; w = 1.0 / ((double)n * n);
        G.assign(2 * xd, 0.0);
        std::vector<double> ga(xd, 0.0), gb(xd, 0.0);
        add_isovec(ga, n, all, all, goal, n, 1.0, 0.0);
        add_isovec(gb, n, all, all, goal, n, 0.0, 1.0);
        std::copy(ga.begin(), ga.end(), G.begin());
        std::copy(gb.begin(), gb.end(), G.begin() + xd);
    } else if (kind == QCB_FID_KET || kind == QCB_FID_COHERENT_KET) {
        if (kind == QCB_FID_KET) count = 1;
        if (count < 1) return fail(nullptr, QCB_ERR_ARG, "count < 1");
        xd = 2LL * n * count; w = 1.0 / ((double)count * count);
        G.assign(2 * xd, 0.0);
        for (int q = 0; q < count; ++q)
            for (int a = 0; a < n; ++a) {
                const double re = goal[2 * (a + (size_t)q * n)], im = goal[2 * (a + (size_t)q * n) + 1];
                G[(size_t)q * 2 * n + a] = re;            G[(size_t)q * 2 * n + n + a] = im;        // psi
                G[xd + (size_t)q * 2 * n + a] = -im;      G[xd + (size_t)q * 2 * n + n + a] = re;   // i*psi
            }
    } else if (kind == QCB_FID_UNITARY_SUBSPACE) {
        if (!subspace || n_sub < 1 || n_sub > n) return fail(nullptr, QCB_ERR_ARG, "bad subspace");
        xd = 2LL * n * n; w = 1.0 / ((double)n_sub * (n_sub + 1));
        R = 2 + 2LL * n_sub * n_sub;
        G.assign(R * xd, 0.0);
        std::vector<int> sub(subspace, subspace + n_sub);
        for (int v : sub)
            if (v < 0 || v >= n) return fail(nullptr, QCB_ERR_ARG, "subspace index out of range");
        std::vector<double> row(xd);
        int64_t r = 0;
        for (int ph = 0; ph < 2; ++ph) {   // tr(M): Re and Im
            std::fill(row.begin(), row.end(), 0.0);
            add_isovec(row, n, sub, sub, goal, n_sub, ph == 0 ? 1.0 : 0.0, ph == 0 ? 0.0 : 1.0);
            std::copy(row.begin(), row.end(), G.begin() + r * xd); ++r;
        }
        // M[a,b] = sum_r conj(Ug[r,a]) U_sub[r,b]: linear forms with V[:,b] = Ug[:,a]
        for (int a = 0; a < n_sub; ++a)
            for (int b = 0; b < n_sub; ++b)
                for (int ph = 0; ph < 2; ++ph) {
                    std::fill(row.begin(), row.end(), 0.0);
                    std::vector<int> colb{sub[b]};
                    add_isovec(row, n, sub, colb, goal + 2 * (size_t)a * n_sub, n_sub, ph == 0 ? 1.0 : 0.0, ph == 0 ? 0.0 : 1.0);
                    std::copy(row.begin(), row.end(), G.begin() + r * xd); ++r;
                }
    } else if (kind == QCB_FID_DENSITY_PURE) {
        // F = Re(psi^dagger rho psi) on rho_iso = [vec(Re rho); vec(Im rho)] (column-major vec):
        //   sum_ij Re(c_ij) Re(rho_ij) - Im(c_ij) Im(rho_ij),  c_ij = conj(psi_i) psi_j
        xd = 2LL * n * n; w = 1.0; R = 1;
        G.assign(xd, 0.0);
        for (int j = 0; j < n; ++j)
            for (int i = 0; i < n; ++i) {
                const double ar = goal[2 * i], ai = goal[2 * i + 1], br = goal[2 * j], bi = goal[2 * j + 1];
                G[(size_t)i + (size_t)j * n] = ar * br + ai * bi;                           // Re(conj(a) b)
                G[(size_t)n * n + i + (size_t)j * n] = -(ar * bi - ai * br);                // -Im(conj(a) b)
            }
    } else {
        return fail(nullptr, QCB_ERR_ARG, "unknown fidelity kind");
    }
    qcb_fidelity_s *f = new qcb_fidelity_s();
    try {
        int ndev = 0;
        if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
            cudaGetLastError();
            delete f;
            return fail(nullptr, QCB_ERR_CUDA, "no usable CUDA device: this library has no CPU fallback");
        }
        f->device = device; f->x_dim = xd; f->R = R; f->w = w; f->linear = kind == QCB_FID_DENSITY_PURE;
        CK(cudaSetDevice(device));
        CK(cudaStreamCreateWithFlags(&f->stream, cudaStreamNonBlocking));
        CK(cudaMalloc(&f->dG, G.size() * 8));
        CK(cudaMemcpy(f->dG, G.data(), G.size() * 8, cudaMemcpyHostToDevice));
        CK(cudaMalloc(&f->dx, xd * 8));
        CK(cudaMalloc(&f->dp, R * 8));
        CK(cudaMalloc(&f->dgrad, xd * 8));
        CK(cudaMalloc(&f->dF, 8));
    } catch (const CudaFail &e) {
        qcb_status st = fail_cuda(nullptr, e);
        qcb_fidelity_destroy(f);
        return st;
    }
    *out = f;
    return QCB_OK;
}

void qcb_fidelity_destroy(qcb_fidelity f) {
    if (!f) return;
    cudaSetDevice(f->device);
    void *ptrs[] = {f->dG, f->dx, f->dp, f->dgrad, f->dF, f->dH};
    for (void *p : ptrs)
        if (p) cudaFree(p);
    if (f->stream) cudaStreamDestroy(f->stream);
    delete f;
}

int64_t qcb_fidelity_x_dim(qcb_fidelity f) { return f ? f->x_dim : 0; }

int64_t qcb_fidelity_rank(qcb_fidelity f) { return f ? (f->linear ? 0 : f->R) : 0; }

qcb_status qcb_fidelity_factor(qcb_fidelity f, double *G, double *w) {
    if (!f || !G || !w) return QCB_ERR_ARG;
    if (cudaSetDevice(f->device) != cudaSuccess ||
        cudaMemcpy(G, f->dG, (size_t)f->R * f->x_dim * 8, cudaMemcpyDeviceToHost) != cudaSuccess) {
        cudaGetLastError();
        return QCB_ERR_CUDA;
    }
    *w = f->w;
    return QCB_OK;
}

qcb_status qcb_fidelity_eval(qcb_fidelity f, const double *x, double *F, double *grad, double *hess) {
    if (!f || !x) return QCB_ERR_ARG;
    try {
        CK(cudaSetDevice(f->device));
        const int64_t n = f->x_dim;
        CK(cudaMemcpyAsync(f->dx, x, n * 8, cudaMemcpyHostToDevice, f->stream));
        qcb::fid_project_kernel<<<(unsigned)f->R, 256, 0, f->stream>>>(f->dG, f->dx, f->dp, n);
        qcb::fid_grad_kernel<<<(unsigned)((n + 255) / 256), 256, 0, f->stream>>>(f->dG, f->dp, f->dgrad, f->dF, n, f->R, f->w, f->linear);
        CK(cudaGetLastError());
        if (F) CK(cudaMemcpyAsync(F, f->dF, 8, cudaMemcpyDeviceToHost, f->stream));
        if (grad) CK(cudaMemcpyAsync(grad, f->dgrad, n * 8, cudaMemcpyDeviceToHost, f->stream));
        if (hess) {
            if (!f->dH) {
                CK(cudaMalloc(&f->dH, (size_t)n * n * 8));
                dim3 grid((unsigned)((n + 255) / 256), (unsigned)n);
                qcb::fid_hess_kernel<<<grid, 256, 0, f->stream>>>(f->dG, f->dH, n, f->R, f->w, f->linear);
                CK(cudaGetLastError());
            }
            CK(cudaMemcpyAsync(hess, f->dH, (size_t)n * n * 8, cudaMemcpyDeviceToHost, f->stream));
        }
        CK(cudaStreamSynchronize(f->stream));
    } catch (const CudaFail &e) {
        char buf[256];
        snprintf(buf, sizeof buf, "CUDA error '%s' in %s", cudaGetErrorString(e.e), e.what);
        f->err = buf;
        cudaGetLastError();
        return QCB_ERR_CUDA;
    }
    return QCB_OK;
}

}  // extern "C"
