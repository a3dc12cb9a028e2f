// qcb_aux.cuh -- the cheap knot-point terms that surround the bilinear integrator in every
// problem template: derivative integrators and leakage objective / constraint.
// One fused elementwise launch per callback; included at the end of qcb200.cu.
#pragma once

struct qcb_aux_s {
    int device = 0, N = 0, z_dim = 0, dt_off = 0;
    int n_deriv = 0, deriv_rows = 0;           // rows per interval = sum of dims
    int leak_x_off = 0, n_leak = 0;
    int *d_row_x = nullptr, *d_row_dx = nullptr;   // [deriv_rows] knot offsets of x[i] and dx[i] per row
    int *d_leak = nullptr;                         // [n_leak] knot offsets of the leakage entries
    std::vector<int> row_x, row_dx, leak;
    double *dZ = nullptr, *dmu = nullptr, *dQ = nullptr, *dout = nullptr;
    int *d_reg_off = nullptr;                      // [z_dim-capped rows] regulariser rows of the last call
    double *d_reg_R = nullptr;
    size_t reg_cap = 0;
    size_t out_cap = 0;
    cudaStream_t stream = nullptr;
    int64_t launches = 0;
    std::string err;
};

namespace qcb {

// delta[k][r] = x_{k+1}[r] - x_k[r] - dt_k dx_k[r];  jac[k][r][0..3] = d/d(x_k, dx_k, dt_k, x_{k+1});
// hess[k][r] = d^2 (mu.delta) / d(dx_k[r]) d(dt_k) = -mu[k][r]
__global__ void aux_derivative_kernel(const double *Z, const double *mu, int K, int z_dim, int dt_off, int rows,
                                      const int *row_x, const int *row_dx, double *delta, double *jac, double *hess) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (long long)K * rows) return;
    const int k = (int)(t / rows), r = (int)(t % rows);
    const double *zk = Z + (long long)k * z_dim, *zn = zk + z_dim;
    const double dt = zk[dt_off], dx = zk[row_dx[r]];
    if (delta) delta[t] = zn[row_x[r]] - zk[row_x[r]] - dt * dx;
    if (jac) { double *j = jac + 4 * t; j[0] = -1.0; j[1] = -dt; j[2] = -dx; j[3] = 1.0; }
    if (hess && mu) hess[t] = -mu[t];
}

// leakage: per knot t and leakage entry i:  v = x_t[idx_i]
//   obj_terms[t] = Q_t * sum_i v^2 / n   (summed on the host side of the ABI: N numbers)
//   grad[t][i] = 2 Q_t v / n ;  con[t][i] = v^2 - value ;  cjac[t][i] = 2 v ;  chess[t][i] = 2 mu[t][i]
__global__ void aux_leakage_kernel(const double *Z, const double *Q, const double *mu, int N, int z_dim, int n_leak,
                                   const int *leak, double value, double *obj_terms, double *grad, double *con,
                                   double *cjac, double *chess) {
    const int t = blockIdx.x;
    const double *z = Z + (long long)t * z_dim;
    const double q = Q ? Q[t] : 1.0;
    double part = 0.0;
    for (int i = threadIdx.x; i < n_leak; i += blockDim.x) {
        const double v = z[leak[i]];
        part = fma(v, v, part);
        const long long o = (long long)t * n_leak + i;
        if (grad) grad[o] = 2.0 * q * v / n_leak;
        if (con) con[o] = v * v - value;
        if (cjac) cjac[o] = 2.0 * v;
        if (chess && mu) chess[o] = 2.0 * mu[o];
    }
    __shared__ double red[8];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = part;
    __syncthreads();
    if (threadIdx.x == 0 && obj_terms) {
        double s = 0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) s += red[w];
        obj_terms[t] = q * s / n_leak;
    }
}

// Quadratic regularisers and the minimum-time term of the objective, one CTA per knot t:
//   J = sum_rows 1/2 R_r (dt_t v_{t,r})^2  (all N knots)  +  D sum_{t<N} dt_t
// (the form the reference's solver logs pin: tests/test_reference_fixtures.py).  Per knot:
//   obj_terms[t]; grad[t][rows + 1] = {dJ/dv_r = R dt^2 v_r per row, then dJ/ddt}; hess[t][2 rows + 1] =
//   {(v_r,v_r) = R dt^2, (v_r,dt) = 2 R dt v_r} per row, then (dt,dt) = sum_r R v_r^2.
// Compact outputs in row order (the caller scatters them: a dense gradient over Z would be
// N * z_dim doubles -- 132 MB for cfg5 -- to carry 2000 * 37 numbers).
__global__ void aux_regularizer_kernel(const double *Z, int N, int z_dim, int dt_off, int rows, const int *row_off,
                                       const double *row_R, double D, double *obj_terms, double *grad, double *hess) {
    const int t = blockIdx.x;
    const double *z = Z + (long long)t * z_dim;
    const double dt = z[dt_off];
    double part = 0.0, gdt = 0.0, hdt = 0.0;
    for (int r = threadIdx.x; r < rows; r += blockDim.x) {
        const double v = z[row_off[r]], R = row_R[r];
        part = fma(0.5 * R * dt * dt, v * v, part);
        gdt = fma(R * dt, v * v, gdt);
        hdt = fma(R, v * v, hdt);
        if (grad) grad[(long long)t * (rows + 1) + r] = R * dt * dt * v;
        if (hess) {
            double *h = hess + (long long)t * (2 * rows + 1) + 2 * r;
            h[0] = R * dt * dt; h[1] = 2.0 * R * dt * v;
        }
    }
    __shared__ double red[3][8];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        part += __shfl_xor_sync(0xffffffffu, part, o);
        gdt += __shfl_xor_sync(0xffffffffu, gdt, o);
        hdt += __shfl_xor_sync(0xffffffffu, hdt, o);
    }
    if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = part; red[1][threadIdx.x >> 5] = gdt; red[2][threadIdx.x >> 5] = hdt; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double s0 = 0, s1 = 0, s2 = 0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) { s0 += red[0][w]; s1 += red[1][w]; s2 += red[2][w]; }
        const bool timed = t < N - 1;
        if (obj_terms) obj_terms[t] = s0 + (timed ? D * dt : 0.0);
        if (grad) grad[(long long)t * (rows + 1) + rows] = s1 + (timed ? D : 0.0);
        if (hess) hess[(long long)t * (2 * rows + 1) + 2 * rows] = s2;
    }
}

// all timesteps equal: rows t = 0..N-2:  dt_t - dt_{N-1} = 0, constant Jacobian (+1, -1)
__global__ void aux_dt_equal_kernel(const double *Z, int N, int z_dim, int dt_off, double *delta, double *jac) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= N - 1) return;
    if (delta) delta[t] = Z[(long long)t * z_dim + dt_off] - Z[(long long)(N - 1) * z_dim + dt_off];
    if (jac) { jac[2 * t] = 1.0; jac[2 * t + 1] = -1.0; }
}

}  // namespace qcb

namespace {
void aux_ensure(qcb_aux_s *a, double **p, size_t n) { if (!*p) CK(cudaMalloc(p, std::max<size_t>(n, 1) * 8)); }
void aux_out(qcb_aux_s *a, size_t n) {
    if (a->out_cap < n) { if (a->dout) CK(cudaFree(a->dout)); CK(cudaMalloc(&a->dout, n * 8)); a->out_cap = n; }
}
// Host -> device copy of the knot columns [lo, hi) only (same layout on the device, so the kernels
// index Z as before).  The control / timestep columns are a few dozen doubles of a knot that may be
// thousands wide (cfg5: 38 of 16 422): a strided copy moves 0.6 MB instead of 132 MB per callback.
void aux_upload_columns(qcb_aux_s *a, const double *Z, int lo, int hi) {
    lo = std::max(0, lo); hi = std::min(a->z_dim, hi);
    if (hi <= lo) return;
    const size_t pitch = (size_t)a->z_dim * 8;
    if ((size_t)(hi - lo) * 2 > (size_t)a->z_dim)
        CK(cudaMemcpyAsync(a->dZ, Z, (size_t)a->N * pitch, cudaMemcpyHostToDevice, a->stream));
    else
        CK(cudaMemcpy2DAsync(a->dZ + lo, pitch, Z + lo, pitch, (size_t)(hi - lo) * 8, (size_t)a->N,
                             cudaMemcpyHostToDevice, a->stream));
}
}  // namespace

extern "C" {

qcb_status qcb_aux_create(qcb_aux *out, const qcb_aux_desc *d) {
    if (!out || !d) return fail(nullptr, QCB_ERR_ARG, "null argument");
    *out = nullptr;
    if (d->struct_bytes != (int32_t)sizeof(qcb_aux_desc)) return fail(nullptr, QCB_ERR_ARG, "qcb_aux_desc.struct_bytes mismatch");
    if (d->N < 2 || d->z_dim < 1 || d->dt_off < 0 || d->dt_off >= d->z_dim || d->n_deriv < 0 || d->n_leak < 0)
        return fail(nullptr, QCB_ERR_ARG, "bad dimensions in qcb_aux_desc");
    qcb_aux_s *a = new qcb_aux_s();
    a->device = d->device; a->N = d->N; a->z_dim = d->z_dim; a->dt_off = d->dt_off; a->n_deriv = d->n_deriv;
    for (int i = 0; i < d->n_deriv; ++i) {
        const int xo = d->deriv_x_off[i], dxo = d->deriv_dx_off[i], dim = d->deriv_dim[i];
        if (dim < 1 || xo < 0 || dxo < 0 || xo + dim > d->z_dim || dxo + dim > d->z_dim) {
            delete a;
            return fail(nullptr, QCB_ERR_ARG, "derivative integrator components outside the knot");
        }
        for (int r = 0; r < dim; ++r) { a->row_x.push_back(xo + r); a->row_dx.push_back(dxo + r); }
    }
    a->deriv_rows = (int)a->row_x.size();
    a->leak_x_off = d->leak_x_off; a->n_leak = d->n_leak;
    for (int i = 0; i < d->n_leak; ++i) {
        const int o = d->leak_x_off + d->leak_idx[i];
        if (d->leak_idx[i] < 0 || o >= d->z_dim) { delete a; return fail(nullptr, QCB_ERR_ARG, "leakage index outside the knot"); }
        a->leak.push_back(o);
    }
    try {
        int ndev = 0;
        if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
            cudaGetLastError();
            delete a;
            return fail(nullptr, QCB_ERR_CUDA, "no usable CUDA device: this library has no CPU fallback");
        }
        CK(cudaSetDevice(a->device));
        CK(cudaStreamCreateWithFlags(&a->stream, cudaStreamNonBlocking));
        if (a->deriv_rows) {
            CK(cudaMalloc(&a->d_row_x, a->deriv_rows * sizeof(int)));
            CK(cudaMalloc(&a->d_row_dx, a->deriv_rows * sizeof(int)));
            CK(cudaMemcpy(a->d_row_x, a->row_x.data(), a->deriv_rows * sizeof(int), cudaMemcpyHostToDevice));
            CK(cudaMemcpy(a->d_row_dx, a->row_dx.data(), a->deriv_rows * sizeof(int), cudaMemcpyHostToDevice));
        }
        if (a->n_leak) {
            CK(cudaMalloc(&a->d_leak, a->n_leak * sizeof(int)));
            CK(cudaMemcpy(a->d_leak, a->leak.data(), a->n_leak * sizeof(int), cudaMemcpyHostToDevice));
        }
        CK(cudaMalloc(&a->dZ, (size_t)a->N * a->z_dim * 8));
    } catch (const CudaFail &f) {
        qcb_status st = fail_cuda(nullptr, f);
        qcb_aux_destroy(a);
        return st;
    }
    *out = a;
    return QCB_OK;
}

void qcb_aux_destroy(qcb_aux a) {
    if (!a) return;
    cudaSetDevice(a->device);
    void *ptrs[] = {a->d_row_x, a->d_row_dx, a->d_leak, a->dZ, a->dmu, a->dQ, a->dout, a->d_reg_off, a->d_reg_R};
    for (void *p : ptrs) if (p) cudaFree(p);
    if (a->stream) cudaStreamDestroy(a->stream);
    delete a;
}

int64_t qcb_aux_derivative_dim(qcb_aux a) { return a ? (int64_t)(a->N - 1) * a->deriv_rows : 0; }
int64_t qcb_aux_leakage_dim(qcb_aux a) { return a ? (int64_t)a->N * a->n_leak : 0; }

qcb_status qcb_aux_derivative_structure(qcb_aux a, int64_t row_base, int64_t *jac_rows, int64_t *jac_cols,
                                        int64_t *hess_rows, int64_t *hess_cols) {
    if (!a) return QCB_ERR_ARG;
    const int K = a->N - 1;
    for (int k = 0; k < K; ++k)
        for (int r = 0; r < a->deriv_rows; ++r) {
            const int64_t t = (int64_t)k * a->deriv_rows + r, zb = (int64_t)k * a->z_dim;
            if (jac_rows && jac_cols) {
                const int64_t cols[4] = {zb + a->row_x[r], zb + a->row_dx[r], zb + a->dt_off, zb + a->z_dim + a->row_x[r]};
                for (int q = 0; q < 4; ++q) { jac_rows[4 * t + q] = row_base + t + 1; jac_cols[4 * t + q] = cols[q] + 1; }
            }
            if (hess_rows && hess_cols) {
                const int64_t p = zb + a->row_dx[r], q = zb + a->dt_off;
                hess_rows[t] = std::min(p, q) + 1; hess_cols[t] = std::max(p, q) + 1;
            }
        }
    return QCB_OK;
}

qcb_status qcb_aux_derivative_eval(qcb_aux a, const double *Z, const double *mu, double *delta, double *jac_vals,
                                   double *hess_vals) {
    if (!a || !Z) return QCB_ERR_ARG;
    if (hess_vals && !mu) { a->err = "Hessian values need the multipliers"; return QCB_ERR_ARG; }
    try {
        CK(cudaSetDevice(a->device));
        const size_t n = (size_t)qcb_aux_derivative_dim(a);
        if (n == 0) return QCB_OK;
        {
            int lo = a->dt_off, hi = a->dt_off + 1;
            for (int r = 0; r < a->deriv_rows; ++r) {
                lo = std::min(lo, std::min(a->row_x[r], a->row_dx[r]));
                hi = std::max(hi, std::max(a->row_x[r], a->row_dx[r]) + 1);
            }
            aux_upload_columns(a, Z, lo, hi);
        }
        if (mu) { aux_ensure(a, &a->dmu, std::max(n, (size_t)qcb_aux_leakage_dim(a))); CK(cudaMemcpyAsync(a->dmu, mu, n * 8, cudaMemcpyHostToDevice, a->stream)); }
        aux_out(a, 6 * n);
        double *dd = a->dout, *dj = a->dout + n, *dh = a->dout + 5 * n;
        qcb::aux_derivative_kernel<<<(unsigned)((n + 255) / 256), 256, 0, a->stream>>>(
            a->dZ, mu ? a->dmu : nullptr, a->N - 1, a->z_dim, a->dt_off, a->deriv_rows, a->d_row_x, a->d_row_dx,
            delta ? dd : nullptr, jac_vals ? dj : nullptr, hess_vals ? dh : nullptr);
        ++a->launches;
        CK(cudaGetLastError());
        if (delta) CK(cudaMemcpyAsync(delta, dd, n * 8, cudaMemcpyDeviceToHost, a->stream));
        if (jac_vals) CK(cudaMemcpyAsync(jac_vals, dj, 4 * n * 8, cudaMemcpyDeviceToHost, a->stream));
        if (hess_vals) CK(cudaMemcpyAsync(hess_vals, dh, n * 8, cudaMemcpyDeviceToHost, a->stream));
        CK(cudaStreamSynchronize(a->stream));
    } catch (const CudaFail &f) {
        char buf[256];
        snprintf(buf, sizeof buf, "CUDA error '%s' in %s", cudaGetErrorString(f.e), f.what);
        a->err = buf; cudaGetLastError();
        return QCB_ERR_CUDA;
    }
    return QCB_OK;
}

qcb_status qcb_aux_leakage_eval(qcb_aux a, const double *Z, const double *Qs, double value, const double *mu,
                                double *objective, double *grad_vals, double *con_vals, double *con_jac_vals,
                                double *con_hess_vals) {
    if (!a || !Z) return QCB_ERR_ARG;
    if (con_hess_vals && !mu) { a->err = "Hessian values need the multipliers"; return QCB_ERR_ARG; }
    try {
        CK(cudaSetDevice(a->device));
        const size_t n = (size_t)qcb_aux_leakage_dim(a), N = (size_t)a->N;
        if (n == 0) { if (objective) *objective = 0.0; return QCB_OK; }
        {
            int lo = a->z_dim, hi = 0;
            for (int o : a->leak) { lo = std::min(lo, o); hi = std::max(hi, o + 1); }
            aux_upload_columns(a, Z, lo, hi);
        }
        if (Qs) { aux_ensure(a, &a->dQ, N); CK(cudaMemcpyAsync(a->dQ, Qs, N * 8, cudaMemcpyHostToDevice, a->stream)); }
        if (mu) { aux_ensure(a, &a->dmu, std::max(n, (size_t)qcb_aux_derivative_dim(a))); CK(cudaMemcpyAsync(a->dmu, mu, n * 8, cudaMemcpyHostToDevice, a->stream)); }
        aux_out(a, std::max(4 * n + N, 6 * (size_t)qcb_aux_derivative_dim(a)));
        double *dobj = a->dout, *dg = dobj + N, *dc = dg + n, *dcj = dc + n, *dch = dcj + n;
        qcb::aux_leakage_kernel<<<(unsigned)N, 128, 0, a->stream>>>(a->dZ, Qs ? a->dQ : nullptr, mu ? a->dmu : nullptr, a->N,
                                                                   a->z_dim, a->n_leak, a->d_leak, value, dobj, grad_vals ? dg : nullptr,
                                                                   con_vals ? dc : nullptr, con_jac_vals ? dcj : nullptr,
                                                                   con_hess_vals ? dch : nullptr);
        ++a->launches;
        CK(cudaGetLastError());
        std::vector<double> terms(N);
        CK(cudaMemcpyAsync(terms.data(), dobj, N * 8, cudaMemcpyDeviceToHost, a->stream));
        if (grad_vals) CK(cudaMemcpyAsync(grad_vals, dg, n * 8, cudaMemcpyDeviceToHost, a->stream));
        if (con_vals) CK(cudaMemcpyAsync(con_vals, dc, n * 8, cudaMemcpyDeviceToHost, a->stream));
        if (con_jac_vals) CK(cudaMemcpyAsync(con_jac_vals, dcj, n * 8, cudaMemcpyDeviceToHost, a->stream));
        if (con_hess_vals) CK(cudaMemcpyAsync(con_hess_vals, dch, n * 8, cudaMemcpyDeviceToHost, a->stream));
        CK(cudaStreamSynchronize(a->stream));
        if (objective) { double s = 0; for (double v : terms) s += v; *objective = s; }   // N knot terms computed on the GPU
    } catch (const CudaFail &f) {
        char buf[256];
        snprintf(buf, sizeof buf, "CUDA error '%s' in %s", cudaGetErrorString(f.e), f.what);
        a->err = buf; cudaGetLastError();
        return QCB_ERR_CUDA;
    }
    return QCB_OK;
}

namespace {
// flattens (off, dim, R) terms into per-row knot offsets and weights; false if a term leaves the knot
bool aux_reg_rows(qcb_aux a, int32_t n_terms, const int32_t *off, const int32_t *dim, const double *R,
                  std::vector<int> &ro, std::vector<double> &rr) {
    size_t w = 0;
    for (int i = 0; i < n_terms; ++i) {
        if (dim[i] < 1 || off[i] < 0 || off[i] + dim[i] > a->z_dim) return false;
        for (int r = 0; r < dim[i]; ++r) { ro.push_back(off[i] + r); rr.push_back(R ? R[w] : 1.0); ++w; }
    }
    return true;
}
}  // namespace

qcb_status qcb_aux_regularizer_structure(qcb_aux a, int32_t n_terms, const int32_t *off, const int32_t *dim,
                                         int64_t *grad_cols, int64_t *hess_rows, int64_t *hess_cols) {
    if (!a || n_terms < 0 || (n_terms > 0 && (!off || !dim))) return QCB_ERR_ARG;
    std::vector<int> ro; std::vector<double> rr;
    if (!aux_reg_rows(a, n_terms, off, dim, nullptr, ro, rr)) { a->err = "regulariser component outside the knot"; return QCB_ERR_ARG; }
    const int64_t rows = (int64_t)ro.size(), per = 2 * rows + 1;
    for (int t = 0; t < a->N; ++t) {
        const int64_t zb = (int64_t)t * a->z_dim, dtc = zb + a->dt_off;
        for (int64_t r = 0; r < rows; ++r) {
            const int64_t v = zb + ro[r], e = t * per + 2 * r;
            if (grad_cols) grad_cols[t * (rows + 1) + r] = v + 1;
            if (hess_rows && hess_cols) {
                hess_rows[e] = v + 1; hess_cols[e] = v + 1;
                hess_rows[e + 1] = std::min(v, dtc) + 1; hess_cols[e + 1] = std::max(v, dtc) + 1;
            }
        }
        if (grad_cols) grad_cols[t * (rows + 1) + rows] = dtc + 1;
        if (hess_rows && hess_cols) { hess_rows[t * per + 2 * rows] = dtc + 1; hess_cols[t * per + 2 * rows] = dtc + 1; }
    }
    return QCB_OK;
}

qcb_status qcb_aux_regularizer_eval(qcb_aux a, const double *Z, int32_t n_terms, const int32_t *off, const int32_t *dim,
                                    const double *R, double min_time_D, double *objective, double *grad_vals,
                                    double *hess_vals) {
    if (!a || !Z || n_terms < 0 || (n_terms > 0 && (!off || !dim || !R))) return QCB_ERR_ARG;
    std::vector<int> ro; std::vector<double> rr;
    if (!aux_reg_rows(a, n_terms, off, dim, R, ro, rr)) { a->err = "regulariser component outside the knot"; return QCB_ERR_ARG; }
    try {
        CK(cudaSetDevice(a->device));
        const size_t rows = ro.size(), N = (size_t)a->N, zd = (size_t)a->z_dim, per = 2 * rows + 1;
        if (a->reg_cap < std::max<size_t>(rows, 1)) {
            if (a->d_reg_off) CK(cudaFree(a->d_reg_off));
            if (a->d_reg_R) CK(cudaFree(a->d_reg_R));
            a->reg_cap = std::max<size_t>(rows, 1);
            CK(cudaMalloc(&a->d_reg_off, a->reg_cap * sizeof(int)));
            CK(cudaMalloc(&a->d_reg_R, a->reg_cap * 8));
        }
        if (rows) {
            CK(cudaMemcpyAsync(a->d_reg_off, ro.data(), rows * sizeof(int), cudaMemcpyHostToDevice, a->stream));
            CK(cudaMemcpyAsync(a->d_reg_R, rr.data(), rows * 8, cudaMemcpyHostToDevice, a->stream));
        }
        {
            int lo = a->dt_off, hi = a->dt_off + 1;
            for (int o : ro) { lo = std::min(lo, o); hi = std::max(hi, o + 1); }
            aux_upload_columns(a, Z, lo, hi);
        }
        aux_out(a, std::max(N + N * (rows + 1) + N * per, std::max(4 * (size_t)qcb_aux_leakage_dim(a) + N, 6 * (size_t)qcb_aux_derivative_dim(a))));
        double *dobj = a->dout, *dg = dobj + N, *dh = dg + N * (rows + 1);
        qcb::aux_regularizer_kernel<<<(unsigned)N, 128, 0, a->stream>>>(a->dZ, a->N, a->z_dim, a->dt_off, (int)rows, a->d_reg_off,
                                                                       a->d_reg_R, min_time_D, dobj, grad_vals ? dg : nullptr,
                                                                       hess_vals ? dh : nullptr);
        ++a->launches;
        CK(cudaGetLastError());
        std::vector<double> terms(N);
        CK(cudaMemcpyAsync(terms.data(), dobj, N * 8, cudaMemcpyDeviceToHost, a->stream));
        if (grad_vals) CK(cudaMemcpyAsync(grad_vals, dg, N * (rows + 1) * 8, cudaMemcpyDeviceToHost, a->stream));
        if (hess_vals) CK(cudaMemcpyAsync(hess_vals, dh, N * per * 8, cudaMemcpyDeviceToHost, a->stream));
        CK(cudaStreamSynchronize(a->stream));
        if (objective) { double s = 0; for (double v : terms) s += v; *objective = s; }   // N knot terms computed on the GPU
    } catch (const CudaFail &f) {
        char buf[256];
        snprintf(buf, sizeof buf, "CUDA error '%s' in %s", cudaGetErrorString(f.e), f.what);
        a->err = buf; cudaGetLastError();
        return QCB_ERR_CUDA;
    }
    return QCB_OK;
}

qcb_status qcb_aux_timesteps_equal(qcb_aux a, const double *Z, int64_t row_base, double *delta, int64_t *jac_rows,
                                   int64_t *jac_cols, double *jac_vals) {
    if (!a) return QCB_ERR_ARG;
    const int K = a->N - 1;
    if (jac_rows && jac_cols)
        for (int t = 0; t < K; ++t) {
            jac_rows[2 * t] = jac_rows[2 * t + 1] = row_base + t + 1;
            jac_cols[2 * t] = (int64_t)t * a->z_dim + a->dt_off + 1;
            jac_cols[2 * t + 1] = (int64_t)K * a->z_dim + a->dt_off + 1;
        }
    if (!delta && !jac_vals) return QCB_OK;
    if (!Z) return QCB_ERR_ARG;
    try {
        CK(cudaSetDevice(a->device));
        const size_t N = (size_t)a->N;
        aux_upload_columns(a, Z, a->dt_off, a->dt_off + 1);
        aux_out(a, 3 * N);
        double *dd = a->dout, *dj = a->dout + N;
        qcb::aux_dt_equal_kernel<<<(unsigned)((K + 127) / 128), 128, 0, a->stream>>>(a->dZ, a->N, a->z_dim, a->dt_off,
                                                                                    delta ? dd : nullptr, jac_vals ? dj : nullptr);
        ++a->launches;
        CK(cudaGetLastError());
        if (delta) CK(cudaMemcpyAsync(delta, dd, (size_t)K * 8, cudaMemcpyDeviceToHost, a->stream));
        if (jac_vals) CK(cudaMemcpyAsync(jac_vals, dj, (size_t)2 * K * 8, cudaMemcpyDeviceToHost, a->stream));
        CK(cudaStreamSynchronize(a->stream));
    } catch (const CudaFail &f) {
        char buf[256];
        snprintf(buf, sizeof buf, "CUDA error '%s' in %s", cudaGetErrorString(f.e), f.what);
        a->err = buf; cudaGetLastError();
        return QCB_ERR_CUDA;
    }
    return QCB_OK;
}

}  // extern "C"
