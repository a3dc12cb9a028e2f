// qcb_global.cuh -- terminal terms that involve global variables or variational states:
//   * free-phase unitary fidelity  F(x, theta) = |tr(U_goal(theta)^dagger U)|^2 / n^2 with
//     U_goal(theta) = diag(e^{i P theta}) U_goal  (or U_goal diag(...)):
//     UnitaryFreePhaseInfidelityObjective, src/quantum_objectives.jl:203-226, and the theta-parametrised
//     FinalUnitaryFidelityConstraint, src/quantum_constraints.jl:123-144;
//   * sensitivity objective  sum_t Q_t scale^4 |tr(U_t' U_t)|^2 / n^2  on a variational state,
//     UnitarySensitivityObjective, src/quantum_objectives.jl:255-279.
// Included at the end of qcb200.cu (shares its error helpers).
#pragma once

struct qcb_free_phase_s {
    int device = 0, n = 0, n_theta = 0, side = 0;
    int64_t x_dim = 0;
    double *dWr = nullptr, *dWi = nullptr;   // [n*n] conj(U_goal[a,b]) at a + b*n
    double *dP = nullptr;                    // [n_theta][n]
    double *dx = nullptr, *dtheta = nullptr;
    double *dwork = nullptr;                 // S(2) | e_r(2n) | dS/dtheta(2 n_theta) | d2S(2 n_theta^2) | s_e(2 x_dim)
    double *dF = nullptr, *dgrad = nullptr, *dH = nullptr;
    cudaStream_t stream = nullptr;
    std::string err;
};

namespace qcb {

// One CTA.  t_r = sum over the entries of group r of conj(Ug) * U (group = row for a left phase,
// column for a right phase); y_r = e^{-i phi_r} t_r, phi = P theta; S = sum_r y_r and its theta
// derivatives dS_k = sum_r -i P_rk y_r, d2S_kl = sum_r -P_rk P_rl y_r.
__global__ void __launch_bounds__(256)
fp_reduce_kernel(const double *Wr, const double *Wi, const double *P, const double *x, const double *theta, int n,
                 int n_theta, int side, double *work) {
    extern __shared__ double sh[];          // y_r: [2n]
    double *yr = sh, *yi = sh + n;
    double *S = work, *er = work + 2, *ei = er + n, *dS = ei + n, *d2S = dS + 2 * n_theta;
    for (int r = threadIdx.x; r < n; r += blockDim.x) {
        double tr = 0.0, ti = 0.0;
        for (int q = 0; q < n; ++q) {
            const int a = side == 0 ? r : q, b = side == 0 ? q : r;
            const double wr = Wr[a + (size_t)b * n], wi = Wi[a + (size_t)b * n];
            const double ur = x[(size_t)b * 2 * n + a], ui = x[(size_t)b * 2 * n + n + a];
            tr += wr * ur - wi * ui;
            ti += wr * ui + wi * ur;
        }
        double phi = 0.0;
        for (int k = 0; k < n_theta; ++k) phi = fma(P[(size_t)k * n + r], theta[k], phi);
        double sn, cs;
        sincos(phi, &sn, &cs);
        er[r] = cs; ei[r] = -sn;                              // e^{-i phi}
        yr[r] = cs * tr + sn * ti; yi[r] = cs * ti - sn * tr;
    }
    __syncthreads();
    // the sums are tiny (n terms each): one thread per output
    const int nout = 1 + n_theta + n_theta * n_theta;
    for (int o = threadIdx.x; o < nout; o += blockDim.x) {
        double ar = 0.0, ai = 0.0;
        if (o == 0) {
            for (int r = 0; r < n; ++r) { ar += yr[r]; ai += yi[r]; }
            S[0] = ar; S[1] = ai;
        } else if (o <= n_theta) {
            const int k = o - 1;
            for (int r = 0; r < n; ++r) { const double p = P[(size_t)k * n + r]; ar += p * yi[r]; ai -= p * yr[r]; }   // -i p y
            dS[2 * k] = ar; dS[2 * k + 1] = ai;
        } else {
            const int kl = o - 1 - n_theta, k = kl / n_theta, l = kl % n_theta;
            for (int r = 0; r < n; ++r) { const double pp = P[(size_t)k * n + r] * P[(size_t)l * n + r]; ar -= pp * yr[r]; ai -= pp * yi[r]; }
            d2S[2 * kl] = ar; d2S[2 * kl + 1] = ai;
        }
    }
}

// Per state entry e: s_e = dS/dx_e (stored), grad_x; the first threads also write F and grad_theta.
__global__ void fp_grad_kernel(const double *Wr, const double *Wi, int n, int n_theta, int side, double *work,
                               double *F, double *grad) {
    const int64_t xd = 2LL * n * n;
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const double *S = work, *er = work + 2, *ei = er + n, *dS = ei + n;
    double *se = work + 2 + 2 * n + 2 * n_theta + 2 * (size_t)n_theta * n_theta;
    const double w = 2.0 / ((double)n * n), Sr = S[0], Si = S[1];
    if (e < xd) {
        const int b = (int)(e / (2 * n)), ra = (int)(e % (2 * n)), a = ra % n, im = ra / n;
        const int r = side == 0 ? a : b;
        const double wr = Wr[a + (size_t)b * n], wi = Wi[a + (size_t)b * n];
        double vr = er[r] * wr - ei[r] * wi, vi = er[r] * wi + ei[r] * wr;      // e^{-i phi_r} conj(Ug_ab)
        if (im) { const double t = vr; vr = -vi; vi = t; }                       // times i for the imaginary part of U_ab
        se[e] = vr; se[xd + e] = vi;
        if (grad) grad[e] = w * (Sr * vr + Si * vi);
    }
    if (e < n_theta && grad) grad[xd + e] = w * (Sr * dS[2 * e] + Si * dS[2 * e + 1]);
    if (e == 0 && F) *F = 0.5 * w * (Sr * Sr + Si * Si);
}

// Dense Hessian over z = [x; theta], column-major (x_dim + n_theta)^2.
__global__ void fp_hess_kernel(const double *P, int n, int n_theta, int side, const double *work, double *H) {
    const int64_t xd = 2LL * n * n, zd = xd + n_theta;
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, j = blockIdx.y;
    if (i >= zd) return;
    const double *S = work, *dS = work + 2 + 2 * n, *d2S = dS + 2 * n_theta;
    const double *se = d2S + 2 * (size_t)n_theta * n_theta;
    const double w = 2.0 / ((double)n * n), Sr = S[0], Si = S[1];
    auto group = [&](int64_t e) { const int b = (int)(e / (2 * n)), a = (int)(e % (2 * n)) % n; return side == 0 ? a : b; };
    double v;
    if (i < xd && j < xd) {
        v = se[i] * se[j] + se[xd + i] * se[xd + j];
    } else if (i >= xd && j >= xd) {
        const int k = (int)(i - xd), l = (int)(j - xd);
        v = dS[2 * k] * dS[2 * l] + dS[2 * k + 1] * dS[2 * l + 1] + Sr * d2S[2 * (k * n_theta + l)] + Si * d2S[2 * (k * n_theta + l) + 1];
    } else {
        const int64_t e = i < xd ? i : j;
        const int k = (int)((i < xd ? j : i) - xd);
        const double sr = se[e], si = se[xd + e], p = P[(size_t)k * n + group(e)];
        // Re(conj(dS_k) s_e) + Re(conj(S) * (-i p) s_e)
        v = dS[2 * k] * sr + dS[2 * k + 1] * si + p * (Sr * si - Si * sr);
    }
    H[i + j * zd] = w * v;
}

// Sensitivity: one CTA per listed knot.  s = ||x_t||^2 (= tr(U'U)); term = q s^2; grad = 4 q s x;
// Hessian = q (4 s I + 8 x x^T), q = Q_t scale^4 / n^2.
__global__ void __launch_bounds__(256)
aux_sensitivity_kernel(const double *Z, int z_dim, int x_off, int x_dim, int n, const int *times, const double *Q,
                       double scale4, double *obj_terms, double *grad, double *hess) {
    const int it = blockIdx.x;
    const double *x = Z + (long long)(times[it] - 1) * z_dim + x_off;
    double part = 0.0;
    for (int i = threadIdx.x; i < x_dim; i += blockDim.x) part = fma(x[i], x[i], part);
    __shared__ double red[8];
    __shared__ double s_sh;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = part;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) s += red[w];
        s_sh = s;
    }
    __syncthreads();
    const double s = s_sh, q = (Q ? Q[it] : 1.0) * scale4 / ((double)n * n);
    if (threadIdx.x == 0 && obj_terms) obj_terms[it] = q * s * s;
    if (grad)
        for (int i = threadIdx.x; i < x_dim; i += blockDim.x) grad[(long long)it * x_dim + i] = 4.0 * q * s * x[i];
    if (hess)
        for (long long e = threadIdx.x; e < (long long)x_dim * x_dim; e += blockDim.x) {
            const int i = (int)(e % x_dim), j = (int)(e / x_dim);
            hess[(long long)it * x_dim * x_dim + e] = q * (8.0 * x[i] * x[j] + (i == j ? 4.0 * s : 0.0));
        }
}

}  // namespace qcb

extern "C" {

qcb_status qcb_free_phase_create(qcb_free_phase *out, int32_t n, const double *goal, int32_t n_theta,
                                 const double *phase_map, int32_t side, int32_t device) {
    if (!out || !goal || n < 1 || n_theta < 0 || (n_theta > 0 && !phase_map) || (side != 0 && side != 1))
        return fail(nullptr, QCB_ERR_ARG, "null/bad argument");
    if (n > 2048) return fail(nullptr, QCB_ERR_ARG, "free-phase fidelity: n > 2048 is not supported");   // 2n doubles of shared memory
    *out = nullptr;
    qcb_free_phase_s *f = new qcb_free_phase_s();
    try {
        int ndev = 0;
        if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
            cudaGetLastError();
            delete f;
            return fail(nullptr, QCB_ERR_CUDA, "no usable CUDA device: this library has no CPU fallback");
        }
        f->device = device; f->n = n; f->n_theta = n_theta; f->side = side; f->x_dim = 2LL * n * n;
        const size_t nn = (size_t)n * n, zd = (size_t)f->x_dim + n_theta;
        std::vector<double> wr(nn), wi(nn);
        for (size_t e = 0; e < nn; ++e) { wr[e] = goal[2 * e]; wi[e] = -goal[2 * e + 1]; }    // conj(U_goal)
        CK(cudaSetDevice(device));
        CK(cudaStreamCreateWithFlags(&f->stream, cudaStreamNonBlocking));
        CK(cudaMalloc(&f->dWr, nn * 8)); CK(cudaMalloc(&f->dWi, nn * 8));
        CK(cudaMemcpy(f->dWr, wr.data(), nn * 8, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(f->dWi, wi.data(), nn * 8, cudaMemcpyHostToDevice));
        CK(cudaMalloc(&f->dP, std::max<size_t>(1, (size_t)n_theta * n) * 8));
        if (n_theta) CK(cudaMemcpy(f->dP, phase_map, (size_t)n_theta * n * 8, cudaMemcpyHostToDevice));
        CK(cudaMalloc(&f->dx, (size_t)f->x_dim * 8));
        CK(cudaMalloc(&f->dtheta, std::max(1, n_theta) * 8));
        CK(cudaMalloc(&f->dwork, (2 + 2 * (size_t)n + 2 * (size_t)n_theta + 2 * (size_t)n_theta * n_theta + 2 * (size_t)f->x_dim) * 8));
        CK(cudaMalloc(&f->dF, 8));
        CK(cudaMalloc(&f->dgrad, zd * 8));
    } catch (const CudaFail &e) {
        qcb_status st = fail_cuda(nullptr, e);
        qcb_free_phase_destroy(f);
        return st;
    }
    *out = f;
    return QCB_OK;
}

void qcb_free_phase_destroy(qcb_free_phase f) {
    if (!f) return;
    cudaSetDevice(f->device);
    void *ptrs[] = {f->dWr, f->dWi, f->dP, f->dx, f->dtheta, f->dwork, f->dF, f->dgrad, f->dH};
    for (void *p : ptrs) if (p) cudaFree(p);
    if (f->stream) cudaStreamDestroy(f->stream);
    delete f;
}

qcb_status qcb_free_phase_eval(qcb_free_phase f, const double *x, const double *theta, double *F, double *grad, double *hess) {
    if (!f || !x || (f->n_theta > 0 && !theta)) return QCB_ERR_ARG;
    try {
        CK(cudaSetDevice(f->device));
        const int n = f->n, nt = f->n_theta;
        const int64_t xd = f->x_dim, zd = xd + nt;
        CK(cudaMemcpyAsync(f->dx, x, xd * 8, cudaMemcpyHostToDevice, f->stream));
        if (nt) CK(cudaMemcpyAsync(f->dtheta, theta, (size_t)nt * 8, cudaMemcpyHostToDevice, f->stream));
        qcb::fp_reduce_kernel<<<1, 256, 2 * (size_t)n * 8, f->stream>>>(f->dWr, f->dWi, f->dP, f->dx, f->dtheta, n, nt, f->side, f->dwork);
        qcb::fp_grad_kernel<<<(unsigned)((std::max<int64_t>(xd, nt) + 255) / 256), 256, 0, f->stream>>>(f->dWr, f->dWi, n, nt, f->side, f->dwork, f->dF, f->dgrad);
        CK(cudaGetLastError());
        if (F) CK(cudaMemcpyAsync(F, f->dF, 8, cudaMemcpyDeviceToHost, f->stream));
        if (grad) CK(cudaMemcpyAsync(grad, f->dgrad, (size_t)zd * 8, cudaMemcpyDeviceToHost, f->stream));
        if (hess) {
            if (!f->dH) CK(cudaMalloc(&f->dH, (size_t)zd * zd * 8));
            dim3 grid((unsigned)((zd + 255) / 256), (unsigned)zd);
            qcb::fp_hess_kernel<<<grid, 256, 0, f->stream>>>(f->dP, n, nt, f->side, f->dwork, f->dH);
            CK(cudaGetLastError());
            CK(cudaMemcpyAsync(hess, f->dH, (size_t)zd * zd * 8, cudaMemcpyDeviceToHost, f->stream));
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

qcb_status qcb_aux_sensitivity_eval(qcb_aux a, const double *Z, int32_t x_off, int32_t x_dim, int32_t n, int32_t n_times,
                                    const int32_t *times, const double *Qs, double scale, double *objective,
                                    double *grad_vals, double *hess_vals) {
    if (!a || !Z || n_times < 0 || (n_times > 0 && !times) || n < 1) return QCB_ERR_ARG;
    if (x_off < 0 || x_dim < 1 || x_off + x_dim > a->z_dim) { a->err = "sensitivity component outside the knot"; return QCB_ERR_ARG; }
    for (int i = 0; i < n_times; ++i)
        if (times[i] < 1 || times[i] > a->N) { a->err = "sensitivity knot index outside 1..N"; return QCB_ERR_ARG; }
    if (n_times == 0) { if (objective) *objective = 0.0; return QCB_OK; }
    int *d_times = nullptr;
    try {
        CK(cudaSetDevice(a->device));
        const size_t T = (size_t)n_times, xd = (size_t)x_dim;
        aux_upload_columns(a, Z, x_off, x_off + x_dim);
        const size_t need = T + T + T * xd + (hess_vals ? T * xd * xd : 0);
        aux_out(a, need);
        double *dobj = a->dout, *dq = dobj + T, *dg = dq + T, *dh = dg + T * xd;
        CK(cudaMalloc(&d_times, T * sizeof(int)));
        CK(cudaMemcpyAsync(d_times, times, T * sizeof(int), cudaMemcpyHostToDevice, a->stream));
        if (Qs) CK(cudaMemcpyAsync(dq, Qs, T * 8, cudaMemcpyHostToDevice, a->stream));
        const double s2 = scale * scale;
        qcb::aux_sensitivity_kernel<<<(unsigned)T, 256, 0, a->stream>>>(a->dZ, a->z_dim, x_off, x_dim, n, d_times, Qs ? dq : nullptr,
                                                                       s2 * s2, dobj, grad_vals ? dg : nullptr, hess_vals ? dh : nullptr);
        ++a->launches;
        CK(cudaGetLastError());
        std::vector<double> terms(T);
        CK(cudaMemcpyAsync(terms.data(), dobj, T * 8, cudaMemcpyDeviceToHost, a->stream));
        if (grad_vals) CK(cudaMemcpyAsync(grad_vals, dg, T * xd * 8, cudaMemcpyDeviceToHost, a->stream));
        if (hess_vals) CK(cudaMemcpyAsync(hess_vals, dh, T * xd * xd * 8, cudaMemcpyDeviceToHost, a->stream));
        CK(cudaStreamSynchronize(a->stream));
        cudaFree(d_times);
        if (objective) { double s = 0; for (double v : terms) s += v; *objective = s; }
    } catch (const CudaFail &f) {
        if (d_times) cudaFree(d_times);
        char buf[256];
        snprintf(buf, sizeof buf, "CUDA error '%s' in %s", cudaGetErrorString(f.e), f.what);
        a->err = buf; cudaGetLastError();
        return QCB_ERR_CUDA;
    }
    return QCB_OK;
}

}  // extern "C"
