// qcb_problem.cuh -- one C call per Ipopt callback (second half).
// The reference hands Ipopt five callbacks built by DirectTrajOpt's MOI evaluator over the lists a
// problem template assembles (src/problem_templates/smooth_pulse_problem.jl:139-165: objective =
// terminal infidelity + QuadraticRegularizers, integrators = [BilinearIntegrator,
// DerivativeIntegrator(u, du), DerivativeIntegrator(du, ddu)]; _problem_templates.jl:38-67 adds the
// timesteps-all-equal rows; the solve is src/quantum_control_problem.jl:163-164).  A qcb_problem is
// that list on this side of the ABI: it borrows integrator handles, one qcb_aux and one
// qcb_fidelity, fixes the row / entry offsets of every piece inside Ipopt's vectors once, and
// serves eval_f, eval_grad_f, eval_g, eval_jac_g and eval_h each as one call that writes the
// caller's full-length arrays.  All arithmetic is done by the pieces' own GPU kernels; this file
// is offsets and scatter.  Included at the end of qcb200.cu.
#pragma once
#include <memory>

struct qcb_problem_s {
    int N = 0, z_dim = 0;
    std::vector<qcb_handle> integ;
    qcb_aux aux = nullptr;
    bool deriv_rows = false, dt_rows = false;
    qcb_fidelity fid = nullptr;
    std::vector<int64_t> fid_pos;            // [fid_states][x_dim] 0-based positions in Z of the terminal states' entries
    int fid_states = 0;                      // terminal states the fidelity is summed over (systems of a sampling problem)
    std::vector<double> fid_w, Fv;           // weights and values per state
    double *dX = nullptr, *dF = nullptr, *dGr = nullptr;      // device buffers of the batched fidelity call
    double Q = 0.0;
    bool fid_hess = false;
    std::vector<int32_t> reg_off, reg_dim;
    std::vector<double> reg_R;
    double min_time_D = 0.0;
    bool has_reg = false;
    // offsets inside Ipopt's vectors
    int64_t n = 0, m = 0, nnz_j = 0, nnz_h = 0;
    std::vector<int64_t> row0, jac0, hes0;   // per integrator
    int64_t d_row0 = 0, d_jac0 = 0, d_hes0 = 0, d_rows = 0;
    int64_t t_row0 = 0, t_jac0 = 0, t_rows = 0;
    int64_t f_hes0 = 0, f_hes_n = 0, r_hes0 = 0, r_hes_n = 0, r_rows = 0;
    // leakage (qcb_aux_leakage_eval): inequality rows x^2 - value <= 0 and / or the objective term leak_Q sum x^2 / n_leak
    bool leak_rows = false, leak_obj = false;
    double leak_Q = 0.0, leak_value = 0.0;
    int64_t l_n = 0, l_row0 = 0, l_jac0 = 0, l_hes0 = 0, lo_hes0 = 0;
    std::vector<double> lbuf;                // [3][l_n]: gradient values | constraint values or Jacobian | Hessian diagonal
    std::vector<int64_t> reg_cols;           // 1-based columns of the compact regulariser gradient
    std::vector<char> stale;                 // integrator i has not seen the current Z yet
    std::vector<double> xN, gF, HF, rg, rh;
    std::string err;
};

namespace {
qcb_status pfail(qcb_problem_s *p, qcb_status st, const std::string &msg) {
    if (p) p->err = msg; else g_create_error = msg;
    return st;
}
// status of a piece -> status of the problem, with the piece's message
qcb_status pstep(qcb_problem_s *p, qcb_status st, const char *what, const std::string &detail) {
    if (st == QCB_OK) return st;
    p->err = std::string(what) + ": " + detail;
    return st;
}
// F_s (p->Fv) and, on request, dF_s/dx (p->gF, [state][x_dim]) of every terminal state: one upload of the states, the
// batched fidelity kernels, one read-back.  The Hessian of these losses is constant in x (p->HF, taken at create).
qcb_status prob_fidelity(qcb_problem_s *p, const double *Z, bool grad) {
    const size_t xd = (size_t)qcb_fidelity_x_dim(p->fid), tot = xd * p->fid_states;
    for (size_t i = 0; i < tot; ++i) p->xN[i] = Z[p->fid_pos[i]];
    cudaError_t e = cudaSetDevice(p->fid->device);
    if (e == cudaSuccess) e = cudaMemcpy(p->dX, p->xN.data(), tot * 8, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        qcb_status st = qcb_fidelity_eval_batch_device_noalloc(p->fid, p->dX, p->fid_states, p->dF, grad ? p->dGr : nullptr);
        if (st != QCB_OK) { p->err = "terminal fidelity: " + p->fid->err; return st; }
        e = cudaMemcpy(p->Fv.data(), p->dF, (size_t)p->fid_states * 8, cudaMemcpyDeviceToHost);
    }
    if (e == cudaSuccess && grad) {
        p->gF.resize(tot);
        e = cudaMemcpy(p->gF.data(), p->dGr, tot * 8, cudaMemcpyDeviceToHost);
    }
    if (e != cudaSuccess) {
        p->err = std::string("terminal fidelity: CUDA error '") + cudaGetErrorString(e) + "'";
        cudaGetLastError();
        return QCB_ERR_CUDA;
    }
    return QCB_OK;
}
// d/dF of w_s Q |1 - F_s|
inline double prob_fid_slope(const qcb_problem_s *p, int s) { return ((1.0 - p->Fv[(size_t)s]) >= 0.0 ? -p->Q : p->Q) * p->fid_w[(size_t)s]; }
}  // namespace

extern "C" {

qcb_status qcb_problem_create(qcb_problem *out, const qcb_problem_desc *d) {
    if (!out || !d) return pfail(nullptr, QCB_ERR_ARG, "null argument");
    *out = nullptr;
    if (d->struct_bytes != (int32_t)sizeof(qcb_problem_desc)) return pfail(nullptr, QCB_ERR_ARG, "qcb_problem_desc.struct_bytes mismatch");
    if (d->N < 2 || d->z_dim < 1 || d->n_integrators < 0 || d->n_reg < 0 || (d->n_integrators > 0 && !d->integrators))
        return pfail(nullptr, QCB_ERR_ARG, "bad dimensions in qcb_problem_desc");
    // every early return below releases what was built so far, device buffers included
    std::unique_ptr<qcb_problem_s, void (*)(qcb_problem_s *)> p(new qcb_problem_s(), qcb_problem_destroy);
    p->N = d->N; p->z_dim = d->z_dim;
    p->n = (int64_t)d->N * d->z_dim;
    for (int i = 0; i < d->n_integrators; ++i) {
        qcb_handle h = d->integrators[i];
        if (!h) return pfail(nullptr, QCB_ERR_ARG, "null integrator handle");
        if (h->N != d->N || h->z_dim != d->z_dim) return pfail(nullptr, QCB_ERR_ARG, "integrator built for another trajectory shape");
        if (h->k0 != 0 || h->k1 != h->K) return pfail(nullptr, QCB_ERR_ARG, "a problem takes integrators that own the whole trajectory (k_begin = 0, k_end = K)");
        p->integ.push_back(h);
        p->row0.push_back(p->m); p->jac0.push_back(p->nnz_j); p->hes0.push_back(p->nnz_h);
        p->m += qcb_defect_dim(h); p->nnz_j += qcb_jacobian_nnz(h); p->nnz_h += qcb_hessian_nnz(h);
    }
    p->stale.assign(p->integ.size(), 1);
    const bool want_aux = d->use_derivative_rows || d->use_dt_equal_rows || d->n_reg > 0 || d->min_time_D != 0.0 ||
                          d->use_leakage_rows || d->use_leakage_objective;
    if (want_aux) {
        if (!d->aux) return pfail(nullptr, QCB_ERR_ARG, "derivative rows, timestep rows and regularisers need a qcb_aux");
        if (d->aux->N != d->N || d->aux->z_dim != d->z_dim) return pfail(nullptr, QCB_ERR_ARG, "qcb_aux built for another trajectory shape");
        p->aux = d->aux;
    }
    if (d->use_derivative_rows) {
        p->deriv_rows = true;
        p->d_rows = qcb_aux_derivative_dim(p->aux);
        p->d_row0 = p->m; p->d_jac0 = p->nnz_j; p->d_hes0 = p->nnz_h;
        p->m += p->d_rows; p->nnz_j += 4 * p->d_rows; p->nnz_h += p->d_rows;
    }
    if (d->use_dt_equal_rows) {
        p->dt_rows = true;
        p->t_rows = d->N - 1;
        p->t_row0 = p->m; p->t_jac0 = p->nnz_j;
        p->m += p->t_rows; p->nnz_j += 2 * p->t_rows;
    }
    if (d->use_leakage_rows || d->use_leakage_objective) {
        if (!d->aux || qcb_aux_leakage_dim(d->aux) == 0) return pfail(nullptr, QCB_ERR_ARG, "leakage terms need a qcb_aux created with leakage indices");
        p->aux = d->aux;
        p->l_n = qcb_aux_leakage_dim(p->aux);
        p->lbuf.resize((size_t)3 * p->l_n);
        if (d->use_leakage_rows) {
            p->leak_rows = true; p->leak_value = d->leak_value;
            p->l_row0 = p->m; p->l_jac0 = p->nnz_j; p->l_hes0 = p->nnz_h;
            p->m += p->l_n; p->nnz_j += p->l_n; p->nnz_h += p->l_n;
        }
        if (d->use_leakage_objective) {
            p->leak_obj = true; p->leak_Q = d->leak_Q;
            p->lo_hes0 = p->nnz_h; p->nnz_h += p->l_n;
        }
    }
    if (d->fidelity) {
        p->fid = d->fidelity;
        const int64_t xd = qcb_fidelity_x_dim(p->fid), base = (int64_t)(d->N - 1) * d->z_dim;
        p->fid_states = d->n_fid_states > 0 ? d->n_fid_states : 1;
        if (!d->fid_index) return pfail(nullptr, QCB_ERR_ARG, "fid_index is required with a fidelity");
        if ((int64_t)p->fid_states * p->fid->R > QCB_NOALLOC_MAX_PROJ || p->fid_states > 65535)
            return pfail(nullptr, QCB_ERR_ARG, "too many terminal states for the batched fidelity (states * projections <= 65536)");
        for (int64_t i = 0; i < xd * p->fid_states; ++i) {
            if (d->fid_index[i] < 0 || d->fid_index[i] >= d->z_dim) return pfail(nullptr, QCB_ERR_ARG, "fid_index outside the knot");
            p->fid_pos.push_back(base + d->fid_index[i]);
        }
        for (int s = 0; s < p->fid_states; ++s) p->fid_w.push_back(d->fid_weights ? d->fid_weights[s] : 1.0);
        p->xN.resize((size_t)(xd * p->fid_states));
        p->Fv.assign((size_t)p->fid_states, 0.0);
        p->Q = d->Q;
        p->fid_hess = d->fid_hessian != 0 && !p->fid->linear;
        if (p->fid_hess) {
            if (xd > 8192) return pfail(nullptr, QCB_ERR_ARG, "dense terminal Hessian block over more than 8192 state entries: set fid_hessian = 0 (qcb_fidelity_factor gives the factored form)");
            p->f_hes0 = p->nnz_h; p->f_hes_n = xd * (xd + 1) / 2 * p->fid_states;
            p->nnz_h += p->f_hes_n;
            // quadratic forms: the Hessian does not depend on x
            p->HF.resize((size_t)(xd * xd));
            std::vector<double> zero((size_t)xd, 0.0);
            if (qcb_fidelity_eval(p->fid, zero.data(), nullptr, nullptr, p->HF.data()) != QCB_OK)
                return pfail(nullptr, QCB_ERR_CUDA, "terminal fidelity Hessian: " + p->fid->err);
        }
        const size_t tot = (size_t)(xd * p->fid_states);
        if (cudaSetDevice(p->fid->device) != cudaSuccess || cudaMalloc(&p->dX, tot * 8) != cudaSuccess ||
            cudaMalloc(&p->dGr, tot * 8) != cudaSuccess || cudaMalloc(&p->dF, (size_t)p->fid_states * 8) != cudaSuccess) {
            cudaGetLastError();
            return pfail(nullptr, QCB_ERR_NOMEM, "device buffers of the terminal fidelity");
        }
    }
    if (d->n_reg > 0 || d->min_time_D != 0.0) {
        if (d->n_reg > 0 && (!d->reg_off || !d->reg_dim || !d->reg_R)) return pfail(nullptr, QCB_ERR_ARG, "regulariser arrays missing");
        p->has_reg = true;
        p->min_time_D = d->min_time_D;
        for (int i = 0; i < d->n_reg; ++i) {
            if (d->reg_dim[i] < 1 || d->reg_off[i] < 0 || d->reg_off[i] + d->reg_dim[i] > d->z_dim)
                return pfail(nullptr, QCB_ERR_ARG, "regulariser component outside the knot");
            p->reg_off.push_back(d->reg_off[i]); p->reg_dim.push_back(d->reg_dim[i]);
            for (int r = 0; r < d->reg_dim[i]; ++r) p->reg_R.push_back(d->reg_R[p->r_rows + r]);
            p->r_rows += d->reg_dim[i];
        }
        p->r_hes0 = p->nnz_h; p->r_hes_n = (int64_t)d->N * (2 * p->r_rows + 1);
        p->nnz_h += p->r_hes_n;
        p->reg_cols.resize((size_t)d->N * (p->r_rows + 1));
        if (qcb_aux_regularizer_structure(p->aux, (int32_t)p->reg_off.size(), p->reg_off.data(), p->reg_dim.data(),
                                          p->reg_cols.data(), nullptr, nullptr) != QCB_OK)
            return pfail(nullptr, QCB_ERR_ARG, "regulariser structure: " + p->aux->err);
        p->rg.resize(p->reg_cols.size());
        p->rh.resize((size_t)p->r_hes_n);
    }
    *out = p.release();
    return QCB_OK;
}

void qcb_problem_destroy(qcb_problem p) {      // the pieces are borrowed: the caller destroys them (after the problem)
    if (!p) return;
    if (p->dX || p->dF || p->dGr) {
        if (p->fid) cudaSetDevice(p->fid->device);
        for (double *q : {p->dX, p->dF, p->dGr}) if (q) cudaFree(q);
    }
    delete p;
}
const char *qcb_problem_last_error(qcb_problem p) { return p ? p->err.c_str() : g_create_error.c_str(); }

qcb_status qcb_problem_dims(qcb_problem p, int64_t *n, int64_t *m, int64_t *nnz_jac, int64_t *nnz_hess) {
    if (!p) return QCB_ERR_ARG;
    if (n) *n = p->n;
    if (m) *m = p->m;
    if (nnz_jac) *nnz_jac = p->nnz_j;
    if (nnz_hess) *nnz_hess = p->nnz_h;
    return QCB_OK;
}

// 1-based COO structure of the constraint Jacobian (rows over the m constraint rows, columns over Z) and of the
// Hessian of the Lagrangian (upper triangle; duplicates are summed by the solver), in the order of the values.
qcb_status qcb_problem_jacobian_structure(qcb_problem p, int64_t *rows, int64_t *cols) {
    if (!p || !rows || !cols) return QCB_ERR_ARG;
    for (size_t i = 0; i < p->integ.size(); ++i) {
        qcb_status st = qcb_jacobian_structure(p->integ[i], p->row0[i], rows + p->jac0[i], cols + p->jac0[i]);
        if (st != QCB_OK) return pstep(p, st, "integrator Jacobian structure", p->integ[i]->err);
    }
    if (p->deriv_rows) {
        qcb_status st = qcb_aux_derivative_structure(p->aux, p->d_row0, rows + p->d_jac0, cols + p->d_jac0, nullptr, nullptr);
        if (st != QCB_OK) return pstep(p, st, "derivative rows", p->aux->err);
    }
    if (p->dt_rows) {
        qcb_status st = qcb_aux_timesteps_equal(p->aux, nullptr, p->t_row0, nullptr, rows + p->t_jac0, cols + p->t_jac0, nullptr);
        if (st != QCB_OK) return pstep(p, st, "timestep rows", p->aux->err);
    }
    if (p->leak_rows) {
        const int nl = p->aux->n_leak;
        for (int64_t e = 0; e < p->l_n; ++e) {
            rows[p->l_jac0 + e] = p->l_row0 + e + 1;
            cols[p->l_jac0 + e] = (e / nl) * p->z_dim + p->aux->leak[(size_t)(e % nl)] + 1;
        }
    }
    return QCB_OK;
}

qcb_status qcb_problem_hessian_structure(qcb_problem p, int64_t *rows, int64_t *cols) {
    if (!p || !rows || !cols) return QCB_ERR_ARG;
    for (size_t i = 0; i < p->integ.size(); ++i) {
        qcb_status st = qcb_hessian_structure(p->integ[i], rows + p->hes0[i], cols + p->hes0[i]);
        if (st != QCB_OK) return pstep(p, st, "integrator Hessian structure", p->integ[i]->err);
    }
    if (p->deriv_rows) {
        qcb_status st = qcb_aux_derivative_structure(p->aux, 0, nullptr, nullptr, rows + p->d_hes0, cols + p->d_hes0);
        if (st != QCB_OK) return pstep(p, st, "derivative rows", p->aux->err);
    }
    if (p->leak_rows || p->leak_obj) {
        const int nl = p->aux->n_leak;
        for (int64_t e = 0; e < p->l_n; ++e) {
            const int64_t c = (e / nl) * p->z_dim + p->aux->leak[(size_t)(e % nl)] + 1;
            if (p->leak_rows) { rows[p->l_hes0 + e] = c; cols[p->l_hes0 + e] = c; }
            if (p->leak_obj) { rows[p->lo_hes0 + e] = c; cols[p->lo_hes0 + e] = c; }
        }
    }
    if (p->fid_hess) {
        const int64_t xd = qcb_fidelity_x_dim(p->fid);
        int64_t e = p->f_hes0;
        for (int s = 0; s < p->fid_states; ++s) {
            const int64_t *pos = p->fid_pos.data() + (size_t)s * xd;
            for (int64_t b = 0; b < xd; ++b)
                for (int64_t a = 0; a <= b; ++a, ++e) {
                    rows[e] = std::min(pos[a], pos[b]) + 1;
                    cols[e] = std::max(pos[a], pos[b]) + 1;
                }
        }
    }
    if (p->has_reg) {
        qcb_status st = qcb_aux_regularizer_structure(p->aux, (int32_t)p->reg_off.size(), p->reg_off.data(), p->reg_dim.data(),
                                                      nullptr, rows + p->r_hes0, cols + p->r_hes0);
        if (st != QCB_OK) return pstep(p, st, "regulariser structure", p->aux->err);
    }
    return QCB_OK;
}

// eval_f: Q |1 - F(x_N)| + regularisers + D sum dt.  new_x as Ipopt passes it (nothing of the objective is cached;
// the flag is recorded so that the next constraint callback uploads Z).
qcb_status qcb_problem_eval_f(qcb_problem p, const double *Z, int32_t new_x, double *f) {
    if (!p || !Z || !f) return QCB_ERR_ARG;
    if (new_x) std::fill(p->stale.begin(), p->stale.end(), 1);
    double v = 0.0;
    if (p->fid) {
        qcb_status st = prob_fidelity(p, Z, false);
        if (st != QCB_OK) return st;
        for (int s = 0; s < p->fid_states; ++s) v += p->fid_w[(size_t)s] * p->Q * std::fabs(1.0 - p->Fv[(size_t)s]);
    }
    if (p->has_reg) {
        double J = 0.0;
        qcb_status st = qcb_aux_regularizer_eval(p->aux, Z, (int32_t)p->reg_off.size(), p->reg_off.data(), p->reg_dim.data(),
                                                 p->reg_R.data(), p->min_time_D, &J, nullptr, nullptr);
        if (st != QCB_OK) return pstep(p, st, "regularisers", p->aux->err);
        v += J;
    }
    if (p->leak_obj) {
        double J = 0.0;
        qcb_status st = qcb_aux_leakage_eval(p->aux, Z, nullptr, 0.0, nullptr, &J, nullptr, nullptr, nullptr, nullptr);
        if (st != QCB_OK) return pstep(p, st, "leakage objective", p->aux->err);
        v += p->leak_Q * J;
    }
    *f = v;
    return QCB_OK;
}

// eval_grad_f: dense gradient over Z (n doubles, every entry written).
qcb_status qcb_problem_eval_grad_f(qcb_problem p, const double *Z, int32_t new_x, double *grad) {
    if (!p || !Z || !grad) return QCB_ERR_ARG;
    if (new_x) std::fill(p->stale.begin(), p->stale.end(), 1);
    if (p->n >= (1 << 20)) host_fill_pool(grad, (size_t)p->n, 0.0);      // cfg5: 263 MB of zeros around 16 k nonzeros
    else std::memset(grad, 0, (size_t)p->n * 8);
    if (p->fid) {
        qcb_status st = prob_fidelity(p, Z, true);
        if (st != QCB_OK) return st;
        const size_t xd = (size_t)qcb_fidelity_x_dim(p->fid);
        for (int s = 0; s < p->fid_states; ++s) {
            const double slope = prob_fid_slope(p, s);
            for (size_t i = 0; i < xd; ++i) grad[p->fid_pos[(size_t)s * xd + i]] += slope * p->gF[(size_t)s * xd + i];
        }
    }
    if (p->has_reg) {
        qcb_status st = qcb_aux_regularizer_eval(p->aux, Z, (int32_t)p->reg_off.size(), p->reg_off.data(), p->reg_dim.data(),
                                                 p->reg_R.data(), p->min_time_D, nullptr, p->rg.data(), nullptr);
        if (st != QCB_OK) return pstep(p, st, "regularisers", p->aux->err);
        for (size_t e = 0; e < p->rg.size(); ++e) grad[p->reg_cols[e] - 1] += p->rg[e];
    }
    if (p->leak_obj) {
        qcb_status st = qcb_aux_leakage_eval(p->aux, Z, nullptr, 0.0, nullptr, nullptr, p->lbuf.data(), nullptr, nullptr, nullptr);
        if (st != QCB_OK) return pstep(p, st, "leakage objective", p->aux->err);
        const int nl = p->aux->n_leak;
        for (int64_t e = 0; e < p->l_n; ++e) grad[(e / nl) * p->z_dim + p->aux->leak[(size_t)(e % nl)]] += p->leak_Q * p->lbuf[(size_t)e];
    }
    return QCB_OK;
}

// eval_g: the m constraint rows [integrators][derivative rows][dt_t - dt_N].
qcb_status qcb_problem_eval_g(qcb_problem p, const double *Z, int32_t new_x, double *g) {
    if (!p || !Z || !g) return QCB_ERR_ARG;
    if (new_x) std::fill(p->stale.begin(), p->stale.end(), 1);
    for (size_t i = 0; i < p->integ.size(); ++i) {
        qcb_status st = qcb_eval_defect_x(p->integ[i], Z, p->stale[i], g + p->row0[i]);
        if (st != QCB_OK) return pstep(p, st, "integrator defect", p->integ[i]->err);
        p->stale[i] = 0;
    }
    if (p->deriv_rows && p->d_rows) {
        qcb_status st = qcb_aux_derivative_eval(p->aux, Z, nullptr, g + p->d_row0, nullptr, nullptr);
        if (st != QCB_OK) return pstep(p, st, "derivative rows", p->aux->err);
    }
    if (p->dt_rows) {
        qcb_status st = qcb_aux_timesteps_equal(p->aux, Z, 0, g + p->t_row0, nullptr, nullptr, nullptr);
        if (st != QCB_OK) return pstep(p, st, "timestep rows", p->aux->err);
    }
    if (p->leak_rows) {
        qcb_status st = qcb_aux_leakage_eval(p->aux, Z, nullptr, p->leak_value, nullptr, nullptr, nullptr, g + p->l_row0, nullptr, nullptr);
        if (st != QCB_OK) return pstep(p, st, "leakage rows", p->aux->err);
    }
    return QCB_OK;
}

// eval_jac_g: values in the order of qcb_problem_jacobian_structure.
qcb_status qcb_problem_eval_jac_g(qcb_problem p, const double *Z, int32_t new_x, double *vals) {
    if (!p || !Z || !vals) return QCB_ERR_ARG;
    if (new_x) std::fill(p->stale.begin(), p->stale.end(), 1);
    for (size_t i = 0; i < p->integ.size(); ++i) {
        qcb_status st = qcb_eval_jacobian_x(p->integ[i], Z, p->stale[i], vals + p->jac0[i]);
        if (st != QCB_OK) return pstep(p, st, "integrator Jacobian", p->integ[i]->err);
        p->stale[i] = 0;
    }
    if (p->deriv_rows && p->d_rows) {
        qcb_status st = qcb_aux_derivative_eval(p->aux, Z, nullptr, nullptr, vals + p->d_jac0, nullptr);
        if (st != QCB_OK) return pstep(p, st, "derivative rows", p->aux->err);
    }
    if (p->dt_rows) {
        qcb_status st = qcb_aux_timesteps_equal(p->aux, Z, 0, nullptr, nullptr, nullptr, vals + p->t_jac0);
        if (st != QCB_OK) return pstep(p, st, "timestep rows", p->aux->err);
    }
    if (p->leak_rows) {
        qcb_status st = qcb_aux_leakage_eval(p->aux, Z, nullptr, p->leak_value, nullptr, nullptr, nullptr, nullptr, vals + p->l_jac0, nullptr);
        if (st != QCB_OK) return pstep(p, st, "leakage rows", p->aux->err);
    }
    return QCB_OK;
}

// eval_h: values of  obj_factor * Hess f + sum_i lambda_i Hess g_i  in the order of qcb_problem_hessian_structure.
// lambda: the m multipliers in the row order of eval_g.
qcb_status qcb_problem_eval_h(qcb_problem p, const double *Z, int32_t new_x, double obj_factor, const double *lambda,
                              double *vals) {
    if (!p || !Z || !vals || (p->m > 0 && !lambda)) return QCB_ERR_ARG;
    if (new_x) std::fill(p->stale.begin(), p->stale.end(), 1);
    for (size_t i = 0; i < p->integ.size(); ++i) {
        qcb_status st = qcb_eval_hessian_x(p->integ[i], Z, p->stale[i], lambda + p->row0[i], vals + p->hes0[i]);
        if (st != QCB_OK) return pstep(p, st, "integrator Hessian", p->integ[i]->err);
        p->stale[i] = 0;
    }
    if (p->deriv_rows && p->d_rows) {
        qcb_status st = qcb_aux_derivative_eval(p->aux, Z, lambda + p->d_row0, nullptr, nullptr, vals + p->d_hes0);
        if (st != QCB_OK) return pstep(p, st, "derivative rows", p->aux->err);
    }
    if (p->leak_rows) {
        qcb_status st = qcb_aux_leakage_eval(p->aux, Z, nullptr, p->leak_value, lambda + p->l_row0, nullptr, nullptr, nullptr, nullptr,
                                             vals + p->l_hes0);
        if (st != QCB_OK) return pstep(p, st, "leakage rows", p->aux->err);
    }
    if (p->leak_obj) {          // Hessian of leak_Q sum_t sum_i x^2 / n_leak: the constant diagonal 2 leak_Q / n_leak
        const double v = obj_factor * 2.0 * p->leak_Q / p->aux->n_leak;
        for (int64_t e = 0; e < p->l_n; ++e) vals[p->lo_hes0 + e] = v;
    }
    if (p->fid_hess) {
        qcb_status st = prob_fidelity(p, Z, false);          // the sign of 1 - F_s at this iterate
        if (st != QCB_OK) return st;
        const int64_t xd = qcb_fidelity_x_dim(p->fid);
        double *o = vals + p->f_hes0;
        for (int s = 0; s < p->fid_states; ++s) {
            const double slope = obj_factor * prob_fid_slope(p, s);
            for (int64_t b = 0; b < xd; ++b)
                for (int64_t a = 0; a <= b; ++a) *o++ = slope * p->HF[(size_t)(a + b * xd)];
        }
    }
    if (p->has_reg) {
        qcb_status st = qcb_aux_regularizer_eval(p->aux, Z, (int32_t)p->reg_off.size(), p->reg_off.data(), p->reg_dim.data(),
                                                 p->reg_R.data(), p->min_time_D, nullptr, nullptr, p->rh.data());
        if (st != QCB_OK) return pstep(p, st, "regularisers", p->aux->err);
        double *o = vals + p->r_hes0;
        for (int64_t e = 0; e < p->r_hes_n; ++e) o[e] = obj_factor * p->rh[(size_t)e];
    }
    return QCB_OK;
}

}  // extern "C"
