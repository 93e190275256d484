// step_driver.cu -- host drivers of the time steppers: projection (solver_projection.c:46-297)
// and RK2/RK4 (solver_rk2.c, solver_rk4.c).  Everything between upload and download stays on
// the device; the host synchronises once per time step (Poisson status + NaN flag).
#include "session.h"

#include <limits.h>
#include <math.h>
#include <string.h>

namespace cfdb {

// Source tables A_u sin(pi y_j), A_v sin(2 pi x_i): host libm, uploaded once per parameter set
// (solver_explicit_euler.c:326-331).  The time factor exp(-lambda*iter*dt) is applied per step.
static cfd_status_t ensure_tables(cfd_b200_session* s, const ns_solver_params_t* prm) {
    if (s->tab_valid && s->tab_au == prm->source_amplitude_u && s->tab_av == prm->source_amplitude_v)
        return CFD_SUCCESS;
    std::vector<double> su(s->desc.ny), sv(s->desc.nx);
    for (size_t j = 0; j < s->desc.ny; j++) su[j] = prm->source_amplitude_u * sin(M_PI * s->ys[j]);
    for (size_t i = 0; i < s->desc.nx; i++) sv[i] = prm->source_amplitude_v * sin(2.0 * M_PI * s->xs[i]);
    CFDB_CUDA_OK(cudaMemcpyAsync(s->su_tab, su.data(), su.size() * sizeof(double), cudaMemcpyHostToDevice, s->st));
    CFDB_CUDA_OK(cudaMemcpyAsync(s->sv_tab, sv.data(), sv.size() * sizeof(double), cudaMemcpyHostToDevice, s->st));
    CFDB_CUDA_OK(cudaStreamSynchronize(s->st));
    s->tab_au = prm->source_amplitude_u;
    s->tab_av = prm->source_amplitude_v;
    s->tab_valid = true;
    return CFD_SUCCESS;
}

static NsConst make_const(const cfd_b200_session* s, const ns_solver_params_t* prm) {
    NsConst c;
    memset(&c, 0, sizeof(c));
    c.dt = prm->dt;
    c.nu = prm->mu;
    c.mu = prm->mu;
    c.rho = (s->rho0 < 1e-10) ? 1.0 : s->rho0;  // solver_projection.c:195-198
    c.dt_over_rho = c.dt / c.rho;
    c.rho_over_dt = c.rho / c.dt;
    c.beta = prm->beta;
    c.T_ref = prm->T_ref;
    c.gx = prm->gravity[0];
    c.gy = prm->gravity[1];
    c.gz = prm->gravity[2];
    c.alpha = prm->alpha;
    c.src_on = (prm->source_amplitude_u != 0.0 || prm->source_amplitude_v != 0.0);
    c.src_decay = 1.0;
    return c;
}

static cfd_status_t check_thermal(const ns_solver_params_t* prm, bool dim3) {
    if (prm->alpha <= 0.0) return CFD_SUCCESS;
    const ns_thermal_bc_config_t& b = prm->thermal_bc;
    const bc_type_t f[6] = {b.left, b.right, b.bottom, b.top, b.back, b.front};
    for (int q = 0; q < (dim3 ? 6 : 4); q++)
        if (f[q] != BC_TYPE_PERIODIC && f[q] != BC_TYPE_NEUMANN && f[q] != BC_TYPE_DIRICHLET) {
            set_error(CFD_ERROR_INVALID, "unsupported thermal BC type on a face "
                                         "(only PERIODIC, NEUMANN, DIRICHLET are valid)");
            return CFD_ERROR_INVALID;
        }
    return CFD_SUCCESS;
}

// One host synchronisation per time step.  Decomposed runs first all-reduce (max) the flag: a NaN in ONE slab
// must take every rank out of the step loop at the same step, or the others block in the next exchange.
static cfd_status_t read_flags(cfd_b200_session* s, int* nonfinite) {
    if (s->desc.nranks > 1) {
        if (allreduce_int_max(s, &s->ctl->nonfinite, 1) != CFD_SUCCESS) return CFD_ERROR;
    }
    CFDB_CUDA_OK(cudaMemcpyAsync(&s->h_ctl[0], s->ctl, sizeof(Ctl), cudaMemcpyDeviceToHost, s->st));
    CFDB_CUDA_OK(cudaStreamSynchronize(s->st));
    *nonfinite = s->h_ctl[0].nonfinite;
    if (s->desc.nranks > 1 && comm_check(s) != CFD_SUCCESS) return CFD_ERROR;
    return CFD_SUCCESS;
}

static void fill_stats(cfd_b200_session* s, ns_solver_stats_t* stats, int iters) {
    if (!stats) return;
    CFDB_KL(s, launch_field_stats)(s->L(), s->u, s->v, s->w, s->p, s->T);
    if (s->desc.nranks > 1) allreduce_produced(s, &s->ctl->stat[0], 3, true);
    cudaMemcpyAsync(&s->h_ctl[0], s->ctl, sizeof(Ctl), cudaMemcpyDeviceToHost, s->st);
    cudaStreamSynchronize(s->st);
    stats->iterations = iters;
    stats->max_velocity = s->h_ctl[0].stat[0];
    stats->max_pressure = s->h_ctl[0].stat[1];
    stats->max_temperature = s->h_ctl[0].stat[2];
}

// Every device allocation a run of `kind` will need, made up front.  The drivers below allocate lazily as well
// (stand-alone session use), but a gang must not allocate while slabs already wait for each other inside kernels:
// cudaMalloc may synchronise the device, and with several slabs on one GPU (virtual ranks) the allocating thread
// would then wait for a kernel that waits for that very thread's next launch.
cfd_status_t prepare_buffers(cfd_b200_session* s, int kind, const ns_solver_params_t* prm,
                             const cfd_b200_poisson_config_t* pc) {
    CFDB_CUDA_OK(cudaSetDevice(s->dev));
    cfd_status_t rc = CFD_SUCCESS;
    const bool energy = prm->alpha > 0.0, thermo = energy || prm->beta != 0.0;
    std::vector<double**> need;
    if (kind == 0) {
        need = {&s->us, &s->vs, &s->ws, &s->pn, &s->rhs};
        if (pc->method == POISSON_METHOD_JACOBI || pc->method == POISSON_METHOD_REDBLACK_SOR) need.push_back(&s->xt);
        const int nk = pc->method == POISSON_METHOD_CG ? 4 : (pc->method == POISSON_METHOD_BICGSTAB ? 8 : 0);
        for (int q = 0; q < nk; q++) need.push_back(&s->kv[q]);
    } else {
        const int nbuf = kind == 1 ? 4 : (kind == 4 ? 12 : 8);
        for (int q = 0; q < nbuf; q++) need.push_back(&s->rk[q]);
    }
    if (thermo) need.push_back(&s->T);
    if (energy) need.push_back(&s->Tn);
    for (double** q : need)
        if ((rc = dalloc(s, q)) != CFD_SUCCESS) return rc;
    if ((rc = ensure_tables(s, prm)) != CFD_SUCCESS) return rc;
    CFDB_CUDA_OK(cudaStreamSynchronize(s->st));
    return CFD_SUCCESS;
}

cfd_status_t run_projection(cfd_b200_session* s, const ns_solver_params_t* prm,
                            const cfd_b200_poisson_config_t* pc, int nsteps, ns_solver_stats_t* stats) {
    if (!s || !prm || !pc) return CFD_ERROR_INVALID;
    if (prm->source_func || prm->heat_source_func) {
        set_error(CFD_ERROR_UNSUPPORTED, "host source/heat-source callbacks cannot run on the GPU");
        return CFD_ERROR_UNSUPPORTED;
    }
    cfd_status_t rc = check_thermal(prm, s->g.dim3);
    if (rc != CFD_SUCCESS) return rc;
    CFDB_CUDA_OK(cudaSetDevice(s->dev));
    t_prof = &s->prof;
    if ((rc = comm_ready(s)) != CFD_SUCCESS) return rc;
    double** need[] = {&s->us, &s->vs, &s->ws, &s->pn, &s->rhs};
    for (auto q : need)
        if ((rc = dalloc(s, q)) != CFD_SUCCESS) return rc;
    if (pc->method == POISSON_METHOD_JACOBI && (rc = dalloc(s, &s->xt)) != CFD_SUCCESS) return rc;
    const bool energy = prm->alpha > 0.0;
    const bool thermo = energy || prm->beta != 0.0;
    if (thermo && (rc = dalloc(s, &s->T)) != CFD_SUCCESS) return rc;  // (zeros unless uploaded)
    if (energy && (rc = dalloc(s, &s->Tn)) != CFD_SUCCESS) return rc;
    if ((rc = ensure_tables(s, prm)) != CFD_SUCCESS) return rc;
    const int R = s->desc.nranks;
    if (R > 1) {
        // the projection uses field->rho[0] of the GLOBAL field (solver_projection.c:195): rank 0 owns it
        const double mine = (s->desc.rank == 0) ? s->rho0 : 0.0;
        CFDB_CUDA_OK(cudaMemcpyAsync(&s->ctl->stat[3], &mine, sizeof(double), cudaMemcpyHostToDevice, s->st));
        if ((rc = allreduce(s, &s->ctl->stat[3], 1, false)) != CFD_SUCCESS) return rc;
        double all = 0.0;
        CFDB_CUDA_OK(cudaMemcpyAsync(&all, &s->ctl->stat[3], sizeof(double), cudaMemcpyDeviceToHost, s->st));
        CFDB_CUDA_OK(cudaStreamSynchronize(s->st));
        s->rho0 = all;
    }
    NsConst c = make_const(s, prm);
    CFDB_CUDA_OK(cudaMemsetAsync(&s->ctl->nonfinite, 0, sizeof(int), s->st));
    for (int iter = 0; iter < nsteps; iter++) {
        const LaunchCtx L = s->L();
        c.src_decay = exp(-prm->source_decay_rate * iter * prm->dt);
        if (R > 1) {
            double* f[4] = {s->u, s->v, s->w, s->T};
            halo_exchange(s, f, thermo ? 4 : 3, false);
        }
        CFDB_KL(s, launch_predictor)(L, c, s->u, s->v, s->w, s->T, s->su_tab, s->sv_tab, s->us, s->vs, s->ws);
        if (R > 1) {
            double* f[3] = {s->us, s->vs, s->ws};
            halo_exchange(s, f, 3, false);
        }
        CFDB_KL(s, launch_divergence)(L, c, s->us, s->vs, s->ws, s->rhs);
        CFDB_KL(s, launch_copy)(L, s->pn, s->p, s->N);  // p_new starts from the previous pressure (:108, :253)
        poisson_solver_stats_t ps;
        rc = poisson_solve(s, pc, &s->pn, &s->xt, s->rhs, &ps);
        if (rc == CFD_ERROR || rc == CFD_ERROR_NOMEM || rc == CFD_ERROR_UNSUPPORTED || rc == CFD_ERROR_INVALID)
            return rc;
        if (rc != CFD_SUCCESS && pc->on_nonconvergence == CFD_B200_NONCONV_ERROR) {
            set_error(CFD_ERROR_MAX_ITER, "pressure Poisson solve did not converge (status %d after %d iterations)",
                      (int)ps.status, ps.iterations);
            return CFD_ERROR_MAX_ITER;  // solver_projection.c:220-224
        }
        CFDB_KL(s, launch_corrector)(L, c, s->us, s->vs, s->ws, s->pn, s->u, s->v, s->w, s->T, s->Tn);
        std::swap(s->p, s->pn);  // memcpy(field->p, p_new) without the copy
        if (energy) {
            std::swap(s->T, s->Tn);
            CFDB_KL(s, launch_thermal_bc)(s->L(), prm->thermal_bc, s->T);
        }
        int bad = 0;
        if ((rc = read_flags(s, &bad)) != CFD_SUCCESS) return rc;
        if (bad) {
            set_error(CFD_ERROR_DIVERGED, "NaN/Inf detected in projection step %d", iter);
            return CFD_ERROR_DIVERGED;
        }
    }
    if (comm_check(s) != CFD_SUCCESS) return CFD_ERROR;
    cudaError_t ce = cudaGetLastError();
    if (ce != cudaSuccess) {
        set_error(CFD_ERROR, "CUDA failure in projection step: %s", cudaGetErrorString(ce));
        return CFD_ERROR;
    }
    fill_stats(s, stats, nsteps);
    return CFD_SUCCESS;
}

// Explicit Euler, cpu/solver_explicit_euler.c:337-582 (the `explicit_euler_gpu` name).  Per step:
// one out-of-place stencil pass over the interior; the caller's boundary velocities (u, v, and w
// when nz > 1) are carried into the new arrays (the reference saves them, wraps, restores: :549-553);
// energy step with the new velocity; periodic images on the boundary cells of p, rho, T (and w when
// nz == 1, which the reference does not restore); thermal BCs; NaN/Inf scan.
static cfd_status_t run_euler(cfd_b200_session* s, const ns_solver_params_t* prm, int nsteps,
                              ns_solver_stats_t* stats) {
    if (prm->source_func || prm->heat_source_func) {
        set_error(CFD_ERROR_UNSUPPORTED, "host source/heat-source callbacks cannot run on the GPU");
        return CFD_ERROR_UNSUPPORTED;
    }
    cfd_status_t rc = check_thermal(prm, s->g.dim3);
    if (rc != CFD_SUCCESS) return rc;
    CFDB_CUDA_OK(cudaSetDevice(s->dev));
    t_prof = &s->prof;
    if ((rc = comm_ready(s)) != CFD_SUCCESS) return rc;
    for (int q = 0; q < 4; q++)
        if ((rc = dalloc(s, &s->rk[q])) != CFD_SUCCESS) return rc;
    if (!s->rho) {  // session extension: no density uploaded => rho = 1 everywhere
        if ((rc = dalloc(s, &s->rho)) != CFD_SUCCESS) return rc;
        CFDB_KL(s, launch_fill)(s->L(), s->rho, 1.0, s->N);
    }
    const bool energy = prm->alpha > 0.0;
    if ((energy || prm->beta != 0.0) && (rc = dalloc(s, &s->T)) != CFD_SUCCESS) return rc;
    if (energy && (rc = dalloc(s, &s->Tn)) != CFD_SUCCESS) return rc;
    if ((rc = ensure_tables(s, prm)) != CFD_SUCCESS) return rc;
    NsConst c = make_const(s, prm);
    EulerArgs ea;
    ea.dt = fmin(prm->dt, 1e-4);  // DT_CONSERVATIVE_LIMIT (:395)
    ea.p_factor = -0.1 * ea.dt;   // -PRESSURE_UPDATE_FACTOR * conservative_dt (:518)
    c.dt = ea.dt;                 // the energy step advances by the same conservative dt (:537-539)
    CFDB_CUDA_OK(cudaMemsetAsync(&s->ctl->nonfinite, 0, sizeof(int), s->st));
    // slabs: the stencil reads the neighbours' edge layers (plain halos; the caller's boundary velocities
    // must survive, so velocities are never ring-wrapped), the periodic images of p / rho / T along the
    // decomposed axis come from the ring exchange, as in the RK steppers
    const int R = s->desc.nranks;
    const int wrap_y = !(R > 1 && s->dec_axis == 1), wrap_z = !(R > 1 && s->dec_axis == 2);
    for (int iter = 0; iter < nsteps; iter++) {
        const LaunchCtx L = s->L();
        ea.src_decay = exp(-prm->source_decay_rate * iter * ea.dt);
        P4 in;
        W4 out;
        double** cur[4] = {&s->u, &s->v, &s->w, &s->p};
        for (int q = 0; q < 4; q++) { in.a[q] = *cur[q]; out.a[q] = s->rk[q]; }
        if (R > 1) {
            double* f[6] = {s->u, s->v, s->w, s->p, s->rho, s->T};
            halo_exchange(s, f, s->T ? 6 : 5, false);
        }
        CFDB_KL(s, launch_euler)(L, c, ea, in, out, s->rho, s->T, s->su_tab, s->sv_tab);
        CFDB_KL(s, launch_boundary_carry3)(L, s->u, s->v, s->w, s->rk[0], s->rk[1], s->rk[2]);
        for (int q = 0; q < 4; q++) std::swap(*cur[q], s->rk[q]);  // memcpy(field <- new) without the copy (:528-531)
        if (energy) {
            CFDB_KL(s, launch_energy)(L, c, s->u, s->v, s->w, s->T, s->Tn);  // (the halo of T is current: exchanged above)
            std::swap(s->T, s->Tn);
        }
        // apply_boundary_conditions (:231-314) on everything the velocity restore does not undo
        double* wrap[4];
        int nw = 0;
        wrap[nw++] = s->p;
        wrap[nw++] = s->rho;
        if (s->T) wrap[nw++] = s->T;  // a T that never reached the device stays on the host untouched
        if (!s->g.dim3) wrap[nw++] = s->w;
        CFDB_KL(s, launch_periodic_faces)(L, wrap, nw, wrap_y, wrap_z);
        if (R > 1) halo_exchange(s, wrap, nw, true);
        if (energy) CFDB_KL(s, launch_thermal_bc)(L, prm->thermal_bc, s->T);
        CFDB_KL(s, launch_boundary_scan)(L, s->u, s->v, s->w, s->p);
        int bad = 0;
        if ((rc = read_flags(s, &bad)) != CFD_SUCCESS) return rc;
        if (bad) {
            set_error(CFD_ERROR_DIVERGED, "NaN/Inf detected in explicit_euler step %d", iter);
            return CFD_ERROR_DIVERGED;
        }
    }
    if (comm_check(s) != CFD_SUCCESS) return CFD_ERROR;
    cudaError_t ce = cudaGetLastError();
    if (ce != cudaSuccess) {
        set_error(CFD_ERROR, "CUDA failure in explicit Euler step: %s", cudaGetErrorString(ce));
        return CFD_ERROR;
    }
    fill_stats(s, stats, nsteps);
    return CFD_SUCCESS;
}

cfd_status_t run_rk(cfd_b200_session* s, int order, const ns_solver_params_t* prm, int nsteps,
                    ns_solver_stats_t* stats) {
    if (!s || !prm) return CFD_ERROR_INVALID;
    if (order == 1) return run_euler(s, prm, nsteps, stats);
    if (order != 2 && order != 4) {
        set_error(CFD_ERROR_UNSUPPORTED, "RK order %d is not implemented on the GPU", order);
        return CFD_ERROR_UNSUPPORTED;
    }
    if (prm->source_func || prm->heat_source_func) {
        set_error(CFD_ERROR_UNSUPPORTED, "host source/heat-source callbacks cannot run on the GPU");
        return CFD_ERROR_UNSUPPORTED;
    }
    cfd_status_t rc = check_thermal(prm, s->g.dim3);
    if (rc != CFD_SUCCESS) return rc;
    CFDB_CUDA_OK(cudaSetDevice(s->dev));
    t_prof = &s->prof;
    if ((rc = comm_ready(s)) != CFD_SUCCESS) return rc;
    const int nbuf = (order == 4) ? 12 : 8;  // B[4], (C[4]), ACC[4]
    for (int q = 0; q < nbuf; q++)
        if ((rc = dalloc(s, &s->rk[q])) != CFD_SUCCESS) return rc;
    if (!s->rho) {  // session extension: no density uploaded => rho = 1 everywhere
        if ((rc = dalloc(s, &s->rho)) != CFD_SUCCESS) return rc;
        CFDB_KL(s, launch_fill)(s->L(), s->rho, 1.0, s->N);
    }
    const bool energy = prm->alpha > 0.0;
    if ((energy || prm->beta != 0.0) && (rc = dalloc(s, &s->T)) != CFD_SUCCESS) return rc;
    if (energy && (rc = dalloc(s, &s->Tn)) != CFD_SUCCESS) return rc;
    if ((rc = ensure_tables(s, prm)) != CFD_SUCCESS) return rc;
    NsConst c = make_const(s, prm);
    const int R = s->desc.nranks;
    const int wrap_y = !(R > 1 && s->dec_axis == 1), wrap_z = !(R > 1 && s->dec_axis == 2);
    double* Aq[4];
    double* B[4] = {s->rk[0], s->rk[1], s->rk[2], s->rk[3]};
    double* C[4] = {s->rk[4], s->rk[5], s->rk[6], s->rk[7]};
    double* ACC[4];
    for (int q = 0; q < 4; q++) ACC[q] = s->rk[(order == 4 ? 8 : 4) + q];
    CFDB_CUDA_OK(cudaMemsetAsync(&s->ctl->nonfinite, 0, sizeof(int), s->st));
    const double dt = prm->dt;
    for (int iter = 0; iter < nsteps; iter++) {
        const LaunchCtx L = s->L();
        Aq[0] = s->u; Aq[1] = s->v; Aq[2] = s->w; Aq[3] = s->p;
        auto P = [](double* const a[4]) { P4 p; for (int q = 0; q < 4; q++) p.a[q] = a[q]; return p; };
        auto W = [](double* const a[4]) { W4 w; for (int q = 0; q < 4; q++) w.a[q] = a[q]; return w; };
        RkArgs ra;
        memset(&ra, 0, sizeof(ra));
        ra.wrap_y = wrap_y;
        ra.wrap_z = wrap_z;
        ra.src_decay = exp(-prm->source_decay_rate * iter * dt);
        // stage inputs carry their periodic images on the boundary cells (the TMA stage kernel reads
        // plain halos; the reference reads wrapped indices -- same values)
        CFDB_KL(s, launch_periodic_faces)(L, Aq, 4, wrap_y, wrap_z);
        if (R > 1) halo_exchange(s, Aq, 4, true);
        if (order == 4) {
            // stage 1: k1 = f(Q0); acc = k1; B = Q0 + dt/2 k1
            ra.first = 1; ra.last = 0; ra.fac_next = 0.5 * dt; ra.w_k = 1.0;
            CFDB_KL(s, launch_rk_stage)(L, c, ra, P(Aq), P(Aq), W(ACC), W(B), s->rho, s->T, s->su_tab, s->sv_tab);
            CFDB_KL(s, launch_periodic_faces)(L, B, 4, wrap_y, wrap_z);
            if (R > 1) halo_exchange(s, B, 4, true);
            // stage 2: k2 = f(B); acc += 2 k2; C = Q0 + dt/2 k2
            ra.first = 0; ra.w_k = 2.0;
            CFDB_KL(s, launch_rk_stage)(L, c, ra, P(B), P(Aq), W(ACC), W(C), s->rho, s->T, s->su_tab, s->sv_tab);
            CFDB_KL(s, launch_periodic_faces)(L, C, 4, wrap_y, wrap_z);
            if (R > 1) halo_exchange(s, C, 4, true);
            // stage 3: k3 = f(C); acc += 2 k3; B = Q0 + dt k3
            ra.fac_next = dt;
            CFDB_KL(s, launch_rk_stage)(L, c, ra, P(C), P(Aq), W(ACC), W(B), s->rho, s->T, s->su_tab, s->sv_tab);
            CFDB_KL(s, launch_periodic_faces)(L, B, 4, wrap_y, wrap_z);
            if (R > 1) halo_exchange(s, B, 4, true);
            // stage 4: k4 = f(B); Q = Q0 + dt/6 (acc + k4)
            ra.last = 1; ra.w_final = dt / 6.0;
            CFDB_KL(s, launch_rk_stage)(L, c, ra, P(B), P(Aq), W(ACC), W(Aq), s->rho, s->T, s->su_tab, s->sv_tab);
        } else {
            ra.first = 1; ra.last = 0; ra.fac_next = dt; ra.w_k = 1.0;
            CFDB_KL(s, launch_rk_stage)(L, c, ra, P(Aq), P(Aq), W(ACC), W(B), s->rho, s->T, s->su_tab, s->sv_tab);
            CFDB_KL(s, launch_periodic_faces)(L, B, 4, wrap_y, wrap_z);
            if (R > 1) halo_exchange(s, B, 4, true);
            ra.first = 0; ra.last = 1; ra.w_final = 0.5 * dt;
            CFDB_KL(s, launch_rk_stage)(L, c, ra, P(B), P(Aq), W(ACC), W(Aq), s->rho, s->T, s->su_tab, s->sv_tab);
        }
        if (energy) {
            if (R > 1) { double* f[1] = {s->T}; halo_exchange(s, f, 1, false); }
            CFDB_KL(s, launch_energy)(L, c, s->u, s->v, s->w, s->T, s->Tn);
            std::swap(s->T, s->Tn);
        }
        // apply_boundary_conditions: periodic wrap of u,v,w,p,rho,T (solver_explicit_euler.c:231-314)
        double* all[6] = {s->u, s->v, s->w, s->p, s->rho, s->T};
        const int nall = s->T ? 6 : 5;  // a T that never reached the device stays on the host untouched
        CFDB_KL(s, launch_periodic_faces)(L, all, nall, wrap_y, wrap_z);
        if (R > 1) halo_exchange(s, all, nall, true);
        if (energy) CFDB_KL(s, launch_thermal_bc)(s->L(), prm->thermal_bc, s->T);
        int bad = 0;
        if ((rc = read_flags(s, &bad)) != CFD_SUCCESS) return rc;
        if (bad) {
            set_error(CFD_ERROR_DIVERGED, "NaN/Inf detected in RK%d step %d", order, iter);
            return CFD_ERROR_DIVERGED;
        }
    }
    if (comm_check(s) != CFD_SUCCESS) return CFD_ERROR;
    cudaError_t ce = cudaGetLastError();
    if (ce != cudaSuccess) {
        set_error(CFD_ERROR, "CUDA failure in RK step: %s", cudaGetErrorString(ce));
        return CFD_ERROR;
    }
    fill_stats(s, stats, nsteps);
    return CFD_SUCCESS;
}

}  // namespace cfdb
