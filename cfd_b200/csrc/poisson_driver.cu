// poisson_driver.cu -- host drivers of the four Poisson methods.  The host only enqueues
// kernels: batches of iterations are launched back to back, the device control block decides
// convergence, and the host looks at a pinned copy of it one batch late (so the GPU never
// waits for the CPU).  Launches after convergence are no-ops.
#include "session.h"

#include <limits.h>
#include <math.h>

namespace cfdb {

namespace {

struct Poll {
    cfd_b200_session* s;
    int batch = 0;
    // enqueue a copy of the control block; returns true when an OLDER copy already shows "done"
    bool step() {
        const int cur = batch & 1;
        cudaMemcpyAsync(&s->h_ctl[cur], s->ctl, sizeof(Ctl), cudaMemcpyDeviceToHost, s->st);
        cudaEventRecord(s->ev[cur], s->st);
        bool done = false;
        if (batch >= 1) {
            cudaEventSynchronize(s->ev[cur ^ 1]);
            done = s->h_ctl[cur ^ 1].done_seq != INT_MAX;
        }
        batch++;
        return done;
    }
    // final: wait for everything and return the latest control block
    const Ctl& final_state() {
        cudaMemcpyAsync(&s->h_ctl[0], s->ctl, sizeof(Ctl), cudaMemcpyDeviceToHost, s->st);
        cudaStreamSynchronize(s->st);
        return s->h_ctl[0];
    }
};

int batch_size(const cfd_b200_session* s) {
    // small grids: launch-bound, poll rarely; big grids: each iteration is milliseconds
    const double cells = (double)s->N;
    if (cells < 1e5) return 64;
    if (cells < 4e6) return 32;
    if (cells < 1e8) return 8;
    return 4;
}

double optimal_omega(const cfd_b200_session_desc_t& d, size_t gnx, size_t gny, size_t gnz) {
    // linear_solver_internal.h:184-220 (global extents)
    const double ix = 1.0 / (d.dx * d.dx), iy = 1.0 / (d.dy * d.dy);
    const double iz = (d.dz > 0.0) ? 1.0 / (d.dz * d.dz) : 0.0;
    double num = cos(M_PI / (double)(gnx - 1)) * ix + cos(M_PI / (double)(gny - 1)) * iy;
    double den = ix + iy;
    if (gnz > 1 && iz > 0.0) {
        num += cos(M_PI / (double)(gnz - 1)) * iz;
        den += iz;
    }
    const double rj = num / den;
    return 2.0 / (1.0 + sqrt(1.0 - (rj * rj)));
}

}  // namespace

cfd_status_t poisson_solve(cfd_b200_session* s, const cfd_b200_poisson_config_t* cfg, double** px,
                           double** pxt, const double* rhs, poisson_solver_stats_t* out) {
    if (cfg->method == POISSON_METHOD_MULTIGRID) return mg_solve(s, cfg, *px, rhs, out);
    if (cfg->method == POISSON_METHOD_CG && (int)cfg->params.preconditioner == (int)CFD_B200_PRECOND_MULTIGRID)
        return mgpcg_solve(s, cfg, *px, rhs, out);
    {
        const cfd_status_t cr = comm_ready(s);
        if (cr != CFD_SUCCESS) return cr;
    }
    const poisson_solver_params_t& pp = cfg->params;
    const LaunchCtx L = s->L();
    const int R = s->desc.nranks;
    KArgs a;
    a.seq = 0;
    a.it = 0;
    a.check = pp.check_interval > 0 ? pp.check_interval : 1;
    a.max_iter = pp.max_iterations;
    a.rel_tol = pp.tolerance;
    a.abs_tol = pp.absolute_tolerance;
    a.dinv = 1.0 / (2.0 * s->g.inv_dx2 + 2.0 * s->g.inv_dy2 + 2.0 * s->g.inv_dz2);
    a.zchunk = s->zchunk;
    {
        // global extents along the decomposed axis (every rank owns n-2 interior layers)
        size_t gnx = s->desc.nx, gny = s->desc.ny, gnz = s->desc.nz;
        if (s->g.dim3) gnz = global_dec_extent(s);  // uneven slabs: cfd_b200_session_set_slab
        else gny = global_dec_extent(s);
        a.omega = pp.omega > 0.0 ? pp.omega : optimal_omega(s->desc, gnx, gny, gnz);
    }
    // Halo exchange of <= 2 fields whose PULL half rides in the next sweep's edge blocks (HaloGate) when the transport
    // (peer memory) and the kernels (TMA engines) allow it: the interior of the sweep then runs under the neighbour's
    // push.  Otherwise the plain exchange (push + pull kernels, or NCCL send/recv).
    auto exchange = [&](double* f0, double* f1) {
        a.gate = HaloGate();
        if (R <= 1) return;
        double* f[2] = {f0, f1};
        const int nf = f1 ? 2 : 1;
        if (s->peer.on && s->gate_enabled && L.tc && L.tc->usable) peer_halo_push_gated(s, f, nf, &a.gate);
        else halo_exchange(s, f, nf, false);
    };
    auto flush_gate = [&]() {  // before a launch that cannot take the gate
        if (a.gate.on) peer_pull_from_gate(s->st, a.gate);
        a.gate = HaloGate();
    };
    // One-pass Red-Black SOR on z-slabs: peer transport (its box holds the second halo layer) and >= 2 own planes on
    // EVERY rank -- agreed once per slab layout with an all-reduce, since ranks that took different paths would wait
    // for exchanges the others never issue.
    const int method = cfg->method;
    double* const* halo2 = nullptr;
    if (method == POISSON_METHOD_REDBLACK_SOR && R > 1 && s->g.dim3 && s->peer.on && s->peer.halo2[0]) {
        if (s->rb_slab_ok < 0) {
            CFDB_CUDA_OK(cudaMemsetAsync(&s->ctl->pad[0], s->g.nz >= 4 ? 0 : 1, sizeof(int), s->st));
            if (allreduce_int_max(s, &s->ctl->pad[0], 1) != CFD_SUCCESS) return CFD_ERROR;
            int bad = 1;
            CFDB_CUDA_OK(cudaMemcpyAsync(&bad, &s->ctl->pad[0], sizeof(int), cudaMemcpyDeviceToHost, s->st));
            CFDB_CUDA_OK(cudaStreamSynchronize(s->st));
            s->rb_slab_ok = bad ? 0 : 1;
        }
        if (s->rb_slab_ok == 1) halo2 = s->peer.halo2;
    }
    const bool rb_fused = method == POISSON_METHOD_REDBLACK_SOR && (R == 1 || halo2) &&
                          CFDB_KL(s, rbsor_fused_available)(L, halo2);
    CFDB_KL(s, launch_zero_ctl)(L);
    Poll poll{s};
    const int B = batch_size(s);
    double* x = *px;
    int seq = 1;
    cfd_status_t rc = CFD_SUCCESS;

    // the one-pass Red-Black SOR iteration is out of place: it needs the ping-pong partner as well
    // (no memory for it => the in-place half-sweeps below)
    if (rb_fused && pxt && !*pxt && dalloc(s, pxt) != CFD_SUCCESS) *pxt = nullptr;

    if (method == POISSON_METHOD_JACOBI && R == 1 && a.max_iter >= 1 && jacobi_cluster_available(L) &&
        (a.seq = seq, launch_jacobi_cluster(L, a, x, rhs))) {
        // small 2-D grid: the whole solve ran in one cluster launch, x updated in place
        seq++;
        (void)poll.final_state();
    } else if (method == POISSON_METHOD_JACOBI) {
        if (!pxt || !*pxt) return CFD_ERROR_INVALID;
        double* buf[2] = {x, *pxt};
        // sweeps q = 0..max_iter (sweep q computes res(x_q)); then the final decision
        for (int q = 0; q <= a.max_iter;) {
            for (int b = 0; b < B && q <= a.max_iter; b++, q++) {
                a.seq = seq++;
                a.it = q;
                CFDB_KL(s, launch_jacobi)(L, a, buf[q & 1], buf[(q + 1) & 1], rhs);  // (pulls the halo of its input)
                if (R > 1) {
                    exchange(buf[(q + 1) & 1], nullptr);
                    allreduce_produced(s, &s->ctl->dot[q & 1], 1, true);
                }
            }
            if (poll.step()) break;
        }
        flush_gate();
        a.seq = seq++;
        a.it = a.max_iter;
        CFDB_KL(s, launch_stationary_final)(L, a);
        const Ctl& c = poll.final_state();
        if (c.result_buf == 1) {  // x lives in the ping-pong partner: swap the caller's pointers
            *px = buf[1];
            *pxt = buf[0];
        }
    } else if (rb_fused && pxt && *pxt && L.tc->get_wide(x) && L.tc->get_wide(*pxt) && L.tc->get(rhs) &&
               (!halo2 || (L.tc->get_wide(halo2[0]) && L.tc->get_wide(halo2[1])))) {
        // one pass per iteration, same protocol as Jacobi: sweep q reads x_q, writes x_{q+1}, leaves res(x_q)
        double* buf[2] = {x, *pxt};
        // slabs: x travels two layers deep, once per iteration; the pull rides in the next sweep's edge blocks
        auto exchange2 = [&](double* f) {
            a.gate = HaloGate();
            if (peer_halo_exchange2(s, f, s->gate_enabled ? &a.gate : nullptr) != CFD_SUCCESS && s->comm_err == CFD_SUCCESS)
                s->comm_err = CFD_ERROR;
        };
        if (halo2) {
            // the first half-sweep on the neighbour's edge plane reads its right-hand side
            double* fr[1] = {const_cast<double*>(rhs)};
            if ((rc = halo_exchange(s, fr, 1, false)) != CFD_SUCCESS) return rc;
            exchange2(x);
        }
        for (int q = 0; q <= a.max_iter;) {
            for (int b = 0; b < B && q <= a.max_iter; b++, q++) {
                a.seq = seq++;
                a.it = q;
                if (!CFDB_KL(s, launch_rbsor_fused)(L, a, buf[q & 1], buf[(q + 1) & 1], rhs, halo2)) return CFD_ERROR;
                if (halo2) {
                    exchange2(buf[(q + 1) & 1]);
                    allreduce_produced(s, &s->ctl->dot[q & 1], 1, true);
                }
            }
            if (poll.step()) break;
        }
        flush_gate();
        a.seq = seq++;
        a.it = a.max_iter;
        CFDB_KL(s, launch_stationary_final)(L, a);
        const Ctl& c = poll.final_state();
        if (c.result_buf == 1) {
            *px = buf[1];
            *pxt = buf[0];
        }
    } else if (method == POISSON_METHOD_REDBLACK_SOR) {
        a.seq = seq++;
        CFDB_KL(s, launch_linf_residual)(L, a, x, rhs, 0);
        if (R > 1) allreduce_produced(s, &s->ctl->dot[0], 1, true);
        for (int it = 0; it < a.max_iter;) {
            for (int b = 0; b < B && it < a.max_iter; b++, it++) {
                a.it = it;
                a.seq = seq++;
                CFDB_KL(s, launch_redblack)(L, a, x, rhs, 0, 1);
                exchange(x, nullptr);
                a.seq = seq++;
                CFDB_KL(s, launch_redblack)(L, a, x, rhs, 1, 0);
                exchange(x, nullptr);
                if (it % a.check == 0) {
                    a.seq = seq++;
                    CFDB_KL(s, launch_linf_residual)(L, a, x, rhs, (it + 1) & 1);
                    a.gate = HaloGate();
                    if (R > 1) allreduce_produced(s, &s->ctl->dot[(it + 1) & 1], 1, true);
                }
            }
            if (poll.step()) break;
        }
        flush_gate();
        a.seq = seq++;
        a.it = a.max_iter;
        CFDB_KL(s, launch_stationary_final)(L, a);
        (void)poll.final_state();
    } else if (method == POISSON_METHOD_CG && R == 1 && cg_cluster_available(L) &&
               (CFDB_KL(s, launch_neumann_fill)(L, x, 0, 0, false), a.seq = seq,
                launch_cg_cluster(L, a, x, rhs, pp.preconditioner == POISSON_PRECOND_JACOBI))) {
        // small 2-D grid: the whole (P)CG solve ran in one cluster launch between the two Neumann passes
        // (a failed launch leaves only the -- idempotent -- initial pass behind and falls through)
        seq++;
        CFDB_KL(s, launch_neumann_fill)(L, x, 0, 1, true);
        (void)poll.final_state();
    } else if (method == POISSON_METHOD_CG) {
        const int pre = (pp.preconditioner == POISSON_PRECOND_JACOBI);
        for (int q = 0; q < 4; q++)
            if ((rc = dalloc(s, &s->kv[q])) != CFD_SUCCESS) return rc;
        double *r = s->kv[0], *Ap = s->kv[1], *pb[2] = {s->kv[2], s->kv[3]};
        CFDB_KL(s, launch_neumann_fill)(L, x, 0, 0, false);
        a.seq = seq++;
        CFDB_KL(s, launch_krylov_init)(L, a, x, rhs, r, pb[0], nullptr, pre);
        if (R > 1) {
            exchange(r, pb[0]);
            allreduce_produced(s, &s->ctl->dot[0], 2, false);
        }
        for (int it = 0; it < a.max_iter;) {
            for (int b = 0; b < B && it < a.max_iter; b++, it++) {
                a.it = it;
                a.seq = seq++;
                // iteration it: p_it lives in pb[it&1]; it == 0 only applies A to p_0
                double* pold = (it == 0) ? pb[0] : pb[(it - 1) & 1];
                double* pnew = pb[it & 1];
                CFDB_KL(s, launch_cg_dir)(L, a, r, pold, pnew, Ap, pre, 0);  // (pulls the halo of r -- and of p_0)
                a.gate = HaloGate();
                if (R > 1) allreduce_produced(s, &s->ctl->dot[4 + (it & 1)], 1, false);
                a.seq = seq++;
                CFDB_KL(s, launch_cg_update)(L, a, x, r, pnew, Ap, pre);
                if (R > 1) {
                    exchange(r, nullptr);
                    allreduce_produced(s, &s->ctl->dot[2 * ((it + 1) & 1)], 2, false);
                }
            }
            if (poll.step()) break;
        }
        a.seq = seq++;
        a.it = a.max_iter;
        CFDB_KL(s, launch_cg_dir)(L, a, r, pb[0], pb[1], Ap, pre, 1);  // (decide only; a pending pull is done stand-alone)
        a.gate = HaloGate();
        CFDB_KL(s, launch_neumann_fill)(L, x, 0, 1, true);
        (void)poll.final_state();
    } else if (method == POISSON_METHOD_BICGSTAB) {
        for (int q = 0; q < 8; q++)
            if ((rc = dalloc(s, &s->kv[q])) != CFD_SUCCESS) return rc;
        double *r = s->kv[0], *rh = s->kv[1], *sv = s->kv[2], *tv = s->kv[3];
        double *pb[2] = {s->kv[4], s->kv[5]}, *vb[2] = {s->kv[6], s->kv[7]};
        CFDB_KL(s, launch_neumann_fill)(L, x, 0, 0, false);
        a.seq = seq++;
        CFDB_KL(s, launch_krylov_init)(L, a, x, rhs, r, nullptr, rh, 0);
        if (R > 1) {
            exchange(r, nullptr);
            allreduce_produced(s, &s->ctl->dot[0], 2, false);
        }
        for (int it = 0; it < a.max_iter;) {
            for (int b = 0; b < B && it < a.max_iter; b++, it++) {
                a.it = it;
                a.seq = seq++;
                double *pold = pb[(it + 1) & 1], *pnew = pb[it & 1];
                double *vold = vb[(it + 1) & 1], *vnew = vb[it & 1];
                CFDB_KL(s, launch_bicg_dir)(L, a, r, pold, vold, rh, pnew, vnew, it == 0, 0);  // (pulls the halo of r)
                if (R > 1) {
                    exchange(vnew, nullptr);
                    allreduce_produced(s, &s->ctl->dot[4 + (it & 1)], 1, false);
                }
                a.seq = seq++;
                CFDB_KL(s, launch_bicg_s)(L, a, r, vnew, sv, tv);  // (pulls the halo of v)
                a.gate = HaloGate();
                if (R > 1) allreduce_produced(s, &s->ctl->dot[8 + 4 * (it & 1)], 3, false);
                a.seq = seq++;
                CFDB_KL(s, launch_bicg_update)(L, a, x, r, pnew, sv, tv, rh);
                if (R > 1) {
                    exchange(r, nullptr);
                    allreduce_produced(s, &s->ctl->dot[2 * ((it + 1) & 1)], 2, false);
                }
            }
            if (poll.step()) break;
        }
        a.seq = seq++;
        a.it = a.max_iter;
        CFDB_KL(s, launch_bicg_dir)(L, a, r, pb[0], vb[0], rh, pb[1], vb[1], 0, 1);  // (decide only)
        a.gate = HaloGate();
        CFDB_KL(s, launch_neumann_fill)(L, x, 0, 1, true);
        (void)poll.final_state();
    } else {
        set_error(CFD_ERROR_UNSUPPORTED, "Poisson method %d has no GPU implementation", method);
        return CFD_ERROR_UNSUPPORTED;
    }

    cudaError_t ce = cudaGetLastError();
    if (ce != cudaSuccess) {
        set_error(CFD_ERROR, "CUDA failure in Poisson solve: %s", cudaGetErrorString(ce));
        return CFD_ERROR;
    }
    const Ctl& c = s->h_ctl[0];
    poisson_solver_stats_t st;
    st.status = (poisson_solver_status_t)c.status;
    st.iterations = c.iterations;
    st.initial_residual = c.r0;
    st.final_residual = c.res;
    st.elapsed_time_ms = 0.0;
    if (c.done_seq == INT_MAX) st.status = POISSON_ERROR;
    s->last_ps = st;
    if (out) *out = st;
    if (R > 1) {  // make the halo of the solution current for the caller
        double* f[1] = {*px};
        halo_exchange(s, f, 1, false);
    }
    if (comm_check(s) != CFD_SUCCESS) return CFD_ERROR;  // an exchange or reduction was lost: no per-slab "success"
    return st.status == POISSON_CONVERGED ? CFD_SUCCESS
                                          : (st.status == POISSON_ERROR ? CFD_ERROR : CFD_ERROR_MAX_ITER);
}

}  // namespace cfdb
