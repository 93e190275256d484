// poisson_mgpc.cu -- CG preconditioned by one multigrid V(1,1) cycle built for the reference's Krylov operator
// (POISSON_METHOD_CG with params.preconditioner = CFD_B200_PRECOND_MULTIGRID; SURVEY 8f n4, the reference's roadmap
// item ROADMAP.md:190-194).  No counterpart in the reference: same system, same stopping rule
// (linear_solver_cg.c:290-461), 7-8 iterations where its Jacobi-PCG needs hundreds to thousands
// (tools/mgpcg_dirichlet_prototype.py), answers equal to the tolerance asked for.
//
// The arithmetic of every kernel is a loop over the functions of mgpc_cell.h, which oracle/mgpc_host.cpp runs on
// the host: that file is the CPU restatement the tests compare with.  The kernels are plain one-thread-per-cell
// sweeps (no TMA pipeline yet): a cycle moves ~5 fine-grid sweeps of 24-40 B/cell, an iteration ~10 passes over the
// grid -- against 2 passes x several hundred iterations for the reference's preconditioner.  One GPU only; the
// iteration is driven from the host (a handful of iterations, one small copy per dot product).
#include "mgpc_cell.h"
#include "session.h"

#include <math.h>

namespace cfdb {

namespace {

constexpr int MT = 256;  // threads per block

struct MgpcState {
    int nl = 0;
    MgpcLv lv[MGPC_MAX_LEVELS];
    double* x[MGPC_MAX_LEVELS] = {};  // levels >= 1 (level 0 lives in the session's Krylov vectors)
    double* b[MGPC_MAX_LEVELS] = {};
    double* t[MGPC_MAX_LEVELS] = {};
    int nx = 0, ny = 0, nz = 0;
    double dx = 0.0, dy = 0.0, dz = 0.0;
};

// interior cell q (0 .. n0*n1*n2) -> (i, j, k)
__device__ __forceinline__ void cell_of(const MgpcLv& L, long long q, int& i, int& j, int& k) {
    i = 1 + (int)(q % L.n[0]);
    const long long row = q / L.n[0];
    j = 1 + (int)(row % L.n[1]);
    k = L.act[2] ? 1 + (int)(row / L.n[1]) : 0;
}
__host__ __device__ __forceinline__ long long interior_cells(const MgpcLv& L) {
    return (long long)L.n[0] * L.n[1] * L.n[2];
}

__global__ void __launch_bounds__(MT) k_mgpc_jacobi0(MgpcLv L, double* __restrict__ x, const double* __restrict__ b) {
    const long long nc = interior_cells(L), stride = (long long)gridDim.x * MT;
    const double om = mgpc_omega(L);
    for (long long q = (long long)blockIdx.x * MT + threadIdx.x; q < nc; q += stride) {
        int i, j, k;
        cell_of(L, q, i, j, k);
        const long long c = mgpc_idx(L, i, j, k);
        x[c] = om * b[c] / mgpc_diag(L, i, j, k);
    }
}
// xn = x + omega D^-1 (b - K x)
__global__ void __launch_bounds__(MT) k_mgpc_jacobi(MgpcLv L, double* __restrict__ xn, const double* __restrict__ x,
                                                    const double* __restrict__ b) {
    const long long nc = interior_cells(L), stride = (long long)gridDim.x * MT;
    const double om = mgpc_omega(L);
    for (long long q = (long long)blockIdx.x * MT + threadIdx.x; q < nc; q += stride) {
        int i, j, k;
        cell_of(L, q, i, j, k);
        double dg;
        const double au = mgpc_apply(L, x, i, j, k, dg);
        const long long c = mgpc_idx(L, i, j, k);
        xn[c] = x[c] + om * (b[c] - au) / dg;
    }
}
__global__ void __launch_bounds__(MT) k_mgpc_residual(MgpcLv L, double* __restrict__ r, const double* __restrict__ x,
                                                      const double* __restrict__ b) {
    const long long nc = interior_cells(L), stride = (long long)gridDim.x * MT;
    for (long long q = (long long)blockIdx.x * MT + threadIdx.x; q < nc; q += stride) {
        int i, j, k;
        cell_of(L, q, i, j, k);
        double dg;
        const double au = mgpc_apply(L, x, i, j, k, dg);
        const long long c = mgpc_idx(L, i, j, k);
        r[c] = b[c] - au;
    }
}
__global__ void __launch_bounds__(MT) k_mgpc_restrict(MgpcLv Lf, MgpcLv Lc, double* __restrict__ bc,
                                                      const double* __restrict__ rf) {
    const long long nc = interior_cells(Lc), stride = (long long)gridDim.x * MT;
    for (long long q = (long long)blockIdx.x * MT + threadIdx.x; q < nc; q += stride) {
        int I, J, K;
        cell_of(Lc, q, I, J, K);
        bc[mgpc_idx(Lc, I, J, K)] = mgpc_restrict(Lf, rf, I, J, K);
    }
}
__global__ void __launch_bounds__(MT) k_mgpc_prolong_add(MgpcLv Lf, MgpcLv Lc, double* __restrict__ x,
                                                         const double* __restrict__ ec) {
    const long long nc = interior_cells(Lf), stride = (long long)gridDim.x * MT;
    for (long long q = (long long)blockIdx.x * MT + threadIdx.x; q < nc; q += stride) {
        int i, j, k;
        cell_of(Lf, q, i, j, k);
        x[mgpc_idx(Lf, i, j, k)] += mgpc_prolong(Lf, Lc, ec, i, j, k);
    }
}
// coarsest level (<= 27 unknowns, <= 125 stored cells): all sweeps in one block
__global__ void __launch_bounds__(128) k_mgpc_coarse(MgpcLv L, double* __restrict__ x, const double* __restrict__ b) {
    __shared__ double xs[2][128];
    const int t = threadIdx.x;
    const long long nc = interior_cells(L);
    xs[0][t] = xs[1][t] = 0.0;  // wall cells stay zero
    __syncthreads();
    int i = 0, j = 0, k = 0;
    long long c = 0;
    double bb = 0.0, dg = 1.0;
    const bool on = t < nc;
    if (on) {
        cell_of(L, t, i, j, k);
        c = mgpc_idx(L, i, j, k);
        bb = b[c];
        dg = mgpc_diag(L, i, j, k);
    }
    const double om = mgpc_omega(L);
    if (on) xs[0][c] = om * bb / dg;
    __syncthreads();
    int cur = 0;
    for (int s = 1; s < MGPC_COARSE_SWEEPS; s++) {
        if (on) {
            double d2;
            const double au = mgpc_apply(L, xs[cur], i, j, k, d2);
            xs[cur ^ 1][c] = xs[cur][c] + om * (bb - au) / dg;
        }
        __syncthreads();
        cur ^= 1;
    }
    if (on) x[c] = xs[cur][c];
}

// ---- the PCG around it, on the session's own layout (Geom), interior cells ----
__device__ __forceinline__ double ref_lap(const Geom& g, const double* __restrict__ u, long long c) {
    // linear_solver_cg.c:103-123 / 134-158: multiplications by the reciprocals, the z term last
    const double idx2 = 1.0 / g.dx2, idy2 = 1.0 / g.dy2;
    double s = ((u[c + 1] - (2.0 * u[c]) + u[c - 1]) * idx2) + ((u[c + g.sy] - (2.0 * u[c]) + u[c - g.sy]) * idy2);
    const double zp = g.dim3 ? u[c + g.sz] : u[c], zm = g.dim3 ? u[c - g.sz] : u[c];
    return s + ((zp + zm - (2.0 * u[c])) * g.inv_dz2);
}
__device__ __forceinline__ bool geom_cell(const Geom& g, long long q, long long& c) {
    const int nxi = g.nx - 2, nyi = g.ny - 2;
    const int i = 1 + (int)(q % nxi);
    const long long row = q / nxi;
    const int j = 1 + (int)(row % nyi), k = g.k0 + (int)(row / nyi);
    c = (long long)k * g.sz + (long long)j * g.sy + i;
    return k < g.k1;
}
__host__ __device__ __forceinline__ long long geom_cells(const Geom& g) {
    return (long long)(g.nx - 2) * (g.ny - 2) * (g.k1 - g.k0);
}

// r = -rhs + lap x ; dot[0] = r.r
__global__ void __launch_bounds__(MT) k_mgpc_init(Geom g, const double* __restrict__ x, const double* __restrict__ rhs,
                                                  double* __restrict__ r, Ctl* ctl, RedBuf rb) {
    const long long nc = geom_cells(g), stride = (long long)gridDim.x * MT;
    double acc[1] = {0.0};
    for (long long q = (long long)blockIdx.x * MT + threadIdx.x; q < nc; q += stride) {
        long long c;
        geom_cell(g, q, c);
        const double rv = -rhs[c] + ref_lap(g, x, c);
        r[c] = rv;
        acc[0] += rv * rv;
    }
    block_reduce_finish<1, false>(acc, rb, &ctl->dot[0]);
}
// Ap = -lap p ; dot[1] = p.Ap
__global__ void __launch_bounds__(MT) k_mgpc_matvec(Geom g, const double* __restrict__ p, double* __restrict__ Ap, Ctl* ctl,
                                                    RedBuf rb) {
    const long long nc = geom_cells(g), stride = (long long)gridDim.x * MT;
    double acc[1] = {0.0};
    for (long long q = (long long)blockIdx.x * MT + threadIdx.x; q < nc; q += stride) {
        long long c;
        geom_cell(g, q, c);
        const double av = -ref_lap(g, p, c);
        Ap[c] = av;
        acc[0] += p[c] * av;
    }
    block_reduce_finish<1, false>(acc, rb, &ctl->dot[1]);
}
// x += alpha p ; r -= alpha Ap ; dot[2] = r.r
__global__ void __launch_bounds__(MT) k_mgpc_update(Geom g, double* __restrict__ x, double* __restrict__ r,
                                                    const double* __restrict__ p, const double* __restrict__ Ap,
                                                    double alpha, Ctl* ctl, RedBuf rb) {
    const long long nc = geom_cells(g), stride = (long long)gridDim.x * MT;
    double acc[1] = {0.0};
    for (long long q = (long long)blockIdx.x * MT + threadIdx.x; q < nc; q += stride) {
        long long c;
        geom_cell(g, q, c);
        x[c] += alpha * p[c];
        const double rv = r[c] - alpha * Ap[c];
        r[c] = rv;
        acc[0] += rv * rv;
    }
    block_reduce_finish<1, false>(acc, rb, &ctl->dot[2]);
}
// dot[3] = a.b
__global__ void __launch_bounds__(MT) k_mgpc_dot(Geom g, const double* __restrict__ a, const double* __restrict__ b, Ctl* ctl,
                                                 RedBuf rb) {
    const long long nc = geom_cells(g), stride = (long long)gridDim.x * MT;
    double acc[1] = {0.0};
    for (long long q = (long long)blockIdx.x * MT + threadIdx.x; q < nc; q += stride) {
        long long c;
        geom_cell(g, q, c);
        acc[0] += a[c] * b[c];
    }
    block_reduce_finish<1, false>(acc, rb, &ctl->dot[3]);
}
// p = z + beta p (first: p = z)
__global__ void __launch_bounds__(MT) k_mgpc_dir(Geom g, double* __restrict__ p, const double* __restrict__ z, double beta,
                                                 int first) {
    const long long nc = geom_cells(g), stride = (long long)gridDim.x * MT;
    for (long long q = (long long)blockIdx.x * MT + threadIdx.x; q < nc; q += stride) {
        long long c;
        geom_cell(g, q, c);
        p[c] = first ? z[c] : z[c] + beta * p[c];
    }
}

int blocks_for(long long cells) {
    long long b = (cells + MT - 1) / MT;
    if (b < 1) b = 1;
    if (b > 148 * 8) b = 148 * 8;
    return (int)b;
}

void state_free(MgpcState* st) {
    if (!st) return;
    for (int l = 1; l < st->nl; l++) {
        if (st->x[l]) cudaFree(st->x[l]);
        if (st->b[l]) cudaFree(st->b[l]);
        if (st->t[l]) cudaFree(st->t[l]);
    }
    delete st;
}

cfd_status_t state_prepare(cfd_b200_session* s, MgpcState** out) {
    MgpcState* st = static_cast<MgpcState*>(s->mgpc);
    const cfd_b200_session_desc_t& d = s->desc;
    if (st && st->nx == (int)d.nx && st->ny == (int)d.ny && st->nz == (int)d.nz && st->dx == d.dx && st->dy == d.dy &&
        st->dz == d.dz) {
        *out = st;
        return CFD_SUCCESS;
    }
    state_free(st);
    s->mgpc = nullptr;
    st = new MgpcState();
    st->nx = (int)d.nx;
    st->ny = (int)d.ny;
    st->nz = (int)d.nz;
    st->dx = d.dx;
    st->dy = d.dy;
    st->dz = d.dz;
    st->nl = mgpc_plan(st->nx, st->ny, st->nz, d.dx, d.dy, d.nz > 1 ? d.dz : 0.0, st->lv);
    for (int l = 1; l < st->nl; l++) {
        const size_t bytes = (size_t)st->lv[l].cells * sizeof(double);
        double** arr[3] = {&st->x[l], &st->b[l], &st->t[l]};
        for (double** a : arr) {
            if (cudaMalloc((void**)a, bytes) != cudaSuccess) {
                cudaGetLastError();
                *a = nullptr;
                state_free(st);
                set_error(CFD_ERROR_NOMEM, "multigrid preconditioner: level %d (%zu bytes) could not be allocated", l, bytes);
                return CFD_ERROR_NOMEM;
            }
            CFDB_CUDA_OK(cudaMemsetAsync(*a, 0, bytes, s->st));  // wall cells are zero and never written
            s->bytes_allocated += bytes;
        }
    }
    if (interior_cells(st->lv[st->nl - 1]) > 128 || st->lv[st->nl - 1].cells > 128) {
        state_free(st);
        set_error(CFD_ERROR, "multigrid preconditioner: coarsest level too large");
        return CFD_ERROR;
    }
    s->mgpc = st;
    *out = st;
    return CFD_SUCCESS;
}

// z = M^-1 r: one V(1,1) cycle.  x0 / b0 / t0: level-0 arrays in the session's layout (b0 = r, x0 = z).
void vcycle(cfd_b200_session* s, MgpcState* st, int l, double* x0, const double* b0, double* t0) {
    const MgpcLv& L = st->lv[l];
    double* x = l ? st->x[l] : x0;
    const double* b = l ? st->b[l] : b0;
    double* t = l ? st->t[l] : t0;
    cudaStream_t q = s->st;
    if (l == st->nl - 1) {
        CFDB_LAUNCH(KK_MG_COARSE, q, k_mgpc_coarse<<<1, 128, 0, q>>>(L, x, b));
        return;
    }
    const MgpcLv& Lc = st->lv[l + 1];
    const int nb = blocks_for(interior_cells(L)), nbc = blocks_for(interior_cells(Lc));
    CFDB_LAUNCH(KK_MG_SMOOTH, q, k_mgpc_jacobi0<<<nb, MT, 0, q>>>(L, x, b));
    CFDB_LAUNCH(KK_MG_RESIDUAL, q, k_mgpc_residual<<<nb, MT, 0, q>>>(L, t, x, b));
    CFDB_LAUNCH(KK_MG_RESTRICT, q, k_mgpc_restrict<<<nbc, MT, 0, q>>>(L, Lc, st->b[l + 1], t));
    vcycle(s, st, l + 1, x0, b0, t0);
    CFDB_LAUNCH(KK_MG_PROLONG, q, k_mgpc_prolong_add<<<nb, MT, 0, q>>>(L, Lc, x, st->x[l + 1]));
    CFDB_LAUNCH(KK_MG_SMOOTH, q, k_mgpc_jacobi<<<nb, MT, 0, q>>>(L, t, x, b));
    cudaMemcpyAsync(x, t, (size_t)L.cells * sizeof(double), cudaMemcpyDeviceToDevice, q);
}

}  // namespace

void mgpc_release(cfd_b200_session* s) {
    state_free(static_cast<MgpcState*>(s->mgpc));
    s->mgpc = nullptr;
}

cfd_status_t mgpcg_solve(cfd_b200_session* s, const cfd_b200_poisson_config_t* cfg, double* x, const double* rhs,
                         poisson_solver_stats_t* out) {
    if (s->desc.nranks > 1) {
        set_error(CFD_ERROR_UNSUPPORTED, "the multigrid preconditioner runs on one GPU only");
        return CFD_ERROR_UNSUPPORTED;
    }
    MgpcState* st = nullptr;
    cfd_status_t rc = state_prepare(s, &st);
    if (rc != CFD_SUCCESS) return rc;
    for (int q = 0; q < 5; q++)
        if ((rc = dalloc(s, &s->kv[q])) != CFD_SUCCESS) return rc;
    double *r = s->kv[0], *p = s->kv[1], *Ap = s->kv[2], *z = s->kv[3], *t0 = s->kv[4];
    const poisson_solver_params_t& pp = cfg->params;
    const LaunchCtx L = s->L();
    const Geom& g = s->g;
    cudaStream_t q = s->st;
    const int nb = blocks_for(geom_cells(g));
    const int check = pp.check_interval > 0 ? pp.check_interval : 1;
    // the Krylov vectors of an earlier solve may hold anything on interior cells; their boundary cells are zero by
    // construction (dalloc zero-fills, kernels write interior cells only) -- z and t0 are rewritten by every cycle
    double h[4] = {0.0, 0.0, 0.0, 0.0};
    auto fetch = [&](int slot) -> cfd_status_t {
        CFDB_CUDA_OK(cudaMemcpyAsync(&h[slot], &s->ctl->dot[slot], sizeof(double), cudaMemcpyDeviceToHost, q));
        CFDB_CUDA_OK(cudaStreamSynchronize(q));
        return CFD_SUCCESS;
    };
    CFDB_KL(s, launch_zero_ctl)(L);
    CFDB_KL(s, launch_neumann_fill)(L, x, 0, 0, false);
    CFDB_LAUNCH(KK_KRYLOV_INIT, q, k_mgpc_init<<<nb, MT, 0, q>>>(g, x, rhs, r, s->ctl, s->rb));
    if ((rc = fetch(0)) != CFD_SUCCESS) return rc;
    const double r0 = sqrt(h[0]);
    double tol = pp.tolerance * r0;
    if (tol < pp.absolute_tolerance) tol = pp.absolute_tolerance;
    poisson_solver_stats_t stt;
    memset(&stt, 0, sizeof(stt));
    stt.initial_residual = stt.final_residual = r0;
    stt.status = POISSON_MAX_ITER;
    double res = r0;
    int iters = 0;
    bool converged = false, stagnated = false;
    if (!(r0 == r0)) {  // NaN in the input
        stt.status = POISSON_DIVERGED;
    } else if (r0 < pp.absolute_tolerance) {
        converged = true;
    } else {
        vcycle(s, st, 0, z, r, t0);
        CFDB_LAUNCH(KK_CG_DIR, q, k_mgpc_dir<<<nb, MT, 0, q>>>(g, p, z, 0.0, 1));
        CFDB_LAUNCH(KK_CG_UPDATE, q, k_mgpc_dot<<<nb, MT, 0, q>>>(g, r, z, s->ctl, s->rb));
        if ((rc = fetch(3)) != CFD_SUCCESS) return rc;
        double rz = h[3];
        int it = 0;
        for (; it < pp.max_iterations; it++) {
            CFDB_LAUNCH(KK_CG_DIR, q, k_mgpc_matvec<<<nb, MT, 0, q>>>(g, p, Ap, s->ctl, s->rb));
            if ((rc = fetch(1)) != CFD_SUCCESS) return rc;
            if (fabs(h[1]) < 1e-30) {
                stagnated = true;
                break;
            }
            const double alpha = rz / h[1];
            CFDB_LAUNCH(KK_CG_UPDATE, q, k_mgpc_update<<<nb, MT, 0, q>>>(g, x, r, p, Ap, alpha, s->ctl, s->rb));
            if ((rc = fetch(2)) != CFD_SUCCESS) return rc;
            res = sqrt(h[2]);
            if (!(res == res)) break;
            if (it % check == 0 && (res < tol || res < pp.absolute_tolerance)) {
                converged = true;
                break;
            }
            vcycle(s, st, 0, z, r, t0);
            CFDB_LAUNCH(KK_CG_UPDATE, q, k_mgpc_dot<<<nb, MT, 0, q>>>(g, r, z, s->ctl, s->rb));
            if ((rc = fetch(3)) != CFD_SUCCESS) return rc;
            if (fabs(rz) < 1e-30) {
                stagnated = true;
                break;
            }
            const double beta = h[3] / rz;
            CFDB_LAUNCH(KK_CG_DIR, q, k_mgpc_dir<<<nb, MT, 0, q>>>(g, p, z, beta, 0));
            rz = h[3];
        }
        iters = it < pp.max_iterations ? it + 1 : it;
        if (!converged && !stagnated && (res < tol || res < pp.absolute_tolerance)) converged = true;
    }
    if (!stagnated) CFDB_KL(s, launch_neumann_fill)(L, x, 0, 0, false);
    CFDB_CUDA_OK(cudaStreamSynchronize(q));
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        set_error(CFD_ERROR, "multigrid-preconditioned CG: %s", cudaGetErrorString(e));
        return CFD_ERROR;
    }
    stt.iterations = iters;
    stt.final_residual = res;
    if (stt.status != POISSON_DIVERGED) stt.status = converged ? POISSON_CONVERGED : (stagnated ? POISSON_STAGNATED : POISSON_MAX_ITER);
    if (!(res == res)) stt.status = POISSON_DIVERGED;
    s->last_ps = stt;
    if (out) *out = stt;
    return stt.status == POISSON_CONVERGED ? CFD_SUCCESS : CFD_ERROR_MAX_ITER;  // (poisson_solve's convention)
}

}  // namespace cfdb
