// capi.cu -- the C-ABI: drop-in symbols with the reference's names and signatures
// (include/cfd_b200.h Part 2) and the cfd_b200_* extension surface (Part 3).
#include "runner.h"

#include <math.h>
#include <mutex>
#include <stdlib.h>
#include <string.h>

using namespace cfdb;

namespace {

std::mutex g_mu;
cfd_b200_poisson_config_t g_pcfg;
bool g_pcfg_init = false;

poisson_solver_params_t default_pparams() {  // linear_solver.c:37-47
    poisson_solver_params_t p;
    memset(&p, 0, sizeof(p));
    p.tolerance = 1e-6;
    p.absolute_tolerance = 1e-10;
    p.max_iterations = 5000;
    p.omega = 0.0;
    p.check_interval = 1;
    p.verbose = false;
    p.preconditioner = POISSON_PRECOND_NONE;
    return p;
}

void init_pcfg_locked() {
    if (g_pcfg_init) return;
    g_pcfg.method = POISSON_METHOD_CG;  // solver_projection.c:217-218
    g_pcfg.params = default_pparams();
    g_pcfg.on_nonconvergence = CFD_B200_NONCONV_ERROR;
    const char* e = getenv("CFD_B200_POISSON");
    if (e) {
        if (!strcmp(e, "jacobi")) { g_pcfg.method = POISSON_METHOD_JACOBI; g_pcfg.params.max_iterations = 2000; }
        else if (!strcmp(e, "rbsor")) g_pcfg.method = POISSON_METHOD_REDBLACK_SOR;
        else if (!strcmp(e, "pcg")) g_pcfg.params.preconditioner = POISSON_PRECOND_JACOBI;
        else if (!strcmp(e, "mgpcg")) g_pcfg.params.preconditioner = CFD_B200_PRECOND_MULTIGRID;
        else if (!strcmp(e, "bicgstab")) g_pcfg.method = POISSON_METHOD_BICGSTAB;
        else if (!strcmp(e, "multigrid")) { g_pcfg.method = POISSON_METHOD_MULTIGRID; g_pcfg.params.max_iterations = 100; }
    }
    const char* c = getenv("CFD_B200_NONCONV");
    if (c && !strcmp(c, "continue")) g_pcfg.on_nonconvergence = CFD_B200_NONCONV_CONTINUE;
    g_pcfg_init = true;
}

}  // namespace

namespace cfdb {
// Argument checks shared by the steppers (solver_projection.c:48-66, energy_solver.c:55-91).
cfd_status_t validate_common(const flow_field* f, const grid* g, const ns_solver_params_t* prm) {
    if (!f || !g || !prm) return CFD_ERROR_INVALID;
    if (f->nx < 3 || f->ny < 3 || (f->nz > 1 && f->nz < 3) || f->nz == 0) return CFD_ERROR_INVALID;
    if (!f->u || !f->v || !f->w || !f->p || !f->rho || !f->T || !g->dx || !g->dy || !g->x || !g->y)
        return CFD_ERROR_INVALID;
    if (f->nz > 1 && !g->dz) return CFD_ERROR_INVALID;
    if (f->nz > 1)
        for (size_t k = 1; k < f->nz - 1; k++)
            if (fabs(g->dz[k] - g->dz[0]) > 1e-14) return CFD_ERROR_INVALID;
    const double tx = 1e-12 * fmax(1.0, fabs(g->dx[0])), ty = 1e-12 * fmax(1.0, fabs(g->dy[0]));
    for (size_t i = 1; i < f->nx - 1; i++)
        if (fabs(g->dx[i] - g->dx[0]) > tx) {
            set_error(CFD_ERROR_UNSUPPORTED, "non-uniform dx is not supported by the GPU backend");
            return CFD_ERROR_UNSUPPORTED;
        }
    for (size_t j = 1; j < f->ny - 1; j++)
        if (fabs(g->dy[j] - g->dy[0]) > ty) {
            set_error(CFD_ERROR_UNSUPPORTED, "non-uniform dy is not supported by the GPU backend");
            return CFD_ERROR_UNSUPPORTED;
        }
    if (prm->source_func || prm->heat_source_func) {
        set_error(CFD_ERROR_UNSUPPORTED, "host source/heat-source callbacks cannot run on the GPU");
        return CFD_ERROR_UNSUPPORTED;
    }
    return CFD_SUCCESS;
}
}  // namespace cfdb

namespace {

cfd_status_t dropin(StepKind kind, flow_field* f, const grid* g, const ns_solver_params_t* prm,
                    const gpu_config_t* config) {
    if (!f || !g || !prm) return CFD_ERROR_INVALID;
    gpu_config_t cfg = config ? *config : gpu_config_default();
    if (!gpu_should_use(&cfg, f->nx, f->ny, f->nz, prm->max_iter)) return CFD_ERROR;
    cfd_status_t rc = validate_common(f, g, prm);
    if (rc != CFD_SUCCESS) return rc;
    // one runner per (shape, spacing, device, GPU count), used by one call at a time; different shapes run
    // concurrently on their own runners (runner_cache_acquire)
    Runner* r = runner_cache_acquire(f, g);
    if (!r) return CFD_ERROR_NOMEM;
    {
        std::lock_guard<std::mutex> lk(r->mu);
        rc = runner_run(r, kind, f, g, prm, prm->max_iter, nullptr, SYNC_FULL);
        if (kind != SK_PROJECTION && (rc == CFD_SUCCESS || rc == CFD_ERROR_DIVERGED) && prm->max_iter > 0) wrap_host_rho(f);
    }
    runner_cache_release(r);
    return rc;
}

}  // namespace

extern "C" {

// ---- device management (gpu_device.h:97-130) ----
gpu_config_t gpu_config_default(void) {
    gpu_config_t c;
    memset(&c, 0, sizeof(c));
    c.enable_gpu = 1;
    c.min_grid_size = 10000;
    c.min_steps = 10;
    c.block_size_x = 16;
    c.block_size_y = 16;
    c.poisson_max_iter = 1000;
    c.poisson_tolerance = 1e-3;
    c.persistent_memory = 1;
    c.async_transfers = 1;
    return c;
}
int gpu_is_available(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) == cudaSuccess && n > 0) return 1;
    cudaGetLastError();
    return 0;
}
int gpu_get_device_info(gpu_device_info_t* info, int max_devices) {
    int n = 0;
    if (!info || cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    if (n > max_devices) n = max_devices;
    int cur = 0;
    cudaGetDevice(&cur);
    for (int i = 0; i < n; i++) {
        cudaDeviceProp p;
        memset(&info[i], 0, sizeof(info[i]));
        if (cudaGetDeviceProperties(&p, i) != cudaSuccess) continue;
        info[i].device_id = i;
        snprintf(info[i].name, sizeof(info[i].name), "%s", p.name);
        info[i].total_memory = p.totalGlobalMem;
        info[i].compute_capability_major = p.major;
        info[i].compute_capability_minor = p.minor;
        info[i].multiprocessor_count = p.multiProcessorCount;
        info[i].max_threads_per_block = p.maxThreadsPerBlock;
        info[i].warp_size = p.warpSize;
        info[i].is_available = 1;
        size_t fr = 0, tot = 0;
        cudaSetDevice(i);
        info[i].free_memory = cudaMemGetInfo(&fr, &tot) == cudaSuccess ? fr : 0;
    }
    cudaSetDevice(cur);
    return n;
}
cfd_status_t gpu_select_device(int device_id) {
    return cudaSetDevice(device_id) == cudaSuccess ? CFD_SUCCESS : CFD_ERROR;
}
int gpu_should_use(const gpu_config_t* c, size_t nx, size_t ny, size_t nz, int num_steps) {
    if (!c || !c->enable_gpu || !gpu_is_available()) return 0;
    if (nx < 3 || ny < 3 || nz == 0 || nz == 2) return 0;
    if (ny > SIZE_MAX / nx || nz > SIZE_MAX / (nx * ny)) return 0;
    if (nx * ny * nz < c->min_grid_size) return 0;
    if (num_steps < c->min_steps) return 0;
    return 1;
}

// ---- Poisson configuration ----
cfd_b200_poisson_config_t cfd_b200_poisson_config_default(void) {
    cfd_b200_poisson_config_t c;
    c.method = POISSON_METHOD_CG;
    c.params = default_pparams();
    c.on_nonconvergence = CFD_B200_NONCONV_ERROR;
    return c;
}
cfd_status_t cfd_b200_set_poisson_config(const cfd_b200_poisson_config_t* cfg) {
    if (!cfg) return CFD_ERROR_INVALID;
    if (cfg->method != POISSON_METHOD_JACOBI && cfg->method != POISSON_METHOD_REDBLACK_SOR &&
        cfg->method != POISSON_METHOD_CG && cfg->method != POISSON_METHOD_BICGSTAB &&
        cfg->method != POISSON_METHOD_MULTIGRID)
        return CFD_ERROR_UNSUPPORTED;
    std::lock_guard<std::mutex> lk(g_mu);
    g_pcfg = *cfg;
    g_pcfg_init = true;
    return CFD_SUCCESS;
}
cfd_b200_poisson_config_t cfd_b200_get_poisson_config(void) {
    std::lock_guard<std::mutex> lk(g_mu);
    init_pcfg_locked();
    return g_pcfg;
}

// ---- stateless steppers (gpu_device.h:209-249, solver_registry.c:86-88) ----
cfd_status_t solve_projection_method_gpu(flow_field* f, const grid* g, const ns_solver_params_t* p,
                                         const gpu_config_t* c) {
    return dropin(SK_PROJECTION, f, g, p, c);
}
cfd_status_t solve_navier_stokes_gpu(flow_field* f, const grid* g, const ns_solver_params_t* p,
                                     const gpu_config_t* c) {
    return dropin(SK_PROJECTION, f, g, p, c);
}
cfd_status_t solve_rk2_method_gpu(flow_field* f, const grid* g, const ns_solver_params_t* p,
                                  const gpu_config_t* c) {
    return dropin(SK_RK2, f, g, p, c);
}
cfd_status_t solve_rk4_method_gpu(flow_field* f, const grid* g, const ns_solver_params_t* p,
                                  const gpu_config_t* c) {
    return dropin(SK_RK4, f, g, p, c);
}
cfd_status_t solve_explicit_euler_method_gpu(flow_field* f, const grid* g, const ns_solver_params_t* p,
                                             const gpu_config_t* c) {
    return dropin(SK_EULER, f, g, p, c);
}

// ---- session steppers ----
cfd_status_t cfd_b200_session_projection(cfd_b200_session_t* s, const ns_solver_params_t* prm,
                                         const cfd_b200_poisson_config_t* pc, int nsteps,
                                         ns_solver_stats_t* stats) {
    cfd_b200_poisson_config_t d = cfd_b200_get_poisson_config();
    return run_projection(s, prm, pc ? pc : &d, nsteps, stats);
}
cfd_status_t cfd_b200_session_rk(cfd_b200_session_t* s, int order, const ns_solver_params_t* prm,
                                 int nsteps, ns_solver_stats_t* stats) {
    return run_rk(s, order, prm, nsteps, stats);
}
cfd_status_t cfd_b200_session_poisson(cfd_b200_session_t* s, const cfd_b200_poisson_config_t* pc,
                                      poisson_solver_stats_t* stats) {
    if (!s || !pc) return CFD_ERROR_INVALID;
    CFDB_CUDA_OK(cudaSetDevice(s->dev));
    t_prof = &s->prof;
    cfd_status_t rc = dalloc(s, &s->rhs);
    if (rc != CFD_SUCCESS) return rc;
    if (pc->method == POISSON_METHOD_JACOBI && (rc = dalloc(s, &s->xt)) != CFD_SUCCESS) return rc;
    return poisson_solve(s, pc, &s->p, &s->xt, s->rhs, stats);
}

#define CFD_B200_ABI_FN cfd_b200_abi_table
#include "abi_table.inc"

}  // extern "C"
