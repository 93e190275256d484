// capi_vtables.cu -- factories returning the reference's vtable objects:
//   poisson_solver_t  (create_{jacobi,redblack,cg,bicgstab}_gpu_solver; linear_solver.c:165-229)
//   ns_solver_t       (run-time registration path; navier_stokes_solver.h:254-277, 333-337)
// and the legacy persistent-context API gpu_solver_* (gpu_device.h:145-194).
// Objects are calloc'ed because the reference frees them with cfd_free == free
// (solver_registry.c:414-423, linear_solver.c:288-298, core/memory.c).
#include "runner.h"

#include <stdlib.h>
#include <string.h>
#include <sys/time.h>
#include <thread>
#include <vector>

using namespace cfdb;

namespace {

double now_ms() {
    struct timeval tv;
    gettimeofday(&tv, nullptr);
    return tv.tv_sec * 1000.0 + tv.tv_usec / 1000.0;
}

// ---------------- Poisson vtable ----------------
cfd_status_t ps_init(poisson_solver_t* sv, size_t nx, size_t ny, size_t nz, double dx, double dy, double dz,
                     const poisson_solver_params_t* params) {
    (void)params;
    if (!gpu_is_available()) {
        set_error(CFD_ERROR_UNSUPPORTED, "CUDA GPU not available at runtime");
        return CFD_ERROR_UNSUPPORTED;
    }
    cfd_b200_session_desc_t d;
    memset(&d, 0, sizeof(d));
    d.nx = nx; d.ny = ny; d.nz = nz ? nz : 1;
    d.dx = dx; d.dy = dy; d.dz = dz;
    d.nranks = 1; d.device = -1;
    cfd_b200_session* s = cfd_b200_session_create(&d);
    if (!s) return CFD_ERROR_NOMEM;
    sv->context = s;
    return CFD_SUCCESS;
}
void ps_destroy(poisson_solver_t* sv) {
    if (sv && sv->context) {
        cfd_b200_session_destroy((cfd_b200_session*)sv->context);
        sv->context = nullptr;
    }
}
cfd_status_t ps_solve(poisson_solver_t* sv, double* x, double* x_temp, const double* rhs,
                      poisson_solver_stats_t* stats) {
    (void)x_temp;
    if (!sv || !sv->context || !x || !rhs) return CFD_ERROR_INVALID;
    cfd_b200_session* s = (cfd_b200_session*)sv->context;
    const double t0 = now_ms();
    cfd_status_t rc = cfd_b200_session_upload_poisson(s, x, rhs);
    if (rc != CFD_SUCCESS) return rc;
    cfd_b200_poisson_config_t pc;
    pc.method = sv->method;
    pc.params = sv->params;
    pc.on_nonconvergence = CFD_B200_NONCONV_ERROR;
    poisson_solver_stats_t st;
    rc = cfd_b200_session_poisson(s, &pc, &st);
    if (rc == CFD_SUCCESS || rc == CFD_ERROR_MAX_ITER) {
        cfd_status_t r2 = cfd_b200_session_download_poisson(s, x);
        if (r2 != CFD_SUCCESS) return r2;
    }
    st.elapsed_time_ms = now_ms() - t0;
    if (stats) *stats = st;
    return rc;
}

poisson_solver_t* make_ps(const char* name, const char* desc, poisson_solver_method_t m, int max_it) {
    poisson_solver_t* sv = (poisson_solver_t*)calloc(1, sizeof(poisson_solver_t));
    if (!sv) return nullptr;
    sv->name = name;
    sv->description = desc;
    sv->method = m;
    sv->backend = POISSON_BACKEND_GPU;
    sv->params.tolerance = 1e-6;
    sv->params.absolute_tolerance = 1e-10;
    sv->params.max_iterations = max_it;
    sv->params.omega = 0.0;
    sv->params.check_interval = 1;
    sv->params.verbose = false;
    sv->params.preconditioner = POISSON_PRECOND_NONE;
    sv->init = ps_init;
    sv->destroy = ps_destroy;
    sv->solve = ps_solve;
    sv->iterate = nullptr;   // as the reference's GPU solvers
    sv->apply_bc = nullptr;  // default Neumann
    return sv;
}

// ---------------- ns_solver_t vtable (persistent context) ----------------
struct NsCtx {
    Runner* r;     // one session, or a gang of slabs when cfd_b200_set_device_count / CFD_B200_NGPU > 1
    int kind;      // StepKind: 0 projection, 1 explicit Euler, 2 rk2, 4 rk4
    int sync;      // cfd_b200_host_sync_t bits (cfd_b200_solver_set_host_sync)
    unsigned magic;
};
constexpr unsigned NS_MAGIC = 0xCFDB2001u;

// compute_max_temperature (solver_registry.c:52-62) over the caller's array; a few threads for big fields
// (1024^3 doubles take ~1 s on one core -- as long as the projection step itself)
double host_max(const double* a, size_t n) {
    unsigned hc = std::thread::hardware_concurrency();
    size_t nt = n < (1u << 22) ? 1 : (hc >= 32 ? 16 : (hc >= 4 ? hc / 2 : 1));
    std::vector<double> part(nt, a[0]);
    auto work = [&](size_t t) {
        double m = a[0];
        for (size_t q = n * t / nt, e = n * (t + 1) / nt; q < e; q++)
            if (a[q] > m) m = a[q];
        part[t] = m;
    };
    std::vector<std::thread> th;
    for (size_t t = 1; t < nt; t++) th.emplace_back(work, t);
    work(0);
    for (std::thread& x : th) x.join();
    double m = part[0];
    for (size_t t = 1; t < nt; t++)
        if (part[t] > m) m = part[t];
    return m;
}

cfd_status_t ns_ensure(ns_solver_t* slv, const flow_field* f, const grid* g) {
    NsCtx* c = (NsCtx*)slv->context;
    int want = configured_device_count();
    if (c->kind == SK_PROJECTION && cfd_b200_get_poisson_config().method == POISSON_METHOD_MULTIGRID) want = 1;
    if (c->r && !runner_matches(c->r, f, g, want)) {
        runner_destroy(c->r);
        c->r = nullptr;
    }
    if (!c->r) {
        c->r = runner_create(f, g, want);
        if (!c->r) return CFD_ERROR_NOMEM;
    }
    return CFD_SUCCESS;
}

cfd_status_t ns_run(ns_solver_t* slv, flow_field* f, const grid* g, const ns_solver_params_t* prm,
                    ns_solver_stats_t* stats, int nsteps) {
    if (!slv || !slv->context) return CFD_ERROR_INVALID;
    cfd_status_t rc = validate_common(f, g, prm);
    if (rc != CFD_SUCCESS) return rc;
    if ((rc = ns_ensure(slv, f, g)) != CFD_SUCCESS) return rc;
    NsCtx* c = (NsCtx*)slv->context;
    const bool thermo = (prm->alpha > 0.0 || prm->beta != 0.0);
    rc = runner_run(c->r, (StepKind)c->kind, f, g, prm, nsteps, stats, c->sync);
    if (c->kind == SK_PROJECTION) {
        if (stats && rc == CFD_SUCCESS && !thermo)  // T never went to the device: max over the host copy
            stats->max_temperature = host_max(f->T, f->nx * f->ny * f->nz);
    } else if ((rc == CFD_SUCCESS || rc == CFD_ERROR_DIVERGED) && nsteps > 0 && !c->r->host_stale) {
        wrap_host_rho(f);
    }
    if (stats) stats->status = rc;
    return rc;
}

cfd_status_t ns_init(ns_solver_t* slv, const grid* g, const ns_solver_params_t* prm) {
    (void)prm;
    if (!slv || !g) return CFD_ERROR_INVALID;
    if (!gpu_is_available()) {
        set_error(CFD_ERROR_UNSUPPORTED, "CUDA GPU not available at runtime");
        return CFD_ERROR_UNSUPPORTED;
    }
    return CFD_SUCCESS;  // the device context is created on first step (needs the field extents)
}
void ns_destroy(ns_solver_t* slv) {
    if (!slv || !slv->context) return;
    NsCtx* c = (NsCtx*)slv->context;
    if (c->r) runner_destroy(c->r);
    c->magic = 0;
    free(c);
    slv->context = nullptr;
}
cfd_status_t ns_step(ns_solver_t* slv, flow_field* f, const grid* g, const ns_solver_params_t* prm,
                     ns_solver_stats_t* stats) {
    const double t0 = now_ms();
    cfd_status_t rc = ns_run(slv, f, g, prm, stats, 1);  // step == exactly one dt (solver_registry.c:1128)
    if (stats) stats->elapsed_time_ms = now_ms() - t0;
    return rc;
}
cfd_status_t ns_solve(ns_solver_t* slv, flow_field* f, const grid* g, const ns_solver_params_t* prm,
                      ns_solver_stats_t* stats) {
    const double t0 = now_ms();
    cfd_status_t rc = ns_run(slv, f, g, prm, stats, prm ? prm->max_iter : 0);
    if (stats) stats->elapsed_time_ms = now_ms() - t0;
    return rc;
}

ns_solver_t* make_ns(const char* name, const char* desc, int kind) {
    if (!gpu_is_available()) {
        set_error(CFD_ERROR_UNSUPPORTED, "CUDA GPU not available at runtime");
        return nullptr;
    }
    ns_solver_t* s = (ns_solver_t*)calloc(1, sizeof(ns_solver_t));
    NsCtx* c = (NsCtx*)calloc(1, sizeof(NsCtx));
    if (!s || !c) { free(s); free(c); return nullptr; }
    c->kind = kind;
    c->magic = NS_MAGIC;
    c->sync = SYNC_FULL;
    s->name = name;
    s->description = desc;
    s->version = "cfd_b200-0.1.0";
    s->capabilities = (ns_solver_capabilities_t)(NS_SOLVER_CAP_INCOMPRESSIBLE | NS_SOLVER_CAP_TRANSIENT | NS_SOLVER_CAP_GPU);
    s->backend = NS_SOLVER_BACKEND_CUDA;
    s->context = c;
    s->init = ns_init;
    s->destroy = ns_destroy;
    s->step = ns_step;
    s->solve = ns_solve;
    return s;
}

}  // namespace

extern "C" {

poisson_solver_t* create_jacobi_gpu_solver(void) {
    return make_ps("jacobi_gpu", "Jacobi iteration (B200, fused residual)", POISSON_METHOD_JACOBI, 2000);
}
poisson_solver_t* create_redblack_gpu_solver(void) {
    return make_ps("redblack_gpu", "Red-Black SOR (B200)", POISSON_METHOD_REDBLACK_SOR, 5000);
}
poisson_solver_t* create_cg_gpu_solver(void) {
    return make_ps("cg_gpu", "Conjugate Gradient / Jacobi-PCG (B200, device-resident scalars)", POISSON_METHOD_CG, 5000);
}
poisson_solver_t* create_bicgstab_gpu_solver(void) {
    return make_ps("bicgstab_gpu", "BiCGSTAB (B200, device-resident scalars)", POISSON_METHOD_BICGSTAB, 5000);
}

// no counterpart in the reference (POISSON_METHOD_MULTIGRID is a placeholder there): same vtable, V-cycles as iterations
poisson_solver_t* cfd_b200_create_multigrid_gpu_solver(void) {
    return make_ps("multigrid_gpu", "Geometric multigrid, V(2,2) Red-Black Gauss-Seidel (B200)", POISSON_METHOD_MULTIGRID, 100);
}

ns_solver_t* cfd_b200_create_projection_gpu_solver(void) {
    return make_ns("projection_gpu", "B200-native projection method (predictor / Poisson / corrector on device)", 0);
}
ns_solver_t* cfd_b200_create_explicit_euler_gpu_solver(void) {
    return make_ns("explicit_euler_gpu", "B200-native explicit Euler stepper (CPU-exact limiters)", 1);
}
ns_solver_t* cfd_b200_create_rk2_gpu_solver(void) {
    return make_ns("rk2_gpu", "B200-native RK2 (Heun) explicit stepper", 2);
}
ns_solver_t* cfd_b200_create_rk4_gpu_solver(void) {
    return make_ns("rk4_gpu", "B200-native classical RK4 explicit stepper", 4);
}

// ---- host-sync policy of the ns_solver_t objects made above ----
static NsCtx* ns_ctx_of(ns_solver_t* slv) {
    if (!slv || !slv->context || slv->step != ns_step) return nullptr;
    NsCtx* c = (NsCtx*)slv->context;
    return c->magic == NS_MAGIC ? c : nullptr;
}
cfd_status_t cfd_b200_solver_set_host_sync(ns_solver_t* slv, int flags) {
    NsCtx* c = ns_ctx_of(slv);
    if (!c || (flags & ~(SYNC_UPLOAD_BOUNDARY | SYNC_DOWNLOAD_LAZY))) {
        set_error(CFD_ERROR_INVALID, "cfd_b200_solver_set_host_sync: not a cfd_b200 solver object, or unknown flag bits");
        return CFD_ERROR_INVALID;
    }
    if (c->r && c->r->host_stale && !(flags & SYNC_DOWNLOAD_LAZY)) {
        set_error(CFD_ERROR_INVALID, "the device is ahead of the host copy: call cfd_b200_solver_sync_host() before "
                                     "leaving CFD_B200_SYNC_DOWNLOAD_LAZY");
        return CFD_ERROR_INVALID;
    }
    c->sync = flags;
    return CFD_SUCCESS;
}
cfd_status_t cfd_b200_solver_sync_host(ns_solver_t* slv, flow_field* f) {
    NsCtx* c = ns_ctx_of(slv);
    if (!c || !f) return CFD_ERROR_INVALID;
    if (!c->r || !c->r->host_stale) return CFD_SUCCESS;  // nothing newer on the device
    if (f->nx != c->r->nx || f->ny != c->r->ny || f->nz != c->r->nz) return CFD_ERROR_INVALID;
    int mask = (c->r->resident_mask & (FM_U | FM_V | FM_W | FM_P | FM_T));
    if (c->kind != SK_PROJECTION) mask = FM_U | FM_V | FM_W | FM_P | FM_T;
    cfd_status_t rc = runner_download(c->r, f, mask);
    if (rc == CFD_SUCCESS && c->kind != SK_PROJECTION) wrap_host_rho(f);
    return rc;
}

// ---- legacy persistent context (gpu_device.h:145-194) ----
struct gpu_solver_context_t {
    cfd_b200_session* s;
    gpu_config_t cfg;
    gpu_solver_stats_t stats;
    size_t nx, ny, nz;
};

gpu_solver_context_t* gpu_solver_create(size_t nx, size_t ny, size_t nz, const gpu_config_t* config) {
    if (nx < 3 || ny < 3 || nz == 0 || nz == 2 || !gpu_is_available()) return nullptr;
    gpu_solver_context_t* c = (gpu_solver_context_t*)calloc(1, sizeof(*c));
    if (!c) return nullptr;
    c->cfg = config ? *config : gpu_config_default();
    c->nx = nx; c->ny = ny; c->nz = nz;
    return c;  // the session needs the grid spacing: created on first upload/step
}
void gpu_solver_destroy(gpu_solver_context_t* c) {
    if (!c) return;
    if (c->s) cfd_b200_session_destroy(c->s);
    free(c);
}
static cfd_status_t legacy_session(gpu_solver_context_t* c, double dx, double dy, double dz) {
    if (c->s) return CFD_SUCCESS;
    cfd_b200_session_desc_t d;
    memset(&d, 0, sizeof(d));
    d.nx = c->nx; d.ny = c->ny; d.nz = c->nz;
    d.dx = dx; d.dy = dy; d.dz = dz;
    d.nranks = 1; d.device = -1;
    c->s = cfd_b200_session_create(&d);
    if (!c->s) return CFD_ERROR_NOMEM;
    c->stats.memory_allocated = c->s->bytes_allocated;
    return CFD_SUCCESS;
}
cfd_status_t gpu_solver_upload(gpu_solver_context_t* c, const flow_field* f) {
    if (!c || !f) return CFD_ERROR_INVALID;
    if (f->nx != c->nx || f->ny != c->ny || f->nz != c->nz) return CFD_ERROR_INVALID;
    // spacing is not known yet: a unit-spacing placeholder session is replaced at step time
    cfd_status_t rc = legacy_session(c, 1.0, 1.0, c->nz > 1 ? 1.0 : 0.0);
    if (rc != CFD_SUCCESS) return rc;
    const double t0 = now_ms();
    rc = session_upload(c->s, f, FM_ALL);
    c->stats.transfer_time_ms += now_ms() - t0;
    return rc;
}
cfd_status_t gpu_solver_download(gpu_solver_context_t* c, flow_field* f) {
    if (!c || !f || !c->s) return CFD_ERROR_INVALID;
    const double t0 = now_ms();
    cfd_status_t rc = session_download(c->s, f, FM_U | FM_V | FM_W | FM_P | FM_T);
    c->stats.transfer_time_ms += now_ms() - t0;
    return rc;
}
cfd_status_t gpu_solver_step(gpu_solver_context_t* c, const grid* g, const ns_solver_params_t* prm,
                             gpu_solver_stats_t* stats) {
    if (!c || !g || !prm || !c->s) return CFD_ERROR_INVALID;
    cfd_b200_session* s = c->s;
    const double dx = g->dx[0], dy = g->dy[0], dz = (c->nz > 1 && g->dz) ? g->dz[0] : 0.0;
    if (s->desc.dx != dx || s->desc.dy != dy || s->desc.dz != dz) {
        // adopt the real spacing; device fields stay as uploaded
        s->desc.dx = dx; s->desc.dy = dy; s->desc.dz = dz;
        s->desc.xmin = g->xmin; s->desc.ymin = g->ymin;
        set_spacing(s->g, dx, dy, dz);
        mg_release(s);  // the levels' coefficients and the captured V-cycle graph were built from the old spacing
    }
    session_set_coords(s, g->x, g->y);
    const unsigned long long l0 = g_launches.load();
    const double t0 = now_ms();
    cfd_b200_poisson_config_t pc = cfd_b200_get_poisson_config();
    cfd_status_t rc = run_projection(s, prm, &pc, 1, nullptr);
    c->stats.kernel_time_ms += now_ms() - t0;
    c->stats.kernels_launched += (int)(g_launches.load() - l0);
    c->stats.poisson_iterations = s->last_ps.iterations;
    c->stats.poisson_residual = s->last_ps.final_residual;
    c->stats.memory_allocated = s->bytes_allocated;
    if (stats) *stats = c->stats;
    return rc;
}
gpu_solver_stats_t gpu_solver_get_stats(const gpu_solver_context_t* c) {
    gpu_solver_stats_t z;
    memset(&z, 0, sizeof(z));
    return c ? c->stats : z;
}
void gpu_solver_reset_stats(gpu_solver_context_t* c) {
    if (c) memset(&c->stats, 0, sizeof(c->stats));
}

}  // extern "C"
