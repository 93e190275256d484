// session.cu -- session lifecycle, host<->device movement, NCCL plumbing, globals.
#include "session.h"
#include "tma.cuh"

#include <dlfcn.h>
#include <limits.h>
#include <math.h>
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

namespace cfdb {

std::atomic<unsigned long long> g_launches{0};
thread_local Profiler* t_prof = nullptr;
static thread_local char t_err[512] = "";

const char* const kKernelNames[KK_COUNT] = {
    "neumann_fill", "jacobi_sweep", "redblack_half_sweep", "linf_residual", "stationary_final",
    "krylov_init", "cg_dir_matvec", "cg_update", "bicg_dir_matvec", "bicg_s_matvec", "bicg_update",
    "predictor", "divergence_rhs", "corrector", "thermal_bc", "periodic_bc", "rk_stage",
    "field_stats", "synth_init", "copy", "euler", "boundary_cells", "fill", "ctl_reset", "rbsor_fused_iteration",
    "jacobi_cluster_solve", "cg_cluster_solve",
    "mg_smooth_colour", "mg_residual", "mg_restrict", "mg_prolong_add", "mg_coarsest_solve", "mg_sweep_two_colours",
    "halo_push_pull_peer", "allreduce_peer"};
// Algorithmic bytes per interior cell per launch.
const double kKernelBytesPerCell[KK_COUNT] = {0, 24, 16, 16, 0, 32, 32, 48, 48, 32, 56,
                                              48, 32, 56, 0, 0, 136, 40, 0, 16, 72, 0, 8, 0, 24, 0, 0,
                                              0, 0, 0, 0, 0, 0,  // multigrid kernels run on every level: no per-cell figure
                                              0, 0};

// The reference keeps a thread-local error string (cfd/core/cfd_status.h:31-43).  When its core
// library is loaded in the process we forward to it; otherwise we keep our own.
extern "C" void cfd_set_error(cfd_status_t, const char*) __attribute__((weak));

void set_error(cfd_status_t st, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(t_err, sizeof(t_err), fmt, ap);
    va_end(ap);
    if (cfd_set_error) cfd_set_error(st, t_err);
}
const char* last_error() { return t_err; }

static std::atomic<int> g_arith{-1};  // -1: environment decides
bool default_fast_arithmetic() {
    const int v = g_arith.load();
    if (v >= 0) return v != 0;
    const char* e = getenv("CFD_B200_ARITH");
    return e && !strcmp(e, "fast");
}

void Profiler::begin(cudaStream_t st) {
    if (used >= (1u << 16)) collect();  // bound the pool on very long profiled runs
    if (used == pool.size()) {
        Pair p;
        cudaEventCreate(&p.a);
        cudaEventCreate(&p.b);
        p.kind = 0;
        pool.push_back(p);
    }
    cudaEventRecord(pool[used].a, st);
}
void Profiler::end(cudaStream_t st, int kind) {
    launches[kind]++;
    if (!on) return;
    pool[used].kind = kind;
    cudaEventRecord(pool[used].b, st);
    used++;
}
void Profiler::collect() {
    for (size_t q = 0; q < used; q++) {
        cudaEventSynchronize(pool[q].b);
        float t = 0.f;
        if (cudaEventElapsedTime(&t, pool[q].a, pool[q].b) == cudaSuccess) ms[pool[q].kind] += t;
    }
    used = 0;
}
void Profiler::release() {
    for (auto& p : pool) {
        cudaEventDestroy(p.a);
        cudaEventDestroy(p.b);
    }
    pool.clear();
    used = 0;
}

cfd_status_t dalloc(cfd_b200_session* s, double** p) {
    if (*p) return CFD_SUCCESS;
    cudaError_t e = cudaMalloc((void**)p, s->N * sizeof(double));
    if (e != cudaSuccess) {
        *p = nullptr;
        cudaGetLastError();
        set_error(CFD_ERROR_NOMEM, "cudaMalloc of %zu bytes failed: %s", s->N * sizeof(double),
                  cudaGetErrorString(e));
        return CFD_ERROR_NOMEM;
    }
    s->bytes_allocated += s->N * sizeof(double);
    CFDB_CUDA_OK(cudaMemsetAsync(*p, 0, s->N * sizeof(double), s->st));
    return CFD_SUCCESS;
}

// Every spacing-derived constant, formed exactly as the reference forms it on the host.
void set_spacing(Geom& g, double dx, double dy, double dz) {
    g.two_dx = 2.0 * dx;
    g.two_dy = 2.0 * dy;
    g.dx2 = dx * dx;
    g.dy2 = dy * dy;
    g.inv_2dx = 1.0 / (2.0 * dx);
    g.inv_2dy = 1.0 / (2.0 * dy);
    g.inv_dx2 = 1.0 / (dx * dx);
    g.inv_dy2 = 1.0 / (dy * dy);
    g.inv_2dz = (g.dim3 && dz > 0.0) ? 1.0 / (2.0 * dz) : 0.0;
    g.inv_dz2 = (g.dim3 && dz > 0.0) ? 1.0 / (dz * dz) : 0.0;
    g.factor = 2.0 * (g.inv_dx2 + g.inv_dy2 + g.inv_dz2);
    g.inv_factor = 1.0 / g.factor;
}

static void fill_geom(cfd_b200_session* s) {
    const cfd_b200_session_desc_t& d = s->desc;
    Geom& g = s->g;
    memset(&g, 0, sizeof(g));
    g.nx = (int)d.nx;
    g.ny = (int)d.ny;
    g.nz = (int)d.nz;
    g.dim3 = d.nz > 1;
    g.sy = (long long)d.nx;
    g.sz = g.dim3 ? (long long)d.nx * d.ny : 0;
    g.k0 = g.dim3 ? 1 : 0;
    g.k1 = g.dim3 ? g.nz - 1 : 1;
    set_spacing(g, d.dx, d.dy, d.dz);
    for (int a = 0; a < 3; a++) g.glo[a] = g.ghi[a] = 1;
    s->dec_axis = g.dim3 ? 2 : 1;
    if (d.nranks > 1) {
        g.glo[s->dec_axis] = (d.rank == 0);
        g.ghi[s->dec_axis] = (d.rank == d.nranks - 1);
        // global index of local layer 0 (uniform slabs: every rank owns n-2 interior layers)
        if (g.dim3) g.gk0 = (long long)d.rank * (long long)(d.nz - 2);
        else g.gj0 = (long long)d.rank * (long long)(d.ny - 2);
    }
    g.ncell = (long long)d.nx * d.ny * d.nz;
}

}  // namespace cfdb

using namespace cfdb;

extern "C" {

const char* cfd_b200_last_error(void) { return cfdb::last_error(); }
const char* cfd_b200_version(void) { return "cfd_b200 0.1.0 (sm_100a, FP64)"; }
unsigned long long cfd_b200_kernel_launch_count(void) { return g_launches.load(); }

cfd_status_t cfd_b200_set_arithmetic(int fast) {
    if (fast < -1 || fast > 1) return CFD_ERROR_INVALID;
    cfdb::g_arith.store(fast);
    return CFD_SUCCESS;
}
int cfd_b200_session_arithmetic(const cfd_b200_session_t* s) { return s && s->fast ? 1 : 0; }

void* cfd_b200_host_alloc(size_t bytes) {
    void* p = nullptr;
    if (cudaHostAlloc(&p, bytes, cudaHostAllocDefault) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    return p;
}
void cfd_b200_host_free(void* p) {
    if (p) cudaFreeHost(p);
}

cfd_b200_session_t* cfd_b200_session_create(const cfd_b200_session_desc_t* d) {
    if (!d || d->nx < 3 || d->ny < 3 || d->nz == 0 || d->nz == 2 || d->dx <= 0.0 || d->dy <= 0.0 ||
        (d->nz > 1 && d->dz <= 0.0) || d->nranks < 1 || d->rank < 0 || d->rank >= d->nranks) {
        set_error(CFD_ERROR_INVALID, "cfd_b200_session_create: invalid descriptor");
        return nullptr;
    }
    if (d->nx > INT_MAX / 2 || d->ny > INT_MAX / 2 || d->nz > INT_MAX / 2) {
        set_error(CFD_ERROR_LIMIT_EXCEEDED, "cfd_b200_session_create: extent too large");
        return nullptr;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        set_error(CFD_ERROR_UNSUPPORTED, "no CUDA device available (libcfd_b200 has no CPU fallback)");
        return nullptr;
    }
    cfd_b200_session* s = new cfd_b200_session();
    s->desc = *d;
    if (d->device >= 0) {
        if (cudaSetDevice(d->device) != cudaSuccess) {
            set_error(CFD_ERROR, "cudaSetDevice(%d) failed", d->device);
            delete s;
            return nullptr;
        }
    }
    cudaGetDevice(&s->dev);
    s->fast = default_fast_arithmetic();
    fill_geom(s);
    s->N = (size_t)s->g.ncell;
    bool ok = cudaStreamCreateWithFlags(&s->st, cudaStreamNonBlocking) == cudaSuccess;
    // z chunking: enough blocks for ~20 waves of 148 SMs x 8 resident blocks
    {
        const long long tiles = (long long)((s->g.nx + 31) / 32) * ((s->g.ny + 7) / 8);
        const int nzi = s->g.k1 - s->g.k0;
        long long want = (148LL * 8 * 20 + tiles - 1) / tiles;
        if (want < 1) want = 1;
        if (want > nzi) want = nzi;
        s->zchunk = (int)((nzi + want - 1) / want);
        if (s->zchunk < 1) s->zchunk = 1;
        const long long nchunks = (nzi + s->zchunk - 1) / s->zchunk;
        s->rb_blocks = (size_t)(tiles * nchunks);
        if (s->rb_blocks < 148 * 16) s->rb_blocks = 148 * 16;
        // TMA kernels: 64x16 tiles, 2 blocks per SM; aim at >= 8 waves but keep chunks >= 32 planes
        // (each chunk re-reads 2 planes in its prologue)
        const long long ttiles = (long long)(Tile::GX(s->g.nx)) * ((s->g.ny - 2 + Tile::BY - 1) / Tile::BY);
        // Short chunks win (measured at 1024^3: 32 planes -> 1.03 of the copy peak for bicg_dir, 341
        // planes -> 0.85): blocks of one x-y neighbourhood stay within a few planes of each other, so
        // the halo rows they share are still in L2.  32 planes cost 6 % prologue re-reads; small grids
        // use shorter chunks to get at least ~4 waves of 148 x 2 blocks.
        const long long nch_target = (148LL * 2 * 4 + ttiles - 1) / ttiles;
        long long zc = (nzi + nch_target - 1) / nch_target;
        if (zc > 32) zc = 32;
        if (zc < 4) zc = 4;
        if (zc > nzi) zc = nzi;
        // Cache-resident 3-D grids (fewer than ~3 waves of blocks): wave quantisation and the per-block pipeline fill
        // decide, not L2 locality.  Measured at 128^3 (profiles/r2l_zchunk_sweep.jsonl): 4-plane chunks = 512 blocks;
        // 7-plane chunks = 288 blocks (two per SM) take the Jacobi step from 1.31 to 1.09 ms, 14-plane chunks = 144
        // blocks (one per SM) take RK4 from 0.405 to 0.333 ms per step.  Cost model: waves x (planes + prologue + fill);
        // kernels that run one block per SM (rk_stage, euler, predictor) get their own chunk length.
        const bool small_grid = ttiles * ((nzi + zc - 1) / zc) < 3 * 148 * 2;  // (a property of the grid, not of a kernel)
        auto small_grid_chunk = [&](long long slots, long long zc_now) -> long long {
            if (!small_grid) return zc_now;
            long long best = zc_now;
            double best_cost = 1e300;
            for (long long z = 4; z <= 32 && z <= nzi; z++) {
                const long long blocks = ttiles * ((nzi + z - 1) / z);
                const double cost = (double)((blocks + slots - 1) / slots) * (double)(z + 5);
                if (cost < best_cost) {
                    best_cost = cost;
                    best = z;
                }
            }
            return best;
        };
        const long long zc1 = s->g.dim3 ? small_grid_chunk(148, zc) : zc;
        if (s->g.dim3) zc = small_grid_chunk(148 * 2, zc);
        s->zchunk_tma = (int)(zc < 1 ? 1 : zc);
        s->zchunk_tma1 = (int)(zc1 < 1 ? 1 : zc1);
        if (!s->g.dim3) {
            // 2-D: the TMA kernels march along y; zchunk_tma = tiles of 16 rows per block
            const long long tx = Tile::GX(s->g.nx), ty = (s->g.ny - 2 + Tile::BY - 1) / Tile::BY;
            const long long nch2 = (148LL * 2 * 4 + tx - 1) / tx;
            long long yc = (ty + nch2 - 1) / nch2;
            if (yc > 32) yc = 32;
            if (yc < 1) yc = 1;
            s->zchunk_tma = (int)yc;
            s->zchunk_tma1 = (int)yc;
        }
        if (const char* e = getenv("CFD_B200_ZCHUNK_TMA")) {  // tuning knob
            const int v = atoi(e);
            if (v >= 1) s->zchunk_tma = s->zchunk_tma1 = v > nzi ? nzi : v;
        }
        const size_t tma_blocks = s->g.dim3 ? (size_t)ttiles * (size_t)((nzi + s->zchunk_tma - 1) / s->zchunk_tma)
                                            : (size_t)ttiles;
        if (s->rb_blocks < tma_blocks) s->rb_blocks = tma_blocks;  // one partial-sum row per block
        s->tma.configure(s->g);
    }
    ok = ok && cudaMalloc((void**)&s->ctl, sizeof(Ctl)) == cudaSuccess;
    ok = ok && cudaMemset(s->ctl, 0, sizeof(Ctl)) == cudaSuccess;
    ok = ok && cudaMalloc((void**)&s->rb.partials, s->rb_blocks * 4 * sizeof(double)) == cudaSuccess;
    ok = ok && cudaMalloc((void**)&s->rb.ticket, sizeof(unsigned)) == cudaSuccess;
    ok = ok && cudaMemset(s->rb.ticket, 0, sizeof(unsigned)) == cudaSuccess;
    ok = ok && cudaHostAlloc((void**)&s->h_ctl, 2 * sizeof(Ctl), cudaHostAllocDefault) == cudaSuccess;
    ok = ok && cudaEventCreateWithFlags(&s->ev[0], cudaEventDisableTiming) == cudaSuccess;
    ok = ok && cudaEventCreateWithFlags(&s->ev[1], cudaEventDisableTiming) == cudaSuccess;
    ok = ok && cudaMalloc((void**)&s->su_tab, d->ny * sizeof(double)) == cudaSuccess;
    ok = ok && cudaMalloc((void**)&s->sv_tab, d->nx * sizeof(double)) == cudaSuccess;
    if (ok) {
        double** prim[] = {&s->u, &s->v, &s->w, &s->p};  // T and rho are allocated on first use
        for (auto pp : prim) ok = ok && dalloc(s, pp) == CFD_SUCCESS;
    }
    if (!ok) {
        if (!t_err[0]) set_error(CFD_ERROR_NOMEM, "cfd_b200_session_create: device allocation failed");
        cudaGetLastError();
        cfd_b200_session_destroy(s);
        return nullptr;
    }
    s->xs.resize(d->nx);
    s->ys.resize(d->ny);
    for (size_t i = 0; i < d->nx; i++) s->xs[i] = d->xmin + ((double)i * d->dx);
    for (size_t j = 0; j < d->ny; j++) s->ys[j] = d->ymin + ((double)j * d->dy);
    cudaStreamSynchronize(s->st);
    return s;
}

void cfd_b200_session_destroy(cfd_b200_session_t* s) {
    if (!s) return;
    cudaSetDevice(s->dev);
    if (s->st) cudaStreamSynchronize(s->st);
    double* all[] = {s->u, s->v, s->w, s->p, s->T, s->rho, s->us, s->vs, s->ws, s->pn, s->rhs, s->Tn, s->xt,
                     s->su_tab, s->sv_tab};
    for (double* q : all) if (q) cudaFree(q);
    for (double* q : s->kv) if (q) cudaFree(q);
    for (double* q : s->rk) if (q) cudaFree(q);
    mg_release(s);
    mgpc_release(s);
    if (s->ctl) cudaFree(s->ctl);
    if (s->rb.partials) cudaFree(s->rb.partials);
    if (s->rb.ticket) cudaFree(s->rb.ticket);
    if (s->h_ctl) cudaFreeHost(s->h_ctl);
    for (auto& e : s->ev) if (e) cudaEventDestroy(e);
    s->prof.release();
    if (s->nccl.comm && s->nccl.CommDestroy) s->nccl.CommDestroy(s->nccl.comm);
    peer_release(s);
    if (s->st) cudaStreamDestroy(s->st);
    if (t_prof == &s->prof) t_prof = nullptr;
    delete s;
}

void* cfd_b200_session_stream(cfd_b200_session_t* s) { return s ? (void*)s->st : nullptr; }
cfd_status_t cfd_b200_session_sync(cfd_b200_session_t* s) {
    if (!s) return CFD_ERROR_INVALID;
    CFDB_CUDA_OK(cudaStreamSynchronize(s->st));
    return CFD_SUCCESS;
}
poisson_solver_stats_t cfd_b200_session_last_poisson_stats(const cfd_b200_session_t* s) {
    return s->last_ps;
}

void cfd_b200_session_profile(cfd_b200_session_t* s, int enable) {
    if (!enable && s->prof.on) s->prof.collect();
    s->prof.on = enable != 0;
}
void cfd_b200_session_reset_kernel_stats(cfd_b200_session_t* s) {
    s->prof.collect();
    for (int k = 0; k < KK_COUNT; k++) {
        s->prof.launches[k] = 0;
        s->prof.ms[k] = 0.0;
    }
}
int cfd_b200_session_kernel_stats(const cfd_b200_session_t* cs, cfd_b200_kernel_stat_t* out, int max) {
    cfd_b200_session_t* s = const_cast<cfd_b200_session_t*>(cs);
    s->prof.collect();
    int n = 0;
    for (int k = 0; k < KK_COUNT && n < max; k++) {
        if (!s->prof.launches[k]) continue;
        out[n].name = kKernelNames[k];
        out[n].launches = s->prof.launches[k];
        out[n].total_ms = s->prof.ms[k];
        out[n].bytes_per_cell = kKernelBytesPerCell[k];
        n++;
    }
    return n;
}

}  // extern "C"
