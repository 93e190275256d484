// gang.cu -- Runner: one reference-facing call executed on 1..G GPUs of this process (runner.h).
//
// The reference's GPU entry points receive ONE host flow_field (gpu_device.h:209-249; called from
// solver_registry.c:1121-1245 behind run_simulation_step, simulation_api.c:185-202).  With
// CFD_B200_NGPU=G / cfd_b200_set_device_count(G) the same call drives G devices: the grid is cut into G slabs
// along z (along y when nz == 1; uneven when the interior layers do not divide), each slab is a contiguous block
// of the caller's arrays so G copy engines move them in parallel straight from / to the caller's memory, one
// host thread per GPU runs the unchanged per-slab driver (step_driver.cu), and the slabs talk through NCCL
// communicators made by ncclCommInitAll (halo planes, Krylov dots, L-inf residuals, NaN flag).
#include "runner.h"

#include <atomic>
#include <stdlib.h>
#include <string.h>
#include <thread>

namespace cfdb {

const char* last_error();

namespace {

std::atomic<int> g_ngpu{0};  // 0: not set by the API -> environment

template <class F>
void parallel(int n, F fn) {
    std::vector<std::thread> th;
    for (int r = 1; r < n; r++) th.emplace_back([&fn, r] { fn(r); });
    fn(0);
    for (std::thread& t : th) t.join();
}

// Local view of the caller's arrays for rank `rank`: layers [lo, lo + nloc) of the decomposed axis.
flow_field slab_view(const Runner* r, int rank, const flow_field* f) {
    flow_field v = *f;
    const bool d3 = r->nz > 1;
    const size_t stride = d3 ? r->nx * r->ny : r->nx;
    const size_t off = r->lo[rank] * stride;
    if (d3) v.nz = r->nloc[rank];
    else v.ny = r->nloc[rank];
    double** src[6] = {&v.u, &v.v, &v.w, &v.p, &v.rho, &v.T};
    for (auto q : src)
        if (*q) *q += off;
    return v;
}

}  // namespace

// Periodic wrap of the HOST density ghost cells: the only change apply_boundary_conditions makes to rho
// (solver_explicit_euler.c:231-314); the device copy is wrapped by the stepper, the caller's copy here.
void wrap_host_rho(flow_field* f) {
    const size_t nx = f->nx, ny = f->ny, nz = f->nz, plane = nx * ny;
    double* r = f->rho;
    for (size_t k = 0; k < nz; k++) {
        double* q = r + k * plane;
        for (size_t j = 0; j < ny; j++) { q[j * nx] = q[j * nx + nx - 2]; q[j * nx + nx - 1] = q[j * nx + 1]; }
        for (size_t i = 0; i < nx; i++) { q[i] = q[(ny - 2) * nx + i]; q[(ny - 1) * nx + i] = q[nx + i]; }
    }
    if (nz > 1)
        for (size_t n = 0; n < plane; n++) { r[n] = r[(nz - 2) * plane + n]; r[(nz - 1) * plane + n] = r[plane + n]; }
}

// CFD_B200_VIRTUAL_RANKS=1: all slabs on the CURRENT device, one stream each (peer transport only).  The
// decomposition, halo and reduction logic is then exercised on a one-GPU box.
static bool virtual_ranks() {
    const char* e = getenv("CFD_B200_VIRTUAL_RANKS");
    return e && e[0] && e[0] != '0';
}

int configured_device_count() {
    int want = g_ngpu.load();
    if (want <= 0) {
        const char* e = getenv("CFD_B200_NGPU");
        want = e ? atoi(e) : 1;
    }
    if (want < 1) want = 1;
    if (want > PEER_MAX_RANKS) want = PEER_MAX_RANKS;
    if (virtual_ranks()) return want;
    int have = 0;
    if (cudaGetDeviceCount(&have) != cudaSuccess) { cudaGetLastError(); have = 0; }
    if (want > have) want = have > 0 ? have : 1;
    return want;
}

// Slabs of one process talk through peer memory: a box per rank, every device mapped into every other.
static bool connect_peers(Runner* r, const std::vector<int>& devs) {
    const int G = r->G;
    for (int q = 0; q < G; q++)
        if (peer_alloc_box(r->s[q]) != CFD_SUCCESS) return false;
    for (int a = 0; a < G; a++) {
        cudaSetDevice(devs[a]);
        for (int b = 0; b < G; b++) {
            if (devs[a] == devs[b]) continue;
            int can = 0;
            if (cudaDeviceCanAccessPeer(&can, devs[a], devs[b]) != cudaSuccess || !can) {
                cudaGetLastError();
                set_error(CFD_ERROR_UNSUPPORTED, "device %d cannot map device %d's memory", devs[a], devs[b]);
                return false;
            }
            const cudaError_t e = cudaDeviceEnablePeerAccess(devs[b], 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
                set_error(CFD_ERROR_UNSUPPORTED, "cudaDeviceEnablePeerAccess(%d -> %d): %s", devs[a], devs[b], cudaGetErrorString(e));
                cudaGetLastError();
                return false;
            }
            cudaGetLastError();
        }
    }
    for (int a = 0; a < G; a++) {
        for (int b = 0; b < G; b++) r->s[a]->peer.peer_box[b] = r->s[b]->peer.box;
        if (peer_connected(r->s[a]) != CFD_SUCCESS) return false;
    }
    return true;
}

static int effective_gpus(const flow_field* f, int want) {
    const size_t layers = (f->nz > 1 ? f->nz : f->ny) - 2;
    // a slab narrower than 4 owned layers is all halo: fewer, wider slabs then
    while (want > 1 && layers / (size_t)want < 4) want--;
    return want;
}

bool runner_matches(const Runner* r, const flow_field* f, const grid* g, int want) {
    const double dz = (f->nz > 1 && g->dz) ? g->dz[0] : 0.0;
    int dev = 0;
    cudaGetDevice(&dev);
    return r->nx == f->nx && r->ny == f->ny && r->nz == f->nz && r->dx == g->dx[0] && r->dy == g->dy[0] &&
           r->dz == dz && r->G == effective_gpus(f, want) && r->dev0 == dev && !r->s.empty() &&
           r->s[0]->fast == default_fast_arithmetic();
}

void runner_destroy(Runner* r) {
    if (!r) return;
    int cur = 0;
    cudaGetDevice(&cur);
    for (size_t q = 0; q < r->s.size(); q++) {
        if (!r->s[q]) continue;
        cudaSetDevice(r->s[q]->dev);
        if (q < r->d_pack.size() && r->d_pack[q]) cudaFree(r->d_pack[q]);
        if (q < r->h_pack.size() && r->h_pack[q]) cudaFreeHost(r->h_pack[q]);
        cfd_b200_session_destroy(r->s[q]);
    }
    cudaSetDevice(cur);
    delete r;
}

Runner* runner_create(const flow_field* f, const grid* g, int want) {
    Runner* r = new Runner();
    r->G = effective_gpus(f, want);
    r->nx = f->nx; r->ny = f->ny; r->nz = f->nz;
    r->dx = g->dx[0]; r->dy = g->dy[0];
    r->dz = (f->nz > 1 && g->dz) ? g->dz[0] : 0.0;
    cudaGetDevice(&r->dev0);
    int ndev = 1;
    cudaGetDeviceCount(&ndev);
    const int G = r->G;
    const bool d3 = f->nz > 1;
    const size_t glayers = d3 ? f->nz : f->ny, interior = glayers - 2;
    r->s.assign(G, nullptr);
    r->lo.resize(G);
    r->nloc.resize(G);
    r->h_pack.assign(G, nullptr);
    r->d_pack.assign(G, nullptr);
    r->pack_cap.assign(G, 0);
    size_t off = 0;
    for (int q = 0; q < G; q++) {  // uneven slabs: the first (interior % G) ranks own one layer more
        const size_t own = interior / G + ((size_t)q < interior % G ? 1 : 0);
        r->lo[q] = off;
        r->nloc[q] = own + 2;
        off += own;
    }
    std::vector<int> devs(G);
    bool ok = true;
    for (int q = 0; q < G && ok; q++) {
        cfd_b200_session_desc_t d;
        memset(&d, 0, sizeof(d));
        d.nx = f->nx;
        d.ny = d3 ? f->ny : r->nloc[q];
        d.nz = d3 ? r->nloc[q] : 1;
        d.dx = r->dx; d.dy = r->dy; d.dz = r->dz;
        d.xmin = g->xmin;
        d.ymin = d3 ? g->ymin : g->ymin + (double)r->lo[q] * r->dy;
        d.zmin = d3 ? g->zmin + (double)r->lo[q] * r->dz : g->zmin;
        d.rank = q;
        d.nranks = G;
        d.device = (G > 1 && !virtual_ranks()) ? (r->dev0 + q) % ndev : -1;
        devs[q] = d.device >= 0 ? d.device : r->dev0;
        r->s[q] = cfd_b200_session_create(&d);
        ok = r->s[q] != nullptr;
        if (ok && G > 1) ok = cfd_b200_session_set_slab(r->s[q], glayers, r->lo[q]) == CFD_SUCCESS;
    }
    const char* tr = getenv("CFD_B200_GANG_TRANSPORT");
    bool want_nccl = tr && !strcmp(tr, "nccl") && !virtual_ranks();
    if (ok && G > 1 && !want_nccl && !connect_peers(r, devs)) {
        if (virtual_ranks()) ok = false;
        else want_nccl = true;  // no peer mapping between these devices: the library transport
    }
    if (ok && G > 1 && want_nccl) {
        for (int q = 0; q < G; q++) {
            r->s[q]->peer.on = false;
            r->s[q]->rb.peer = nullptr;
        }
        Nccl n;
        ok = load_nccl(n, getenv("CFD_B200_NCCL_LIB")) == CFD_SUCCESS;
        std::vector<void*> comms(G, nullptr);
        if (ok && n.CommInitAll(comms.data(), G, devs.data()) != 0) {
            set_error(CFD_ERROR, "ncclCommInitAll over %d devices failed", G);
            ok = false;
        }
        for (int q = 0; q < G && ok; q++) {
            r->s[q]->nccl = n;
            r->s[q]->nccl.comm = comms[q];
        }
    }
    cudaSetDevice(r->dev0);
    if (!ok) {
        runner_destroy(r);
        return nullptr;
    }
    return r;
}

static int upload_mask(StepKind kind, const ns_solver_params_t* prm) {
    if (kind != SK_PROJECTION) return FM_ALL;
    const bool thermo = (prm->alpha > 0.0 || prm->beta != 0.0);
    return FM_U | FM_V | FM_W | FM_P | (thermo ? FM_T : 0);
}
static int download_mask(StepKind kind, const ns_solver_params_t* prm) {
    if (kind != SK_PROJECTION) return FM_U | FM_V | FM_W | FM_P | FM_T;
    return FM_U | FM_V | FM_W | FM_P | (prm->alpha > 0.0 ? FM_T : 0);
}

// Owned layers of every rank (plus the physical boundary layers on the end ranks) back into the caller's arrays.
cfd_status_t runner_download(Runner* r, flow_field* f, int mask) {
    const int G = r->G;
    const bool d3 = r->nz > 1;
    const size_t stride = d3 ? r->nx * r->ny : r->nx;
    std::vector<cfd_status_t> rc(G, CFD_SUCCESS);
    std::vector<std::string> err(G);
    parallel(G, [&](int q) {
        cfd_b200_session* s = r->s[q];
        if (G == 1) {
            rc[q] = session_download(s, f, mask);
        } else {
            const size_t a = (q == 0) ? 0 : 1, b = (q == G - 1) ? r->nloc[q] : r->nloc[q] - 1;
            double* dv[6] = {s->u, s->v, s->w, s->p, s->rho, s->T};
            double* hv[6] = {f->u, f->v, f->w, f->p, f->rho, f->T};
            for (int k = 0; k < 6; k++) {
                if (!(mask & (1 << k)) || !dv[k] || !hv[k]) { dv[k] = nullptr; continue; }
                dv[k] += a * stride;
                hv[k] += (r->lo[q] + a) * stride;
            }
            rc[q] = xfer_fields(s, dv, hv, 6, (b - a) * stride * sizeof(double), false);
        }
        if (rc[q] != CFD_SUCCESS) err[q] = last_error();
    });
    for (int q = 0; q < G; q++)
        if (rc[q] != CFD_SUCCESS) {
            set_error(rc[q], "%s", err[q].c_str());
            return rc[q];
        }
    r->host_stale = false;
    return CFD_SUCCESS;
}

cfd_status_t runner_run(Runner* r, StepKind kind, flow_field* f, const grid* g, const ns_solver_params_t* prm,
                        int nsteps, ns_solver_stats_t* stats, int sync) {
    const int G = r->G;
    const bool d3 = r->nz > 1;
    int cur = 0;
    cudaGetDevice(&cur);
    std::vector<cfd_status_t> rc(G, CFD_SUCCESS);
    std::vector<std::string> err(G);
    auto first_failure = [&]() -> cfd_status_t {
        for (int q = 0; q < G; q++)
            if (rc[q] != CFD_SUCCESS) {
                set_error(rc[q], "%s", err[q].c_str());
                return rc[q];
            }
        return CFD_SUCCESS;
    };
    const int umask = upload_mask(kind, prm);
    const cfd_b200_poisson_config_t pc = cfd_b200_get_poisson_config();
    // ---- host -> device: whole slabs, or (caller's promise: only boundary cells edited since the last call's
    //      result) the boundary faces; nothing at all while the host copy is the stale one (DOWNLOAD_LAZY)
    const bool state_kept = r->resident && (r->resident_mask & umask) == umask;
    const bool skip_upload = state_kept && r->host_stale;
    const bool faces_only = state_kept && !r->host_stale && (sync & SYNC_UPLOAD_BOUNDARY);
    parallel(G, [&](int q) {
        cfd_b200_session* s = r->s[q];
        session_set_coords(s, g->x, d3 ? g->y : g->y + r->lo[q]);
        if (skip_upload) {
            if (G > 1 && (rc[q] = prepare_buffers(s, (int)kind, prm, &pc)) != CFD_SUCCESS) err[q] = last_error();
            return;
        }
        const flow_field v = slab_view(r, q, f);
        rc[q] = faces_only ? upload_faces(r, q, &v, umask) : session_upload(s, &v, umask);
        // all allocations of the run now, before any slab can be waiting for another inside a kernel
        if (rc[q] == CFD_SUCCESS && G > 1) rc[q] = prepare_buffers(s, (int)kind, prm, &pc);
        if (rc[q] != CFD_SUCCESS) err[q] = last_error();
    });
    cudaSetDevice(cur);
    r->resident = false;
    cfd_status_t st = first_failure();
    if (st != CFD_SUCCESS) return st;
    // ---- step: one host thread per GPU runs the per-slab driver
    std::vector<ns_solver_stats_t> rs(G);
    parallel(G, [&](int q) {
        cfd_b200_session* s = r->s[q];
        if (getenv("CFD_B200_PROFILE_GANG")) s->prof.on = true;
        ns_solver_stats_t* ps = stats ? &rs[q] : nullptr;
        if (ps) *ps = *stats;
        rc[q] = (kind == SK_PROJECTION) ? run_projection(s, prm, &pc, nsteps, ps) : run_rk(s, (int)kind, prm, nsteps, ps);
        if (rc[q] != CFD_SUCCESS) err[q] = last_error();
    });
    cudaSetDevice(cur);
    if (getenv("CFD_B200_PROFILE_GANG")) {  // debugging aid: per-kernel event times of every slab
        for (int q = 0; q < G; q++) {
            cfd_b200_kernel_stat_t ks[CFD_B200_MAX_KERNEL_KINDS];
            const int nk = cfd_b200_session_kernel_stats(r->s[q], ks, CFD_B200_MAX_KERNEL_KINDS);
            for (int k = 0; k < nk; k++)
                fprintf(stderr, "[gang] rank %d %-24s x%-4llu %10.3f ms\n", q, ks[k].name, ks[k].launches, ks[k].total_ms);
        }
    }
    if (stats) *stats = rs[0];  // maxima are all-reduced: identical on every rank
    st = first_failure();
    const bool usable = st == CFD_SUCCESS || st == CFD_ERROR_DIVERGED || (kind == SK_PROJECTION && st == CFD_ERROR_MAX_ITER);
    if (!usable) return st;
    // ---- device -> host
    if (st == CFD_SUCCESS) {
        r->resident = true;
        r->resident_mask = umask;
    }
    if ((sync & SYNC_DOWNLOAD_LAZY) && st == CFD_SUCCESS) {
        r->host_stale = true;
        return st;
    }
    const cfd_status_t r2 = runner_download(r, f, download_mask(kind, prm));
    cudaSetDevice(cur);
    if (r2 != CFD_SUCCESS) return r2;
    if (st != CFD_SUCCESS) set_error(st, "%s", err[0].c_str());
    return st;
}

// ---------------------------------------------------------------------------------------------
// Runner cache of the stateless drop-in symbols: device buffers persist across calls (the reference
// re-allocates 22 buffers per call, solver_projection_gpu.cu:417-436).  A small LRU keyed by shape, spacing,
// device and GPU count, so two simulations of different sizes in one process do not evict each other on every
// call and run concurrently on their own runners; a runner is used by one call at a time.
// ---------------------------------------------------------------------------------------------
namespace {
std::mutex g_cache_mu;
std::vector<Runner*> g_cache;
unsigned long long g_stamp = 0;
bool g_hooked = false;

size_t cache_capacity() {
    static const size_t n = [] {
        const char* e = getenv("CFD_B200_SESSION_CACHE");
        const int v = e ? atoi(e) : 2;
        return (size_t)(v < 1 ? 1 : (v > 16 ? 16 : v));
    }();
    return n;
}
}  // namespace

void runner_cache_clear() {
    std::vector<Runner*> dead;
    {
        std::lock_guard<std::mutex> lk(g_cache_mu);
        for (size_t q = 0; q < g_cache.size();) {
            if (g_cache[q]->busy) { q++; continue; }
            dead.push_back(g_cache[q]);
            g_cache.erase(g_cache.begin() + q);
        }
    }
    for (Runner* r : dead) runner_destroy(r);
}

Runner* runner_cache_acquire(const flow_field* f, const grid* g) {
    int want = configured_device_count();
    if (cfd_b200_get_poisson_config().method == POISSON_METHOD_MULTIGRID) want = 1;  // single-GPU hierarchy
    Runner* hit = nullptr;
    std::vector<Runner*> dead;
    {
        std::lock_guard<std::mutex> lk(g_cache_mu);
        if (!g_hooked) {
            g_hooked = true;
            atexit([] { runner_cache_clear(); });
        }
        for (Runner* r : g_cache)
            if (!r->busy && runner_matches(r, f, g, want)) { hit = r; break; }
        if (hit) {
            hit->busy = true;
            hit->stamp = ++g_stamp;
        } else {
            // make room first: idle runners beyond capacity - 1, least recently used first
            while (g_cache.size() + 1 > cache_capacity()) {
                int lru = -1;
                for (size_t q = 0; q < g_cache.size(); q++)
                    if (!g_cache[q]->busy && (lru < 0 || g_cache[q]->stamp < g_cache[lru]->stamp)) lru = (int)q;
                if (lru < 0) break;
                dead.push_back(g_cache[lru]);
                g_cache.erase(g_cache.begin() + lru);
            }
        }
    }
    for (Runner* r : dead) runner_destroy(r);
    if (hit) return hit;
    Runner* r = runner_create(f, g, want);
    if (!r) {  // out of device memory with idle runners around: drop them and try once more
        runner_cache_clear();
        r = runner_create(f, g, want);
    }
    if (!r) return nullptr;
    std::lock_guard<std::mutex> lk(g_cache_mu);
    r->busy = true;
    r->stamp = ++g_stamp;
    g_cache.push_back(r);
    return r;
}

void runner_cache_release(Runner* r) {
    std::lock_guard<std::mutex> lk(g_cache_mu);
    r->busy = false;
}

}  // namespace cfdb

using namespace cfdb;

extern "C" {

cfd_status_t cfd_b200_set_device_count(int ngpu) {
    if (ngpu < 0) return CFD_ERROR_INVALID;
    g_ngpu.store(ngpu);
    return CFD_SUCCESS;
}
int cfd_b200_get_device_count(void) { return configured_device_count(); }
void cfd_b200_release_cached_sessions(void) { runner_cache_clear(); }

}  // extern "C"
