// runner.h -- what stands behind ONE reference-facing solver object or stateless drop-in call:
// one session on one GPU, or a gang of G sessions (one z-slab -- y-slab when nz == 1 -- per GPU) driven by
// one host thread per GPU inside this process (SURVEY 8e "process model").  The caller always hands over the
// full host flow_field; the runner scatters it into slabs, steps, and gathers the owned layers back.
#pragma once
#include "session.h"

#include <mutex>
#include <string>
#include <vector>

namespace cfdb {

enum StepKind { SK_PROJECTION = 0, SK_EULER = 1, SK_RK2 = 2, SK_RK4 = 4 };

// host-sync policy bits (include/cfd_b200.h: cfd_b200_host_sync_t)
enum { SYNC_FULL = 0, SYNC_UPLOAD_BOUNDARY = 1, SYNC_DOWNLOAD_LAZY = 2 };

struct Runner {
    int G = 1;
    size_t nx = 0, ny = 0, nz = 0;      // global extents
    double dx = 0, dy = 0, dz = 0;
    int dev0 = 0;                        // first device (gang: dev0 .. dev0+G-1)
    std::vector<cfd_b200_session*> s;    // one per rank
    std::vector<size_t> lo, nloc;        // global first layer / local layer count (halo layers included) per rank
    bool resident = false;               // device fields equal the host fields of the last call (interior cells)
    int resident_mask = 0;               // ... for these fields
    bool host_stale = false;             // DOWNLOAD_LAZY: the device is ahead of the caller's arrays
    std::mutex mu;                       // one call at a time per runner
    bool busy = false;                   // cache bookkeeping (guarded by the cache mutex)
    unsigned long long stamp = 0;
    // pinned staging of the packed x-faces (UPLOAD_BOUNDARY)
    std::vector<double*> h_pack, d_pack;
    std::vector<size_t> pack_cap;        // doubles allocated per rank
};

int configured_device_count();            // cfd_b200_set_device_count / CFD_B200_NGPU, clamped to visible devices
Runner* runner_create(const flow_field* f, const grid* g, int want_gpus);
void runner_destroy(Runner* r);
bool runner_matches(const Runner* r, const flow_field* f, const grid* g, int want_gpus);
cfd_status_t runner_run(Runner* r, StepKind kind, flow_field* f, const grid* g, const ns_solver_params_t* prm,
                        int nsteps, ns_solver_stats_t* stats, int sync_flags);
cfd_status_t runner_download(Runner* r, flow_field* f, int mask);

void wrap_host_rho(flow_field* f);  // periodic images of the caller's density (RK / Euler steppers)

// cache of runners for the stateless drop-in symbols (capi.cu)
Runner* runner_cache_acquire(const flow_field* f, const grid* g);
void runner_cache_release(Runner* r);
void runner_cache_clear();

// boundary-face upload of one slab (host_sync.cu)
cfd_status_t upload_faces(Runner* r, int rank, const flow_field* f, int mask);

}  // namespace cfdb
