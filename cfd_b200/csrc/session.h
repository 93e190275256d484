// session.h -- device-resident state of one slab on one GPU and the host-side drivers.
#pragma once
#include "kernels.h"

#include <vector>

namespace cfdb {

// Minimal NCCL surface, bound with dlopen at run time (no link-time dependency).  The enum
// values are those of nccl.h 2.27/2.28: ncclDouble = 8, ncclSum = 0, ncclMax = 2.
struct NcclId {
    char internal[128];
};
struct Nccl {
    void* lib = nullptr;
    void* comm = nullptr;
    int (*GetUniqueId)(void*) = nullptr;
    int (*CommInitRank)(void**, int, /*ncclUniqueId by value*/ NcclId, int) = nullptr;
    int (*CommDestroy)(void*) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    int (*Send)(const void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    int (*Recv)(void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    int (*CommInitAll)(void**, int, const int*) = nullptr;
};

// (PeerBoxHeader and the PEER_* limits live in common.cuh: the reducing kernels use them)
struct Peer {
    bool on = false;
    char* box = nullptr;                     // my box (device memory, cudaMalloc)
    char* peer_box[PEER_MAX_RANKS] = {};     // every rank's box as mapped here (peer_box[rank] == box)
    bool ipc_mapped[PEER_MAX_RANKS] = {};    // opened with cudaIpcOpenMemHandle (to be closed)
    size_t cnt = 0;                          // doubles per halo layer
    double* halo2[2] = {};                   // SECOND halo layer (lo, hi) of the one-pass Red-Black SOR sweep on slabs:
                                             // two planes at the end of the box (poisson_tma3.inc)
    size_t box_bytes = 0;
    unsigned long long hseq = 0, rseq = 0;   // exchanges / reductions issued so far (identical on every rank)
};

struct MgState;  // multigrid hierarchy of a session (poisson_mg.cu)

enum FieldMask { FM_U = 1, FM_V = 2, FM_W = 4, FM_P = 8, FM_RHO = 16, FM_T = 32, FM_ALL = 63 };

}  // namespace cfdb

struct cfd_b200_session {
    cfd_b200_session_desc_t desc;
    cfdb::Geom g;
    int dev = 0;
    cudaStream_t st = nullptr;
    size_t N = 0;             // cells of the local slab
    int dec_axis = 2;         // decomposed axis: 2 = z (3-D), 1 = y (2-D)
    // primary fields
    double *u = nullptr, *v = nullptr, *w = nullptr, *p = nullptr, *T = nullptr, *rho = nullptr;
    // projection work arrays
    double *us = nullptr, *vs = nullptr, *ws = nullptr, *pn = nullptr, *rhs = nullptr, *Tn = nullptr;
    double* xt = nullptr;     // Jacobi ping-pong buffer
    double* kv[8] = {};       // Krylov work vectors
    cfdb::MgState* mg = nullptr;  // multigrid levels, built on the first POISSON_METHOD_MULTIGRID solve
    void* mgpc = nullptr;         // levels of the multigrid PRECONDITIONER of CG (poisson_mgpc.cu), built on first use
    double* rk[12] = {};      // RK: B[4], C[4], ACC[4]
    double *su_tab = nullptr, *sv_tab = nullptr;  // device source tables [ny], [nx]
    std::vector<double> xs, ys;                   // host coordinates of local i, j
    double tab_au = 0.0, tab_av = 0.0;            // amplitudes the tables were built for
    bool tab_valid = false;
    double rho0 = 1.0;        // field->rho[0] (solver_projection.c:195)
    cfdb::Ctl* ctl = nullptr;     // device
    cfdb::Ctl* h_ctl = nullptr;   // pinned host mirror, 2 entries
    cudaEvent_t ev[2] = {};
    cfdb::RedBuf rb = {};
    size_t rb_blocks = 0;
    int zchunk = 1;
    cfdb::Profiler prof;
    cfdb::TmaCache tma;
    int zchunk_tma = 1;
    int zchunk_tma1 = 1;      // ... of the TMA kernels that run ONE block per SM (rk_stage, euler, predictor)
    cfdb::Nccl nccl;
    cfdb::Peer peer;
    poisson_solver_stats_t last_ps = {};
    size_t bytes_allocated = 0;
    bool fast = false;        // arithmetic policy of this session's stencil / Poisson kernels (common.cuh)
    // slab bookkeeping: global extent of the decomposed axis (boundary layers included; 0 = uniform slabs,
    // derived from nranks) -- cfd_b200_session_set_slab
    size_t gdec = 0;
    // first failure of a halo exchange / all-reduce since the last check (comm_check): a decomposed solve
    // that lost an exchange must not report success on a per-slab answer
    cfd_status_t comm_err = CFD_SUCCESS;
    int rb_slab_ok = -1;        // one-pass Red-Black SOR on slabs: every rank has >= 2 own planes (-1: not agreed yet)
    bool gate_enabled = false;  // peer transport: the pull of an exchange may ride in the consuming kernel (HaloGate)
    int kernels_at_create = 0;

    cfdb::LaunchCtx L() const {
        cfdb::LaunchCtx l;
        l.g = g;
        l.st = st;
        l.ctl = ctl;
        l.rb = rb;
        l.zchunk = zchunk;
        l.tc = const_cast<cfdb::TmaCache*>(&tma);
        l.zchunk_tma = zchunk_tma;
        l.zchunk_tma1 = zchunk_tma1;
        l.rb_blocks = rb_blocks;
        return l;
    }
};

// launcher of the session's arithmetic build
#define CFDB_KL(s, fn) ((s)->fast ? cfdb::arith_fast::fn : cfdb::fn)

namespace cfdb {
bool default_fast_arithmetic();   // cfd_b200_set_arithmetic / CFD_B200_ARITH
void set_spacing(Geom& g, double dx, double dy, double dz);
cfd_status_t dalloc(cfd_b200_session* s, double** p);           // zero-filled N doubles
cfd_status_t session_upload(cfd_b200_session* s, const flow_field* f, int mask);
cfd_status_t session_download(cfd_b200_session* s, flow_field* f, int mask);
// whole-field host <-> device copies: pinned host memory directly, pageable memory through a multi-threaded
// pinned staging pipeline (xfer.cu); returns when the data has arrived
cfd_status_t xfer_fields(cfd_b200_session* s, double* const* dev, double* const* host, int n, size_t bytes,
                         bool to_device);
cfd_status_t session_set_coords(cfd_b200_session* s, const double* x, const double* y);
cfd_status_t halo_exchange(cfd_b200_session* s, double* const* fields, int nf, bool periodic_ring);
cfd_status_t allreduce(cfd_b200_session* s, double* dev, int n, bool is_max);
cfd_status_t allreduce_int_max(cfd_b200_session* s, int* dev, int n);
// nranks > 1 needs a communicator; returns (and clears) the first exchange failure since the last call
cfd_status_t comm_ready(cfd_b200_session* s);
cfd_status_t comm_check(cfd_b200_session* s);
size_t global_dec_extent(const cfd_b200_session* s);  // global layers along the decomposed axis
cfd_status_t load_nccl(Nccl& n, const char* path);
// peer transport (peer.cu)
cfd_status_t peer_alloc_box(cfd_b200_session* s);
void peer_release(cfd_b200_session* s);
cfd_status_t peer_halo_exchange(cfd_b200_session* s, double* const* fields, int nf, bool periodic_ring);
cfd_status_t peer_allreduce(cfd_b200_session* s, void* dev, int n, int op);  // op 0 sum f64, 1 max f64, 2 max i32
// after the boxes are mapped: fill the device-side header (box table, rank, time-out flag) and switch the fused
// reductions of the session on (CFD_B200_NO_FUSED_REDUCE=1: keep the stand-alone all-reduce kernel)
cfd_status_t peer_connected(cfd_b200_session* s);
cfd_status_t peer_halo_push_gated(cfd_b200_session* s, double* const* fields, int nf, HaloGate* gate);
// TWO layers of one field per side: the outer one into the field's own halo layer, the one behind it into
// Peer::halo2 (gate may be null: push + stand-alone pull)
cfd_status_t peer_halo_exchange2(cfd_b200_session* s, double* field, HaloGate* gate);
// all-reduce of values a reducing KERNEL has just left in Ctl: nothing to do when the kernel finished it itself
cfd_status_t allreduce_produced(cfd_b200_session* s, double* dev, int n, bool is_max);
cfd_status_t poisson_solve(cfd_b200_session* s, const cfd_b200_poisson_config_t* cfg, double** x,
                           double** xtmp, const double* rhs, poisson_solver_stats_t* st);
// geometric multigrid (poisson_mg.cu): x in place, boundary cells filled at the end; one GPU only
cfd_status_t mg_solve(cfd_b200_session* s, const cfd_b200_poisson_config_t* cfg, double* x, const double* rhs,
                      poisson_solver_stats_t* st);
void mg_release(cfd_b200_session* s);
// CG preconditioned by a multigrid cycle for the Krylov operator (poisson_mgpc.cu): x in place; one GPU only
cfd_status_t mgpcg_solve(cfd_b200_session* s, const cfd_b200_poisson_config_t* cfg, double* x, const double* rhs,
                         poisson_solver_stats_t* st);
void mgpc_release(cfd_b200_session* s);
int mg_level_count(cfd_b200_session* s);
cfd_status_t run_projection(cfd_b200_session* s, const ns_solver_params_t* prm,
                            const cfd_b200_poisson_config_t* pc, int nsteps, ns_solver_stats_t* stats);
cfd_status_t run_rk(cfd_b200_session* s, int order, const ns_solver_params_t* prm, int nsteps,
                    ns_solver_stats_t* stats);
cfd_status_t prepare_buffers(cfd_b200_session* s, int kind, const ns_solver_params_t* prm,
                             const cfd_b200_poisson_config_t* pc);
cfd_status_t validate_common(const flow_field* f, const grid* g, const ns_solver_params_t* prm);
}  // namespace cfdb
