// kernels.h -- host-side launchers of every CUDA kernel in libcfd_b200.
#pragma once
#include "common.cuh"
#include "ns_args.h"

#include <cuda.h>
#include <unordered_map>
#include <vector>

namespace cfdb {

// Tensor maps of a session's device arrays for the TMA kernels (tma_host.cu).
struct TmaCache {
    int nx = 0, ny = 0, nz = 0;
    bool usable = false;
    std::unordered_map<const double*, CUtensorMap> maps, wide_maps;
    void configure(const Geom& g);
    const CUtensorMap* get(const double* p);   // nullptr => take the direct (non-TMA) kernel
    const CUtensorMap* get_wide(const double* p);  // {BX+4, BY+4, 1} box (two-cell halo) for the fused sweeps
};

// Everything a launcher needs besides the per-kernel operands.
struct LaunchCtx {
    Geom g;
    cudaStream_t st;
    Ctl* ctl;       // device
    RedBuf rb;      // device
    int zchunk;     // planes marched per block
    TmaCache* tc;   // may be null
    int zchunk_tma; // planes marched per block of the TMA kernels
    int zchunk_tma1; // ... of those that run one block per SM (small grids: one wave of 148 blocks)
    size_t rb_blocks;  // rows of the partial-sum buffer (upper bound on reducing grids)
    dim3 grid() const;   // tile grid over (nx, ny, z-chunks)
};

// Kernel families, for launch accounting and per-kernel timing (cfd_b200_kernel_stat_t).
enum KernelKind {
    KK_NEUMANN = 0, KK_JACOBI, KK_REDBLACK, KK_LINF, KK_STAT_FINAL, KK_KRYLOV_INIT, KK_CG_DIR,
    KK_CG_UPDATE, KK_BICG_DIR, KK_BICG_S, KK_BICG_UPDATE, KK_PREDICTOR, KK_DIVERGENCE,
    KK_CORRECTOR, KK_THERMAL_BC, KK_PERIODIC_BC, KK_RK_STAGE, KK_FIELD_STATS, KK_SYNTH_INIT,
    KK_COPY, KK_EULER, KK_BOUNDARY, KK_FILL, KK_CTL_RESET, KK_RBSOR_FUSED, KK_JACOBI_CLUSTER, KK_CG_CLUSTER,
    KK_MG_SMOOTH, KK_MG_RESIDUAL, KK_MG_RESTRICT, KK_MG_PROLONG, KK_MG_COARSE, KK_MG_SWEEP2, KK_HALO, KK_ALLREDUCE, KK_COUNT
};
extern const char* const kKernelNames[KK_COUNT];
extern const double kKernelBytesPerCell[KK_COUNT];

// Optional per-launch timing (cfd_b200_session_profile): one CUDA event pair per launch from a
// growing pool, recorded on the launching stream and NOT synchronised, so a timed region can
// run with it switched on; collect() resolves the pairs after the stream has drained.
struct Profiler {
    bool on = false;
    unsigned long long launches[KK_COUNT] = {};
    double ms[KK_COUNT] = {};
    struct Pair { cudaEvent_t a, b; int kind; };
    std::vector<Pair> pool;
    size_t used = 0;
    void begin(cudaStream_t st);
    void end(cudaStream_t st, int kind);
    void collect();   // adds the elapsed time of every recorded pair to ms[]
    void release();
};
extern thread_local Profiler* t_prof;

struct NsConst {            // per-step constants of the projection / RK kernels
    double dt, nu, rho, dt_over_rho, rho_over_dt;
    double beta, T_ref, gx, gy, gz;   // Boussinesq (energy_solver.c:185-196)
    double alpha;                      // thermal diffusivity, <= 0 disables the energy step
    int src_on;                        // default sinusoidal source enabled (amplitudes non-zero)
    double mu;                         // RK: nu_eff = min(mu/rho, 1)
    double src_decay;                  // exp(-lambda*iter*dt) (solver_explicit_euler.c:326-331)
};

// stand-alone pull of a pending halo gate (peer.cu): for a launcher that ends up on a kernel without the in-kernel gate
void peer_pull_from_gate(cudaStream_t st, const HaloGate& hg);

// launchers of the two arithmetic builds (common.cuh): exact ones in cfdb::, fast twins in cfdb::arith_fast::
#ifdef CFDB_FAST
inline namespace arith_fast {
#include "launchers.inc"
}
#else
#include "launchers.inc"
namespace arith_fast {
#include "launchers.inc"
}
#endif
// exact build only (poisson_mini.cu): whole solves of small 2-D grids in one thread-block-cluster launch
// whole Jacobi solve of a small 2-D grid in one thread-block-cluster launch (poisson_mini.cu); x in place
bool jacobi_cluster_available(const LaunchCtx& L);
bool launch_jacobi_cluster(const LaunchCtx& L, const KArgs& a, double* x, const double* rhs);
// whole (P)CG solve of a small 2-D grid in one cluster launch; the caller runs the Neumann passes around it
bool cg_cluster_available(const LaunchCtx& L);
bool launch_cg_cluster(const LaunchCtx& L, const KArgs& a, double* x, const double* rhs, int precond);

#define CFDB_LAUNCH(kind, st, ...)                         \
    do {                                                   \
        cfdb::Profiler* pf_ = cfdb::t_prof;                \
        if (pf_ && pf_->on) pf_->begin(st);                \
        __VA_ARGS__;                                       \
        cfdb::g_launches++;                                \
        if (pf_) pf_->end(st, kind);                       \
    } while (0)

}  // namespace cfdb
