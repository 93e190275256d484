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
    const CUtensorMap* get_wide(const double* p);  // {BX+6, BY+4, 1} box (two-cell halo, even x start) for the fused sweeps
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

// ---- Poisson ----
void launch_neumann_fill(const LaunchCtx& L, double* x, int seq, int honour_skip, bool use_ctl);
void launch_jacobi(const LaunchCtx& L, const KArgs& a, const double* xin, double* xout, const double* rhs);
void launch_redblack(const LaunchCtx& L, const KArgs& a, double* x, const double* rhs, int colour, int decide);
bool rbsor_fused_available(const LaunchCtx& L);
// whole Red-Black SOR iteration x_q -> x_{q+1} out of place (+ res(x_q)); false => not launched, use the half-sweeps
bool launch_rbsor_fused(const LaunchCtx& L, const KArgs& a, const double* xin, double* xout, const double* rhs);
// whole Jacobi solve of a small 2-D grid in one thread-block-cluster launch (poisson_mini.cu); x in place
bool jacobi_cluster_available(const LaunchCtx& L);
bool launch_jacobi_cluster(const LaunchCtx& L, const KArgs& a, double* x, const double* rhs);
// whole (P)CG solve of a small 2-D grid in one cluster launch; the caller runs the Neumann passes around it
bool cg_cluster_available(const LaunchCtx& L);
bool launch_cg_cluster(const LaunchCtx& L, const KArgs& a, double* x, const double* rhs, int precond);
void launch_linf_residual(const LaunchCtx& L, const KArgs& a, const double* x, const double* rhs, int slot);
void launch_stationary_final(const LaunchCtx& L, const KArgs& a);
void launch_krylov_init(const LaunchCtx& L, const KArgs& a, const double* x, const double* rhs, double* r,
                        double* p0, double* rhat, int precond);
void launch_cg_dir(const LaunchCtx& L, const KArgs& a, const double* r, const double* pold, double* pnew,
                   double* Ap, int precond, int decide_only);
void launch_cg_update(const LaunchCtx& L, const KArgs& a, double* x, double* r, const double* p,
                      const double* Ap, int precond);
void launch_bicg_dir(const LaunchCtx& L, const KArgs& a, const double* r, const double* pold,
                     const double* vold, const double* rhat, double* pnew, double* vnew, int first,
                     int decide_only);
void launch_bicg_s(const LaunchCtx& L, const KArgs& a, const double* r, const double* v, double* s, double* t);
void launch_bicg_update(const LaunchCtx& L, const KArgs& a, double* x, double* r, const double* p,
                        const double* s, const double* t, const double* rhat);

// ---- Navier-Stokes step ----
struct NsConst {            // per-step constants of the projection / RK kernels
    double dt, nu, rho, dt_over_rho, rho_over_dt;
    double beta, T_ref, gx, gy, gz;   // Boussinesq (energy_solver.c:185-196)
    double alpha;                      // thermal diffusivity, <= 0 disables the energy step
    int src_on;                        // default sinusoidal source enabled (amplitudes non-zero)
    double mu;                         // RK: nu_eff = min(mu/rho, 1)
    double src_decay;                  // exp(-lambda*iter*dt) (solver_explicit_euler.c:326-331)
};
// su_tab[j] = A_u sin(pi y_j) exp(-lambda*iter*dt), sv_tab[i] = A_v sin(2 pi x_i) exp(...): evaluated
// on the host with libm so device libm differences cannot enter.
void launch_predictor(const LaunchCtx& L, const NsConst& c, const double* u, const double* v, const double* w,
                      const double* T, const double* su_tab, const double* sv_tab, double* us, double* vs,
                      double* ws);
void launch_divergence(const LaunchCtx& L, const NsConst& c, const double* us, const double* vs,
                       const double* ws, double* rhs);
void launch_corrector(const LaunchCtx& L, const NsConst& c, const double* us, const double* vs,
                      const double* ws, const double* p, double* u, double* v, double* w, const double* T,
                      double* Tn);
void launch_energy(const LaunchCtx& L, const NsConst& c, const double* u, const double* v, const double* w,
                   const double* T, double* Tn);
void launch_thermal_bc(const LaunchCtx& L, const ns_thermal_bc_config_t& bc, double* T);
void launch_periodic_bc(const LaunchCtx& L, double* f, int wrap_y, int wrap_z);
void launch_periodic_faces(const LaunchCtx& L, double* const* f, int nf, int wrap_y, int wrap_z);
void launch_field_stats(const LaunchCtx& L, const double* u, const double* v, const double* w,
                        const double* p, const double* T);
void launch_rk_stage(const LaunchCtx& L, const NsConst& c, const RkArgs& ra, const P4& qin, const P4& q0,
                      const W4& acc, const W4& qout, const double* rho, const double* T,
                      const double* su_tab, const double* sv_tab);
// explicit Euler step (interior cells, out of place) and the O(N^(2/3)) boundary helpers
void launch_euler(const LaunchCtx& L, const NsConst& c, const EulerArgs& ea, const P4& qin, const W4& qout,
                  const double* rho, const double* T, const double* su_tab, const double* sv_tab);
void launch_boundary_carry3(const LaunchCtx& L, const double* u, const double* v, const double* w, double* us,
                            double* vs, double* ws);
void launch_boundary_scan(const LaunchCtx& L, const double* u, const double* v, const double* w, const double* p);
void launch_synth_init(const LaunchCtx& L, uint64_t seed, long long gk0, long long gnz, double* u, double* v,
                       double* w, double* p, double* T, double* rho);
void launch_copy(const LaunchCtx& L, double* dst, const double* src, size_t n);
void launch_fill(const LaunchCtx& L, double* dst, double v, size_t n);
void launch_zero_ctl(const LaunchCtx& L);


#define CFDB_LAUNCH(kind, st, ...)                         \
    do {                                                   \
        cfdb::Profiler* pf_ = cfdb::t_prof;                \
        if (pf_ && pf_->on) pf_->begin(st);                \
        __VA_ARGS__;                                       \
        cfdb::g_launches++;                                \
        if (pf_) pf_->end(st, kind);                       \
    } while (0)

}  // namespace cfdb
