// common.cuh -- shared device/host definitions of libcfd_b200 (sm_100a, FP64).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <math.h>
#include <atomic>

#include "../../include/cfd_b200.h"

// Two builds of the stencil / Poisson kernel translation units (cfd_b200/build.py):
//   exact (default)  -fmad=false, the reference's divisions reproduced bit for bit (ediv);
//   fast (CFDB_FAST) -fmad=true (FMA contraction), quotients as reciprocal multiplies -- same formulas, last-bit
//                    different; selected per session (CFD_B200_ARITH=fast / cfd_b200_set_arithmetic).
// The fast twins live in the inline namespace cfdb::arith_fast so both sets of kernels and launchers link into
// one library under distinct symbols.
#ifdef CFDB_FAST
#define CFDB_ARITH_BEGIN inline namespace arith_fast {
#define CFDB_ARITH_END }
#else
#define CFDB_ARITH_BEGIN
#define CFDB_ARITH_END
#endif

namespace cfdb {

// ---------------------------------------------------------------------------------------
// Geometry of one slab resident on one GPU.  Layout is the reference's: SoA, x fastest,
// idx = k*nx*ny + j*nx + i (cfd/core/indexing.h:17-20); the outermost layer of cells on every
// side is either a physical boundary (ghost) layer or, along the decomposed axis, a halo
// plane owned by the neighbouring rank.
// ---------------------------------------------------------------------------------------
struct Geom {
    int nx, ny, nz;          // local extents including boundary / halo layers
    long long sy, sz;        // strides in doubles: sy = nx, sz = nx*ny (0 when nz == 1)
    int k0, k1;              // updated k range [k0,k1): 1..nz-1 in 3-D, 0..1 in 2-D
    int dim3;                // nz > 1
    double inv_2dx, inv_2dy, inv_2dz, inv_dx2, inv_dy2, inv_dz2;
    // divisors in the reference's own form: it DIVIDES by 2.0*dx, dx*dx (x, y) and multiplies by
    // the z reciprocals (solver_projection.c:80-84,120-160; linear_solver_jacobi.c:100-108), so
    // bit-exact parity needs the same operations.
    double two_dx, two_dy, dx2, dy2;
    double factor, inv_factor;  // 2(1/dx2+1/dy2+1/dz2) and its reciprocal (linear_solver_jacobi.c:62-63)
    // physical-boundary flags: glo[a]/ghi[a] == 1 when the low/high side of axis a (0=x,1=y,2=z)
    // is a true domain boundary (Neumann ghost / BC applies); 0 when it is a halo.
    int glo[3], ghi[3];
    // global index of local (0,0,0) -- for coordinate tables and red/black colouring
    long long gi0, gj0, gk0;
    long long ncell;         // nx*ny*nz
};

// Peer-memory transport (peer.cu): every rank owns one "box" in its device memory that its neighbours write
// into over NVLink (cudaDeviceEnablePeerAccess inside one process, cudaIpc handles across processes): halo
// mailboxes, reduction slots and sequence flags.  No library call on the data path.
constexpr int PEER_MAX_RANKS = 16;
constexpr int PEER_MAX_FIELDS = 6;
constexpr int PEER_RED_WORDS = 8;
struct PeerBoxHeader {                       // lives at offset 0 of a box; flags are monotone sequence numbers
    unsigned long long hflag[2][2];          // [side lo/hi][buffer]: halo planes of exchange #seq have landed
    unsigned long long rflag[PEER_MAX_RANKS][2];  // [source rank][buffer]: its reduction payload #seq has landed
    double rval[2][PEER_MAX_RANKS][PEER_RED_WORDS];
    unsigned ticket;                         // last-block detection of the push kernel (self-resetting)
    unsigned pad[15];
    // ---- reductions finished INSIDE the producing kernel (block_reduce_finish): the block that completes the local
    //      reduction publishes it to every box, waits for the others' and combines in rank order ----
    unsigned long long fseq;                 // fused reductions this rank has issued (device-side counter)
    unsigned long long fflag[PEER_MAX_RANKS][2];
    double fval[2][PEER_MAX_RANKS][4];
    char* boxes[PEER_MAX_RANKS];             // every rank's box as mapped on THIS device (filled at connect time)
    int* timeout_flag;                       // &Ctl::comm_timeout of this rank
    int R, r;
};

// One reduction "slot" : up to 4 simultaneous sums (or maxes) finished by the last block.
struct RedBuf {
    double* partials;   // [max_blocks][4]
    unsigned* ticket;   // zero-initialised, self-resetting
    PeerBoxHeader* peer;  // non-null: finish the reduction ACROSS ranks as well (all-reduce fused into the kernel)
};

// Device-resident control block of an iterative solve: kernels read/derive everything they
// need from here, so the host never synchronises inside an iteration.
struct Ctl {
    double dot[32];       // raw reduced values; slot meaning is solver specific (see poisson.cuh)
    double tol;           // max(rel*r0, abs)
    double r0;            // initial residual norm
    double res;           // last residual norm the reference would report
    int done_seq;         // kernels launched with seq > done_seq are no-ops; INT_MAX while running
    int status;           // poisson_solver_status_t
    int iterations;       // as the reference reports them
    int result_buf;       // which ping-pong buffer holds x (Jacobi)
    int skip_final_bc;    // reference returned before its final Neumann BC (stagnation paths)
    int nonfinite;        // NaN/Inf flag of the time step (solver_projection.c:281-289)
    int comm_timeout;     // a peer-transport wait gave up (peer.cu): the step's answer is not to be trusted
    int pad[1];
    double stat[4];       // max |vel|, max |p|, max T, spare (solver_registry.c:31-62)
};

// Halo exchange overlapped with interior compute (peer transport): the exchange's PULL half handed to the
// first kernel that consumes the exchanged field.  The blocks whose chunk touches a halo layer wait for the neighbour's
// sequence flag and copy their part of the mailbox layer into the field's halo layer before their first load; all
// other blocks start at once, and the edge chunks are scheduled last (gate_chunk), so the neighbour's push and the
// NVLink transfer run under the interior of the sweep.  Without a gate the exchange is push + pull kernels.
struct HaloGate {
    int on = 0;                                   // 0: nothing pending
    int nf = 0;                                   // fields in this exchange (<= 2 on the Poisson path)
    long long cnt = 0;                            // doubles per layer (a plane in 3-D, a row in 2-D)
    unsigned long long seq = 0;                   // the exchange's sequence number
    const double* box[2] = {nullptr, nullptr};    // my lo / hi mailbox of this exchange (null: no neighbour there)
    const unsigned long long* flag[2] = {nullptr, nullptr};
    double* halo[2][2] = {{nullptr, nullptr}, {nullptr, nullptr}};  // [side][field]: where the mailbox layer belongs
    int* timeout_flag = nullptr;
};

// Per-launch arguments of the Poisson kernels (everything the host knows at launch time).
struct KArgs {
    int seq;              // launch sequence number inside this solve (monotone)
    int it;               // iteration (or sweep) index
    int check;            // check_interval
    int max_iter;
    double rel_tol, abs_tol;
    double omega;         // SOR relaxation factor
    double dinv;          // Jacobi preconditioner 1/diag (linear_solver_cg.c:216)
    int zchunk;           // planes marched per block
    HaloGate gate;        // pending pull of the operand's halo layers (this launch only)
};

#define CFDB_CUDA_OK(expr)                                                                  \
    do {                                                                                    \
        cudaError_t e_ = (expr);                                                            \
        if (e_ != cudaSuccess) {                                                            \
            cfdb::set_error(CFD_ERROR, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), \
                            __FILE__, __LINE__);                                            \
            return CFD_ERROR;                                                               \
        }                                                                                   \
    } while (0)

void set_error(cfd_status_t st, const char* fmt, ...);

// Function attributes (dynamic shared-memory opt-in, carveout, cluster size) are PER DEVICE: a process that drives
// several GPUs -- the in-process gang, or two solver objects on two devices -- must set them on each one.  Runs the
// statements once per device ordinal (re-running them is harmless, so no lock: two threads may both do it).
#define CFDB_ONCE_PER_DEVICE(...)                                                   \
    do {                                                                            \
        static std::atomic<unsigned long long> done_{0};                            \
        int dev_ = 0;                                                               \
        cudaGetDevice(&dev_);                                                       \
        const unsigned long long bit_ = 1ull << (dev_ & 63);                        \
        if (!(done_.load(std::memory_order_acquire) & bit_)) {                      \
            __VA_ARGS__;                                                            \
            done_.fetch_or(bit_, std::memory_order_release);                        \
        }                                                                           \
    } while (0)
extern std::atomic<unsigned long long> g_launches;  // kernels launched by this library (any thread)

// ---------------------------------------------------------------------------------------
// Device helpers
// ---------------------------------------------------------------------------------------
#ifdef __CUDACC__

// The reference clamps with fmax(-lim, fmin(lim, v)) (NaN -> +lim).  sm_100a has no FP64 min/max
// instruction: each fmin/fmax expands to ~7 instructions, two of them on the FP64 pipe, plus a live
// predicate (measured: a quarter of the RK stage's instruction stream).  For limits whose low word
// is zero (1, 10, 100, 1000) the same function is 6 instructions (5 integer + one FP64 compare):
// |v| >= lim or NaN  <=>  (high word without sign) >= (high word of lim); the result is then +-lim
// with v's sign, +lim for NaN.  Identical results for every input (tests/test_gpu_steppers.py).
// (Quotients come from ediv(), which maps an INFINITE numerator to NaN where a / b is +-inf: with
// +-inf in a field the clamp can return +lim where the reference returns -lim.  NaN inputs are exact.)
constexpr unsigned HI_1 = 0x3ff00000u, HI_10 = 0x40240000u, HI_100 = 0x40590000u, HI_1000 = 0x408f4000u, HI_INF = 0x7ff00000u;
__device__ __forceinline__ unsigned abs_hi(double v) { return (unsigned)__double2hiint(v) & 0x7fffffffu; }
template <unsigned LIM_HI>
__device__ __forceinline__ double clampb(double v) {
    const unsigned hi = (unsigned)__double2hiint(v), lo = (unsigned)__double2loint(v);
    const bool over = (hi & 0x7fffffffu) >= LIM_HI;                     // |v| >= lim, or NaN
    const unsigned lim = (v < 0.0) ? (LIM_HI | 0x80000000u) : LIM_HI;  // -lim only for a real negative: NaN -> +lim
    return __hiloint2double((int)(over ? lim : hi), (int)(over ? 0u : lo));
}
__device__ __forceinline__ double clamp100(double v) { return clampb<HI_100>(v); }
// N clamps against the same limit.  In a converged flow nothing is near a limit, so the usual case is
// "no value of any lane of the warp is over": 2 integer instructions per value (mask + running max)
// and one warp-uniform branch instead of 6 per value; the exact clamps run only when some lane needs
// them.  Same results as clampb<> on every value (it is only skipped where it would return v).
template <unsigned LIM_HI, int N>
__device__ __forceinline__ void clamp_group(double (&v)[N]) {
    unsigned m = 0u;
#pragma unroll
    for (int q = 0; q < N; q++) m = max(m, abs_hi(v[q]));
    if (__any_sync(__activemask(), m >= LIM_HI)) {
#pragma unroll
        for (int q = 0; q < N; q++) v[q] = clampb<LIM_HI>(v[q]);
    }
}
// fmin(v, 1.0) of the reference (NaN -> 1.0) without the FP64 min expansion
__device__ __forceinline__ double min_one(double v) { return (v <= 1.0) ? v : 1.0; }
// running maximum in the reference's own form `if (r > m) m = r;` (linear_solver.c:304-346): a NaN
// candidate is ignored like with fmax, without fmax's NaN-selection sequence
__device__ __forceinline__ void max_into(double& m, double r) {
    if (r > m) m = r;
}

// a / b, correctly rounded, from y = RN(1/b) computed once on the host: q = a*y, one exact
// residual, one correction (Markstein).  Same result as the IEEE division the reference performs
// (checked on 7.6e8 random operands, tests/test_oracle.py::test_exact_division_sequence) at 3
// instructions instead of the ~14 of the generic FP64 division sequence.  Differs only for
// non-finite a (NaN instead of Inf; both are caught by the same scans) and denormal quotients.
__device__ __forceinline__ double ediv(double a, double b, double y) {
#ifdef CFDB_FAST
    (void)b;
    return a * y;
#endif
    const double q = a * y;
    const double r = fma(-b, q, a);
    return fma(r, y, q);
}

// Quotients a / rho for several numerators of one cell (RK / Euler right-hand side: mu/rho and the
// three pressure-gradient terms, ns_momentum_rhs_scalar.h:118,160-175): one IEEE reciprocal
// y = RN(1/rho), then the exact-division sequence per numerator -- 1 division + 3 instructions each
// instead of 4 divisions (~14 instructions each).  Densities outside (1e-10, 1e100] take the plain
// division (the reciprocal could lose bits there).
struct RhoDiv {
    double rh, y;
    bool plain;
    __device__ __forceinline__ explicit RhoDiv(double rho) : rh(rho), y(1.0 / rho), plain(!(rho <= 1e100)) {}
#ifdef CFDB_FAST
    __device__ __forceinline__ double operator()(double a) const { return a * y; }
#else
    __device__ __forceinline__ double operator()(double a) const { return plain ? a / rh : ediv(a, rh, y); }
#endif
};

// ---- cross-GPU flags (peer transport): system-scope release / acquire, bounded spinning ----
constexpr long long SPIN_LIMIT_CYCLES = 8000000000LL;  // ~4 s at 2 GHz
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
// (timeout_flag[0] = which wait gave up first: 1 halo lo, 2 halo hi, 16 + r reduction from rank r, 48 + r fused
//  reduction from rank r; [1] = its sequence number)
__device__ __forceinline__ bool wait_flag(const unsigned long long* p, unsigned long long want, int* timeout_flag,
                                          int code) {
    if (*(volatile int*)timeout_flag) return false;  // an earlier wait of this step already gave up
    const long long t0 = clock64();
    while (ld_acquire_sys(p) < want) {
        if (clock64() - t0 > SPIN_LIMIT_CYCLES) {
            if (atomicCAS(timeout_flag, 0, code) == 0) timeout_flag[1] = (int)want;
            return false;
        }
        __nanosleep(64);
    }
    return true;
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// Block-wide reduction of NV values followed by the deterministic "last block finishes"
// pattern: every block stores its partial, the block that draws the last ticket adds all
// partials in a fixed order and writes out[0..NV).  No atomics on doubles, no host sync.
// IS_MAX selects max instead of sum.  All threads of the block must call it.
template <int NV, bool IS_MAX>
__device__ __forceinline__ void block_reduce_finish(double (&v)[NV], const RedBuf rb, double* out) {
    __shared__ double sh[NV][32];
    __shared__ bool last;
    const int tid = threadIdx.y * blockDim.x + threadIdx.x;
    const int nthr = blockDim.x * blockDim.y;
    const int lane = tid & 31, wid = tid >> 5, nw = (nthr + 31) >> 5;
#pragma unroll
    for (int q = 0; q < NV; q++) {
        double r = IS_MAX ? warp_max(v[q]) : warp_sum(v[q]);
        if (lane == 0) sh[q][wid] = r;
    }
    __syncthreads();
    const unsigned nblocks = gridDim.x * gridDim.y * gridDim.z;
    const unsigned bid = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
    if (wid == 0) {
#pragma unroll
        for (int q = 0; q < NV; q++) {
            double r = (lane < nw) ? sh[q][lane] : (IS_MAX ? -INFINITY : 0.0);
            r = IS_MAX ? warp_max(r) : warp_sum(r);
            if (lane == 0) rb.partials[(size_t)bid * 4 + q] = r;
        }
        if (lane == 0) {
            __threadfence();
            unsigned t = atomicAdd(rb.ticket, 1u);
            last = (t == nblocks - 1);
        }
    }
    __syncthreads();
    if (!last) return;
    __threadfence();
    double acc[NV];
#pragma unroll
    for (int q = 0; q < NV; q++) acc[q] = IS_MAX ? -INFINITY : 0.0;
    for (unsigned b = tid; b < nblocks; b += nthr) {
#pragma unroll
        for (int q = 0; q < NV; q++) {
            double pv = __ldcg(&rb.partials[(size_t)b * 4 + q]);
            acc[q] = IS_MAX ? fmax(acc[q], pv) : acc[q] + pv;
        }
    }
    __syncthreads();  // sh reuse
#pragma unroll
    for (int q = 0; q < NV; q++) {
        double r = IS_MAX ? warp_max(acc[q]) : warp_sum(acc[q]);
        if (lane == 0) sh[q][wid] = r;
    }
    __syncthreads();
    if (wid == 0) {
        double loc[NV];
#pragma unroll
        for (int q = 0; q < NV; q++) {
            double r = (lane < nw) ? sh[q][lane] : (IS_MAX ? -INFINITY : 0.0);
            loc[q] = IS_MAX ? warp_max(r) : warp_sum(r);
        }
        if (rb.peer) {
            // All-reduce fused into the producing kernel: lane t serves rank t.  My local result goes into slot [me] of
            // every box (NVLink stores + system-scope release of a sequence flag); then the R flags in my own box are
            // awaited and the R values combined in RANK ORDER -- the same numbers in the same order on every rank, so
            // all ranks hold the same bits (same result as the stand-alone k_allreduce_peer, minus its launch).
            PeerBoxHeader* me = rb.peer;
            const int R = me->R, rk = me->r;
            unsigned long long seq = 0;
            if (lane == 0) seq = ++me->fseq;
            seq = __shfl_sync(0xffffffffu, seq, 0);
            const int buf = (int)(seq & 1);
            if (lane < R) {
                PeerBoxHeader* h = reinterpret_cast<PeerBoxHeader*>(me->boxes[lane]);
#pragma unroll
                for (int q = 0; q < NV; q++) h->fval[buf][rk][q] = loc[q];
                __threadfence_system();
                st_release_sys(&h->fflag[rk][buf], seq);
            }
            bool ok = true;
            if (lane < R) ok = wait_flag(&me->fflag[lane][buf], seq, me->timeout_flag, 48 + lane);
            ok = __all_sync(0xffffffffu, ok);
            if (ok && lane == 0) {
                const volatile double* fv = &me->fval[buf][0][0];
#pragma unroll
                for (int q = 0; q < NV; q++) {
                    double acc = fv[q];
                    for (int t = 1; t < R; t++) {
                        const double x = fv[t * 4 + q];
                        if (IS_MAX) { if (x > acc) acc = x; }
                        else acc = acc + x;
                    }
                    loc[q] = acc;
                }
            }
        }
        if (lane == 0) {
#pragma unroll
            for (int q = 0; q < NV; q++) out[q] = loc[q];
            *rb.ticket = 0u;
            __threadfence();
        }
    }
}

// ---- boundary-cell enumeration: x faces, y faces, physical z faces (edges visited twice) ----
__device__ __forceinline__ long long face_cell(const Geom& g, long long t) {
    const long long nyz = (long long)g.ny * g.nz, nxz = (long long)g.nx * g.nz, nxy = (long long)g.nx * g.ny;
    if (t < 2 * nyz) {  // x faces
        const int side = (int)(t / nyz);
        const long long r = t % nyz;
        return (r / g.ny) * nxy + (r % g.ny) * g.nx + (side ? g.nx - 1 : 0);
    }
    t -= 2 * nyz;
    if (t < 2 * nxz) {  // y faces
        const int side = (int)(t / nxz);
        const long long r = t % nxz;
        return (r / g.nx) * nxy + (long long)(side ? g.ny - 1 : 0) * g.nx + (r % g.nx);
    }
    t -= 2 * nxz;
    if (g.dim3 && t < 2 * nxy) {  // z faces (physical ones only: halo planes belong to the exchange)
        const int side = (int)(t / nxy);
        if ((side == 0 && !g.glo[2]) || (side == 1 && !g.ghi[2])) return -1;
        return (long long)(side ? g.nz - 1 : 0) * nxy + (t % nxy);
    }
    return -1;
}
__host__ __device__ __forceinline__ long long face_cells(const Geom& g) {
    return 2LL * ((long long)g.ny * g.nz + (long long)g.nx * g.nz + (g.dim3 ? (long long)g.nx * g.ny : 0));
}


#endif  // __CUDACC__

}  // namespace cfdb
