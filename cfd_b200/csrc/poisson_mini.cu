// poisson_mini.cu -- whole Jacobi solve of a SMALL 2-D grid in one launch (sm_100a thread-block cluster).
//
// The 129 x 129 Ghia cavity (BASELINE config C1) runs up to 2000 Jacobi sweeps per time step on
// 16 641 cells: one kernel per sweep costs ~10 us of launch + last-block reduction for ~1 us of
// work.  Here ONE cluster of up to 8 CTAs keeps the whole problem on chip for the whole solve:
//   * every CTA owns a strip of rows; x (two buffers) and rhs of the strip live in its shared memory;
//   * a sweep reads buffer q&1 and writes buffer (q+1)&1 with the Neumann ghosts (same closed form as
//     store_with_ghosts), then pushes its first / last row into the neighbouring CTAs' halo rows
//     through distributed shared memory and meets the others at ONE cluster barrier;
//   * the L-inf residual of x_q comes from the same loads (only on the sweeps where the reference
//     computes it); every CTA writes its partial into every CTA's slot array, so after the barrier
//     all of them take the reference's decision (linear_solver.c:415-484) on identical numbers.
// Arithmetic, operation order, stopping rule, iteration count and status are those of k_jacobi /
// stationary_decide (hence of linear_solver_jacobi.c:76-129): results are bit-identical to the
// per-sweep kernels (tests/test_gpu_poisson.py::test_jacobi_cluster_solve_is_bit_exact).
#include "common.cuh"
#include "kernels.h"
#include "tma.cuh"

#include <cooperative_groups.h>
#include <limits.h>
#include <stdlib.h>

namespace cg = cooperative_groups;

namespace cfdb {

namespace {

constexpr int MINI_THREADS = 512;
constexpr int MINI_MAX_CTAS = 16;          // slot-array stride; clusters above 8 CTAs are the non-portable size (opt-in)
constexpr size_t MINI_MAX_SMEM = 200 * 1024;

struct MiniPlan {
    int nc;        // CTAs in the cluster
    int rows_per;  // interior rows per CTA (the last one may own fewer, never zero)
    int bx, by;    // thread block shape
    int cpt;       // interior cells per thread
    size_t smem;
};

// Shared memory of one CTA.
//   Jacobi: A[(rows_per+2)*nx], B[(rows_per+2)*nx], slot[2][MINI_MAX_CTAS]   (rhs lives in registers)
//   CG    : p[(rows_per+2)*nx], x[rows_per*nx], r[rows_per*nx], Ap[rows_per*nx], slot[3][MINI_MAX_CTAS]
static size_t mini_smem(int nx, int rows_per, bool cg) {
    if (cg) return ((size_t)(4 * rows_per + 2) * nx + 3 * MINI_MAX_CTAS) * sizeof(double);
    return ((size_t)(2 * (rows_per + 2)) * nx + 2 * MINI_MAX_CTAS) * sizeof(double);
}

static bool mini_plan(const Geom& g, MiniPlan& p, bool cg = false) {
    if (g.dim3 || g.nx < 3 || g.ny < 3) return false;
    if (!(g.glo[0] && g.ghi[0] && g.glo[1] && g.ghi[1])) return false;  // single rank only
    const int ni = g.ny - 2;
    // measured on B200: a cluster barrier costs ~1.2 us per sweep, a strip ~0.3 us per cell-per-thread round:
    // grids up to ~1500 interior cells are faster on ONE CTA (block barrier only)
    // 16 CTAs (the non-portable cluster size, opted into per launch) halve the work per SM: measured 25-30 %
    // faster than 8 on 129^2 / 200^2.  CFD_B200_CLUSTER_CTAS=8|4|2 selects a smaller cluster (read per solve).
    int want = 16;
    if (const char* e = getenv("CFD_B200_CLUSTER_CTAS")) {
        const int v = atoi(e);
        if (v == 16 || v == 8 || v == 4 || v == 2) want = v;
    }
    int nc = ((long long)ni * (g.nx - 2) <= 1536) ? 1 : want;
    while (nc > 1 && ((ni + nc - 1) / nc) * (nc - 1) >= ni) nc >>= 1;  // every CTA owns at least one row
    p.nc = nc;
    p.rows_per = (ni + nc - 1) / nc;
    p.smem = mini_smem(g.nx, p.rows_per, cg);
    if (p.smem > MINI_MAX_SMEM) return false;
    p.cpt = (p.rows_per * (g.nx - 2) + MINI_THREADS - 1) / MINI_THREADS;
    if (!cg && p.cpt > 16) return false;  // Jacobi keeps a thread's cells in registers
    int bx = 32;
    while (bx < 128 && bx < g.nx - 2) bx <<= 1;
    p.bx = bx;
    p.by = MINI_THREADS / bx;
    return true;
}

// Barrier over the cluster; a one-CTA "cluster" only needs the block barrier (the hardware cluster barrier
// costs about a microsecond even then).
__device__ __forceinline__ void cluster_barrier(cg::cluster_group& cluster, int nc) {
    if (nc > 1) cluster.sync();
    else __syncthreads();
}

__device__ __forceinline__ double block_max(double v, double* scratch) {
    // max in the reference's form (NaN candidates are ignored), warp shuffle -> shared -> warp 0
    const int tid = threadIdx.y * blockDim.x + threadIdx.x, lane = tid & 31, wid = tid >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) max_into(v, __shfl_xor_sync(0xffffffffu, v, o));
    if (lane == 0) scratch[wid] = v;
    __syncthreads();
    double r = 0.0;
    if (wid == 0) {
        r = (lane < (int)((blockDim.x * blockDim.y) >> 5)) ? scratch[lane] : 0.0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) max_into(r, __shfl_xor_sync(0xffffffffu, r, o));
    }
    return r;  // valid in thread 0
}

// One Jacobi sweep over a thread's cells: reads A (ghosts / halos valid), writes B with its Neumann ghosts,
// optionally the L-inf residual of A.  Same expressions as k_jacobi (poisson_kernels.cu).
template <int CPT>
__device__ __forceinline__ void jacobi_cells(const Geom& g, const double* A, double* B, const int (&cell)[CPT],
                                             const double (&f)[CPT], bool checked, double& rmax) {
    const int nx = g.nx;
    constexpr int BATCH = CPT < 4 ? CPT : 4;  // cells whose loads are in flight together
#pragma unroll
        for (int c0 = 0; c0 < CPT; c0 += BATCH) {
            double xe[BATCH], xw[BATCH], yn[BATCH], ys[BATCH], zc[BATCH];
#pragma unroll
            for (int b = 0; b < BATCH; b++) {
                const int n = cell[c0 + b] < 0 ? nx + 1 : (cell[c0 + b] & 0xffffff);  // idle slots read a valid cell, store nothing
                zc[b] = A[n];
                xe[b] = A[n + 1];
                xw[b] = A[n - 1];
                yn[b] = A[n + nx];
                ys[b] = A[n - nx];
            }
#pragma unroll
            for (int b = 0; b < BATCH; b++) {
                const int c = c0 + b;
                const double zm = 0.0, zp = 0.0;
                if (checked && cell[c] >= 0) {
                    const double lap = ediv(xe[b] - 2.0 * zc[b] + xw[b], g.dx2, g.inv_dx2) +
                                       ediv(yn[b] - 2.0 * zc[b] + ys[b], g.dy2, g.inv_dy2) + (zp + zm - 2.0 * zc[b]) * g.inv_dz2;
                    max_into(rmax, fabs(lap - f[c]));
                }
                const double v = -(f[c] - ediv(xe[b] + xw[b], g.dx2, g.inv_dx2) - ediv(yn[b] + ys[b], g.dy2, g.inv_dy2) -
                                   (zp + zm) * g.inv_dz2) * g.inv_factor;
                if (cell[c] >= 0) {
                    // interior cell + every Neumann ghost whose nearest interior cell it is
                    const int n = cell[c] & 0xffffff, d = cell[c] >> 24;
                    B[n] = v;
                    if (d) {
                        const bool xl = d & 1, xh = d & 2, yl = d & 4, yh = d & 8;
                        if (xl) B[n - 1] = v;
                        if (xh) B[n + 1] = v;
                        if (yl) {
                            B[n - nx] = v;
                            if (xl) B[n - nx - 1] = v;
                            if (xh) B[n - nx + 1] = v;
                        }
                        if (yh) {
                            B[n + nx] = v;
                            if (xl) B[n + nx - 1] = v;
                            if (xh) B[n + nx + 1] = v;
                        }
                    }
                }
            }
        }
}

// CPT = cells per thread (compile time): a thread's cells, their ghost duties and their rhs values are fixed
// for the whole solve and live in registers, so a sweep is straight-line code -- all shared loads of the
// thread's cells in flight together, no index arithmetic, no inner loops (the loop-nest version spent ~500
// instructions per warp and sweep on a 129^2 grid; ncu: latency-bound at 2.4 us per sweep).
template <int CPT>
__global__ void __launch_bounds__(MINI_THREADS, 1)
k_jacobi_cluster(Geom g, double* __restrict__ x, const double* __restrict__ rhs_g, Ctl* ctl, KArgs a, int rows_per) {
    extern __shared__ __align__(16) double sm[];
    __shared__ double red_scratch[MINI_THREADS / 32];
    cg::cluster_group cluster = cg::this_cluster();
    const int rank = (int)cluster.block_rank(), nc = (int)cluster.num_blocks();
    const int nx = g.nx, ny = g.ny;
    const int jb = 1 + rank * rows_per, je = min(jb + rows_per, ny - 1);  // owned interior rows [jb, je)
    const int myrows = je - jb;
    const size_t bufsz = (size_t)(rows_per + 2) * nx;
    double* const buf0 = sm;
    double* const buf1 = sm + bufsz;
    double* slot = sm + 2 * bufsz;  // [2][MINI_MAX_CTAS]
    const int tid = threadIdx.x, nthr = blockDim.x;

    // ---- load the strip: rows jb-1 .. je of x as the caller left them (ghosts included)
    for (int t = tid; t < (myrows + 2) * nx; t += nthr) buf0[t] = x[(size_t)(jb - 1) * nx + t];
    // ---- this thread's cells: shared offset (bits 0..23; < 0: none) + ghost duties (bit 24 xl, 25 xh, 26 yl, 27 yh), rhs
    int cell[CPT];
    double f[CPT];
    const int wi = nx - 2, mine = myrows * wi;
#pragma unroll
    for (int c = 0; c < CPT; c++) {
        const int idx = tid + c * nthr;
        cell[c] = -1;
        f[c] = 0.0;
        if (idx < mine) {
            const int l = 1 + idx / wi, i = 1 + idx % wi, j = jb + l - 1;
            cell[c] = (l * nx + i) | ((i == 1 ? 1 : 0) | (i == nx - 2 ? 2 : 0) | (j == 1 ? 4 : 0) | (j == ny - 2 ? 8 : 0)) << 24;
            f[c] = rhs_g[(size_t)j * nx + i];
        }
    }
    // (the second buffer needs no initialisation: each sweep produces every cell of the strip it later
    // reads -- own rows with their ghost columns, the ghost rows, the halo rows pushed by the neighbours)
    cluster_barrier(cluster, nc);  // all CTAs of the cluster are running before anyone addresses a neighbour's shared memory

    double* const remote_lo = (rank > 0) ? cluster.map_shared_rank(sm, rank - 1) : nullptr;       // CTA below
    double* const remote_hi = (rank < nc - 1) ? cluster.map_shared_rank(sm, rank + 1) : nullptr;  // CTA above
    double r0 = 0.0, tol = 0.0, res_rep = 0.0;
    int status = POISSON_ERROR, iters = 0, result = 0;
    int until_check = 0;  // sweeps until the next one whose residual the reference looks at: s = 0, 1, 1+check, ...
    for (int s = 0;; s++) {
        const double* A = (s & 1) ? buf1 : buf0;
        double* B = (s & 1) ? buf0 : buf1;
        const bool checked = (until_check == 0);
        if (checked) until_check = (s == 0) ? 1 : a.check;
        until_check--;
        double rmax = 0.0;
        jacobi_cells<CPT>(g, A, B, cell, f, checked, rmax);
        if (checked) {
            const double m = block_max(rmax, red_scratch);
            if (tid == 0)
                for (int r = 0; r < nc; r++) cluster.map_shared_rank(slot, r)[(s & 1) * MINI_MAX_CTAS + rank] = m;
        }
        __syncthreads();  // the strip's rows of B are complete: hand the edge rows to the neighbours
        const size_t boff = (s & 1) ? 0 : bufsz;  // offset of B inside a CTA's shared memory
        if (remote_lo)    // my first row -> upper halo row of the CTA below (it owns rows_per rows)
            for (int t = tid; t < nx; t += nthr) remote_lo[boff + (size_t)(rows_per + 1) * nx + t] = B[nx + t];
        if (remote_hi)    // my last row -> lower halo row of the CTA above
            for (int t = tid; t < nx; t += nthr) remote_hi[boff + t] = B[(size_t)myrows * nx + t];
        cluster_barrier(cluster, nc);  // release/acquire at cluster scope: halo rows and residual slots are visible

        // ---- the reference's decision about x_s (linear_solver.c:415-484), identical in every CTA
        bool fin = false;
        if (checked) {
            double res = 0.0;
            for (int r = 0; r < nc; r++) max_into(res, slot[(s & 1) * MINI_MAX_CTAS + r]);
            res_rep = res;
            if (s == 0) {
                r0 = res;
                tol = fmax(a.rel_tol * res, a.abs_tol);
                if (res < a.abs_tol) { fin = true; status = POISSON_CONVERGED; iters = 0; }
            } else if (res < tol || res < a.abs_tol) {
                fin = true; status = POISSON_CONVERGED; iters = s;
            }
        }
        if (!fin && s > 0 && s >= a.max_iter) {
            fin = true; status = POISSON_MAX_ITER; iters = a.max_iter + 1;  // reference reports iter+1 (:478)
        }
        if (fin) { result = s & 1; break; }
    }

    // ---- x_s back to global memory: own rows, plus the ghost rows on the first / last CTA
    {
        const double* X = result ? buf1 : buf0;
        const int l0 = (rank == 0) ? 0 : 1, l1 = (rank == nc - 1) ? myrows + 1 : myrows;
        for (int t = tid + l0 * nx; t < (l1 + 1) * nx; t += nthr) x[(size_t)(jb - 1) * nx + t] = X[t];
    }
    if (rank == 0 && tid == 0) {
        ctl->r0 = r0;
        ctl->tol = tol;
        ctl->res = res_rep;
        ctl->status = status;
        ctl->iterations = iters;
        ctl->result_buf = 0;
        ctl->skip_final_bc = 0;
        ctl->done_seq = 0;
    }
    cluster_barrier(cluster, nc);  // no CTA exits while a neighbour may still address its shared memory
}

// ---- neighbour-only signalling instead of a cluster barrier per sweep (default; CFD_B200_CLUSTER_SYNC=1 = barrier) ----
// cluster.sync() releases with MEMBAR.ALL.GPU and acquires with CCTL.IVALL: most of a microsecond per sweep.
// Here every halo value travels as one st.async (STAS) that lands in the neighbour's shared memory AND
// completes 8 bytes of transaction count on the neighbour's mbarrier; a CTA starts sweep s+1 when the two
// halo rows of sweep s have arrived.  No other synchronisation is needed: a neighbour can only overwrite a
// halo row of buffer b after it received my push of the sweep in which I last read b.  Residual partials of
// the checked sweeps go the same way to every CTA's slot array (second mbarrier pair).
__device__ __forceinline__ uint32_t mapa_u32(uint32_t addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void st_async_f64(uint32_t remote_addr, double v, uint32_t remote_bar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b64 [%0], %1, [%2];" ::"r"(remote_addr),
                 "l"(__double_as_longlong(v)), "r"(remote_bar)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAITC_%=:\n"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONEC_%=;\n"
        "bra WAITC_%=;\n"
        "DONEC_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}

template <int CPT>
__global__ void __launch_bounds__(MINI_THREADS, 1)
k_jacobi_cluster_async(Geom g, double* __restrict__ x, const double* __restrict__ rhs_g, Ctl* ctl, KArgs a, int rows_per) {
    extern __shared__ __align__(16) double sm[];
    __shared__ double red_scratch[MINI_THREADS / 32];
    __shared__ __align__(8) uint64_t bars[4];  // [0,1] halo rows of even / odd sweeps, [2,3] residual partials
    cg::cluster_group cluster = cg::this_cluster();
    const int rank = (int)cluster.block_rank(), nc = (int)cluster.num_blocks();
    const int nx = g.nx, ny = g.ny;
    const int jb = 1 + rank * rows_per, je = min(jb + rows_per, ny - 1);
    const int myrows = je - jb;
    const size_t bufsz = (size_t)(rows_per + 2) * nx;
    double* const buf0 = sm;
    double* const buf1 = sm + bufsz;
    double* slot = sm + 2 * bufsz;  // [2][MINI_MAX_CTAS]
    const int tid = threadIdx.x, nthr = blockDim.x;
    const uint32_t halo_bytes = (uint32_t)(((rank > 0) ? nx : 0) + ((rank < nc - 1) ? nx : 0)) * 8u;
    if (tid == 0) {
        for (int q = 0; q < 4; q++) mbar_init(&bars[q], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        mbar_expect_tx(&bars[0], halo_bytes);
        mbar_expect_tx(&bars[1], halo_bytes);
        mbar_expect_tx(&bars[2], (uint32_t)nc * 8u);
        mbar_expect_tx(&bars[3], (uint32_t)nc * 8u);
    }
    for (int t = tid; t < (myrows + 2) * nx; t += nthr) buf0[t] = x[(size_t)(jb - 1) * nx + t];
    int cell[CPT];
    double f[CPT];
    const int wi = nx - 2, mine = myrows * wi;
#pragma unroll
    for (int c = 0; c < CPT; c++) {
        const int idx = tid + c * nthr;
        cell[c] = -1;
        f[c] = 0.0;
        if (idx < mine) {
            const int l = 1 + idx / wi, i = 1 + idx % wi, j = jb + l - 1;
            cell[c] = (l * nx + i) | ((i == 1 ? 1 : 0) | (i == nx - 2 ? 2 : 0) | (j == 1 ? 4 : 0) | (j == ny - 2 ? 8 : 0)) << 24;
            f[c] = rhs_g[(size_t)j * nx + i];
        }
    }
    cluster.sync();  // strips loaded, barriers initialised and armed in every CTA

    const uint32_t sm_lo = (rank > 0) ? mapa_u32(smem_u32(sm), rank - 1) : 0u;
    const uint32_t sm_hi = (rank < nc - 1) ? mapa_u32(smem_u32(sm), rank + 1) : 0u;
    const uint32_t bar_lo = (rank > 0) ? mapa_u32(smem_u32(&bars[0]), rank - 1) : 0u;
    const uint32_t bar_hi = (rank < nc - 1) ? mapa_u32(smem_u32(&bars[0]), rank + 1) : 0u;
    double r0 = 0.0, tol = 0.0, res_rep = 0.0;
    int status = POISSON_ERROR, iters = 0, result = 0;
    int until_check = 0, nchk = 0;
    for (int s = 0;; s++) {
        const double* A = (s & 1) ? buf1 : buf0;
        double* B = (s & 1) ? buf0 : buf1;
        const bool checked = (until_check == 0);
        if (checked) until_check = (s == 0) ? 1 : a.check;
        until_check--;
        double rmax = 0.0;
        jacobi_cells<CPT>(g, A, B, cell, f, checked, rmax);
        if (checked) {
            const double m = block_max(rmax, red_scratch);
            if (tid == 0) {
                const uint32_t so = smem_u32(slot + (nchk & 1) * MINI_MAX_CTAS + rank), bo = smem_u32(&bars[2 + (nchk & 1)]);
                for (int r = 0; r < nc; r++) st_async_f64(mapa_u32(so, r), m, mapa_u32(bo, r));
            }
        }
        __syncthreads();  // every thread is done reading A and writing B
        {
            const uint32_t boff = (uint32_t)(((s & 1) ? 0 : bufsz) * 8);
            const uint32_t hb = (uint32_t)((s & 1) * 8);  // which halo barrier: sweep parity
            if (rank > 0)
                for (int t = tid; t < nx; t += nthr)
                    st_async_f64(sm_lo + boff + (uint32_t)(((size_t)(rows_per + 1) * nx + t) * 8), B[nx + t], bar_lo + hb);
            if (rank < nc - 1)
                for (int t = tid; t < nx; t += nthr)
                    st_async_f64(sm_hi + boff + (uint32_t)(t * 8), B[(size_t)myrows * nx + t], bar_hi + hb);
        }
        if (halo_bytes) {
            mbar_wait_cluster(&bars[s & 1], (uint32_t)((s >> 1) & 1));
            if (tid == 0) mbar_expect_tx(&bars[s & 1], halo_bytes);  // re-arm for sweep s + 2
        }
        bool fin = false;
        if (checked) {
            mbar_wait_cluster(&bars[2 + (nchk & 1)], (uint32_t)((nchk >> 1) & 1));
            double res = 0.0;
            for (int r = 0; r < nc; r++) max_into(res, slot[(nchk & 1) * MINI_MAX_CTAS + r]);
            __syncwarp();
            if (tid == 0) mbar_expect_tx(&bars[2 + (nchk & 1)], (uint32_t)nc * 8u);
            nchk++;
            res_rep = res;
            if (s == 0) {
                r0 = res;
                tol = fmax(a.rel_tol * res, a.abs_tol);
                if (res < a.abs_tol) { fin = true; status = POISSON_CONVERGED; iters = 0; }
            } else if (res < tol || res < a.abs_tol) {
                fin = true; status = POISSON_CONVERGED; iters = s;
            }
        }
        if (!fin && s > 0 && s >= a.max_iter) {
            fin = true; status = POISSON_MAX_ITER; iters = a.max_iter + 1;
        }
        if (fin) { result = s & 1; break; }
    }
    __syncthreads();
    {
        const double* X = result ? buf1 : buf0;
        const int l0 = (rank == 0) ? 0 : 1, l1 = (rank == nc - 1) ? myrows + 1 : myrows;
        for (int t = tid + l0 * nx; t < (l1 + 1) * nx; t += nthr) x[(size_t)(jb - 1) * nx + t] = X[t];
    }
    if (rank == 0 && tid == 0) {
        ctl->r0 = r0;
        ctl->tol = tol;
        ctl->res = res_rep;
        ctl->status = status;
        ctl->iterations = iters;
        ctl->result_buf = 0;
        ctl->skip_final_bc = 0;
        ctl->done_seq = 0;
    }
    cluster.sync();  // no CTA exits while a neighbour may still address its shared memory
}

// Sum of NV per-thread partials over the block (fixed order: lane tree, then warps 0..15), valid in thread 0.
template <int NV>
__device__ __forceinline__ void block_sum(double (&v)[NV], double* scratch) {
    const int tid = threadIdx.y * blockDim.x + threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int nw = (int)((blockDim.x * blockDim.y) >> 5);
#pragma unroll
    for (int q = 0; q < NV; q++) {
        v[q] = warp_sum(v[q]);
        if (lane == 0) scratch[q * 32 + wid] = v[q];
    }
    __syncthreads();
    if (tid == 0) {
#pragma unroll
        for (int q = 0; q < NV; q++) {
            double t = 0.0;
            for (int w = 0; w < nw; w++) t += scratch[q * 32 + w];
            v[q] = t;
        }
    }
}

// ---------------------------------------------------------------------------------------
// Whole (P)CG solve of a small 2-D grid in one cluster launch.  Same recurrences, decisions, status
// codes and iteration counts as k_krylov_init / k_cg_dir / k_cg_update (linear_solver_cg.c:290-461);
// the dot products are summed thread -> warp -> block -> cluster ranks in a fixed order (deterministic;
// another order than the per-kernel path, so equal to the reference within the CG tolerance, not bitwise).
// x must carry its Neumann ghosts on entry (the host runs k_neumann_fill before, and the final
// BC pass -- which honours Ctl::skip_final_bc -- after).  Three cluster barriers per iteration: halo rows of
// the new direction, p.Ap, {r.r, r.z}.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(MINI_THREADS, 1)
k_cg_cluster(Geom g, double* __restrict__ x_g, const double* __restrict__ rhs_g, Ctl* ctl, KArgs a, int rows_per,
             int precond) {
    extern __shared__ __align__(16) double sm[];
    __shared__ double red_scratch[2 * 32];
    cg::cluster_group cluster = cg::this_cluster();
    const int rank = (int)cluster.block_rank(), nc = (int)cluster.num_blocks();
    const int nx = g.nx, ny = g.ny;
    const int jb = 1 + rank * rows_per, je = min(jb + rows_per, ny - 1);
    const int myrows = je - jb;
    double* P = sm;                                   // rows jb-1 .. je (local l = 0 .. myrows+1), ghosts / halos included
    double* X = P + (size_t)(rows_per + 2) * nx;      // own rows only: local row l-1
    double* R = X + (size_t)rows_per * nx;
    double* AP = R + (size_t)rows_per * nx;
    double* slot = AP + (size_t)rows_per * nx;        // [3][MINI_MAX_CTAS]: p.Ap, r.r, r.z
    const int tid = threadIdx.y * blockDim.x + threadIdx.x, nthr = blockDim.x * blockDim.y;

    auto share = [&](int which, double v) {           // thread 0: my partial into every CTA's slot
        for (int r = 0; r < nc; r++) cluster.map_shared_rank(slot, r)[which * MINI_MAX_CTAS + rank] = v;
    };
    auto total = [&](int which) {                     // after the barrier: same sum, same order, in every CTA
        double t = 0.0;
        for (int r = 0; r < nc; r++) t += slot[which * MINI_MAX_CTAS + r];
        return t;
    };
    auto push_halo = [&]() {                          // edge rows of P -> the neighbours' halo rows
        if (rank > 0) {
            double* dst = cluster.map_shared_rank(P, rank - 1) + (size_t)(rows_per + 1) * nx;
            for (int t = tid; t < nx; t += nthr) dst[t] = P[nx + t];
        }
        if (rank < nc - 1) {
            double* dst = cluster.map_shared_rank(P, rank + 1);
            for (int t = tid; t < nx; t += nthr) dst[t] = P[(size_t)myrows * nx + t];
        }
    };

    // ---- r = -rhs + lap(x), z = r or dinv r, p0 = z (boundary cells of p stay 0), r.r, r.z
    for (int t = tid; t < (rows_per + 2) * nx; t += nthr) P[t] = 0.0;
    cluster_barrier(cluster, nc);  // every CTA runs and has cleared its strip before a neighbour pushes a halo row into it
    double acc[2] = {0.0, 0.0};
    for (int l = 1 + (int)threadIdx.y; l <= myrows; l += blockDim.y) {
        const size_t gn0 = (size_t)(jb + l - 1) * nx;
        for (int i = 1 + (int)threadIdx.x; i < nx - 1; i += blockDim.x) {
            const size_t n = gn0 + i;
            const double zc = x_g[n], zm = 0.0, zp = 0.0;
            const double lap = (x_g[n + 1] - 2.0 * zc + x_g[n - 1]) * g.inv_dx2 +
                               (x_g[n + nx] - 2.0 * zc + x_g[n - nx]) * g.inv_dy2 + (zp + zm - 2.0 * zc) * g.inv_dz2;
            const double rv = -rhs_g[n] + lap;
            const double zv = precond ? a.dinv * rv : rv;
            const int m = (l - 1) * nx + i;
            X[m] = zc;
            R[m] = rv;
            P[l * nx + i] = zv;
            acc[0] += rv * rv;
            acc[1] += rv * zv;
        }
    }
    block_sum<2>(acc, red_scratch);
    if (tid == 0) { share(1, acc[0]); share(2, acc[1]); }
    __syncthreads();
    push_halo();
    cluster_barrier(cluster, nc);
    double rr = total(1), rz = total(2);

    double r0 = 0.0, tol = 0.0, res_fin = 0.0, rho_old = 0.0;
    int status = POISSON_ERROR, iters = 0, skip_bc = 0;
    for (int it = 0;; it++) {
        // ---- decisions at the start of iteration `it` (cg_dir_prologue)
        const double rho_new = precond ? rz : rr;
        const double res = sqrt(rr);
        double beta = 0.0;
        if (it == 0) {
            r0 = res;
            tol = fmax(a.rel_tol * res, a.abs_tol);
            if (res < a.abs_tol) { status = POISSON_CONVERGED; iters = 0; res_fin = res; skip_bc = 1; break; }
        } else {
            const bool conv = (res < tol || res < a.abs_tol);
            if (((it - 1) % a.check) == 0 && conv) { status = POISSON_CONVERGED; iters = it; res_fin = res; skip_bc = 0; break; }
            if (fabs(rho_old) < 1e-30) { status = POISSON_STAGNATED; iters = it; res_fin = res; skip_bc = 1; break; }
            if (it >= a.max_iter) {
                status = conv ? POISSON_CONVERGED : POISSON_MAX_ITER; iters = a.max_iter; res_fin = res; skip_bc = 0;
                break;
            }
            beta = rho_new / rho_old;
            // ---- p = z + beta p on the own cells, then the halo rows
            for (int l = 1 + (int)threadIdx.y; l <= myrows; l += blockDim.y)
                for (int i = 1 + (int)threadIdx.x; i < nx - 1; i += blockDim.x) {
                    const double rv = R[(l - 1) * nx + i];
                    P[l * nx + i] = (precond ? a.dinv * rv : rv) + beta * P[l * nx + i];
                }
            __syncthreads();
            push_halo();
            cluster_barrier(cluster, nc);
        }
        // ---- Ap = -lap(p), p.Ap
        double pap[1] = {0.0};
        for (int l = 1 + (int)threadIdx.y; l <= myrows; l += blockDim.y)
            for (int i = 1 + (int)threadIdx.x; i < nx - 1; i += blockDim.x) {
                const int n = l * nx + i;
                const double cc = P[n], cm = 0.0, cp = 0.0;
                const double ce = P[n + 1], cw = P[n - 1], cn = P[n + nx], cs = P[n - nx];
                const double Ac = -((ce - 2.0 * cc + cw) * g.inv_dx2 + (cn - 2.0 * cc + cs) * g.inv_dy2 +
                                    (cp + cm - 2.0 * cc) * g.inv_dz2);
                AP[n - nx] = Ac;
                pap[0] += cc * Ac;
            }
        block_sum<1>(pap, red_scratch);
        if (tid == 0) share(0, pap[0]);
        cluster_barrier(cluster, nc);
        const double pAp = total(0);
        // ---- x += alpha p, r -= alpha Ap, r.r, r.z (k_cg_update)
        if (fabs(pAp) < 1e-30) {
            status = POISSON_STAGNATED; iters = it + 1; res_fin = (it == 0) ? r0 : sqrt(rr); skip_bc = 1;
            break;
        }
        const double alpha = rho_new / pAp;
        acc[0] = acc[1] = 0.0;
        for (int l = 1 + (int)threadIdx.y; l <= myrows; l += blockDim.y)
            for (int i = 1 + (int)threadIdx.x; i < nx - 1; i += blockDim.x) {
                const int m = (l - 1) * nx + i;
                X[m] += alpha * P[l * nx + i];
                const double rv = R[m] - alpha * AP[m];
                R[m] = rv;
                acc[0] += rv * rv;
                acc[1] += rv * (a.dinv * rv);
            }
        block_sum<2>(acc, red_scratch);
        if (tid == 0) { share(1, acc[0]); share(2, acc[1]); }
        cluster_barrier(cluster, nc);
        rho_old = rho_new;
        rr = total(1);
        rz = total(2);
    }

    // ---- interior of x back to global memory; the ghosts follow in the host's final BC pass
    for (int l = 1 + (int)threadIdx.y; l <= myrows; l += blockDim.y)
        for (int i = 1 + (int)threadIdx.x; i < nx - 1; i += blockDim.x)
            x_g[(size_t)(jb + l - 1) * nx + i] = X[(l - 1) * nx + i];
    if (rank == 0 && tid == 0) {
        ctl->r0 = r0;
        ctl->tol = tol;
        ctl->res = res_fin;
        ctl->status = status;
        ctl->iterations = iters;
        ctl->result_buf = 0;
        ctl->skip_final_bc = skip_bc;
        ctl->done_seq = 0;
    }
    cluster_barrier(cluster, nc);
}

template <class K, class... Args>
static bool launch_cluster(const LaunchCtx& L, const MiniPlan& p, int kind, K kernel, Args... args) {
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3(p.nc, 1, 1);
    cfg.blockDim = dim3(p.bx, p.by, 1);
    cfg.dynamicSmemBytes = p.smem;
    cfg.stream = L.st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = p.nc;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    cudaError_t e = cudaSuccess;
    if (p.nc > 8 && cudaFuncSetAttribute(kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    CFDB_LAUNCH(kind, L.st, (e = cudaLaunchKernelEx(&cfg, kernel, args...)));
    if (e != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return true;
}

}  // namespace

bool cg_cluster_available(const LaunchCtx& L) {
    if (getenv("CFD_B200_NO_CLUSTER_CG")) return false;  // read per solve: tests toggle it
    MiniPlan p;
    return mini_plan(L.g, p, true);
}

// Whole (P)CG solve in one launch (between the host's two Neumann passes); false => not launched.
bool launch_cg_cluster(const LaunchCtx& L, const KArgs& a, double* x, const double* rhs, int precond) {
    MiniPlan p;
    if (!mini_plan(L.g, p, true)) return false;
    CFDB_ONCE_PER_DEVICE(cudaFuncSetAttribute(k_cg_cluster, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MINI_MAX_SMEM));
    if (cudaGetLastError() != cudaSuccess) return false;
    return launch_cluster(L, p, KK_CG_CLUSTER, k_cg_cluster, L.g, x, rhs, L.ctl, a, p.rows_per, precond);
}

bool jacobi_cluster_available(const LaunchCtx& L) {
    if (getenv("CFD_B200_NO_CLUSTER_JACOBI")) return false;  // read per solve: tests toggle it
    MiniPlan p;
    return mini_plan(L.g, p);
}

// Whole solve in one launch; x is updated in place.  false => not launched (use the per-sweep kernels).
template <int CPT>
static bool launch_jacobi_cluster_cpt(const LaunchCtx& L, MiniPlan p, const KArgs& a, double* x, const double* rhs) {
    CFDB_ONCE_PER_DEVICE(cudaFuncSetAttribute(k_jacobi_cluster<CPT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MINI_MAX_SMEM));
    if (cudaGetLastError() != cudaSuccess) return false;
    p.bx = MINI_THREADS;  // 1-D block: cells are dealt out by linear index
    p.by = 1;
    if (p.nc > 1 && !getenv("CFD_B200_CLUSTER_SYNC")) {  // neighbour-only signalling (default); _SYNC=1: cluster.sync() per sweep
        CFDB_ONCE_PER_DEVICE(cudaFuncSetAttribute(k_jacobi_cluster_async<CPT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MINI_MAX_SMEM));
        if (cudaGetLastError() != cudaSuccess) return false;
        return launch_cluster(L, p, KK_JACOBI_CLUSTER, k_jacobi_cluster_async<CPT>, L.g, x, rhs, L.ctl, a, p.rows_per);
    }
    return launch_cluster(L, p, KK_JACOBI_CLUSTER, k_jacobi_cluster<CPT>, L.g, x, rhs, L.ctl, a, p.rows_per);
}
bool launch_jacobi_cluster(const LaunchCtx& L, const KArgs& a, double* x, const double* rhs) {
    MiniPlan p;
    if (!mini_plan(L.g, p)) return false;
    if (p.cpt <= 1) return launch_jacobi_cluster_cpt<1>(L, p, a, x, rhs);
    if (p.cpt <= 2) return launch_jacobi_cluster_cpt<2>(L, p, a, x, rhs);
    if (p.cpt <= 4) return launch_jacobi_cluster_cpt<4>(L, p, a, x, rhs);
    if (p.cpt <= 8) return launch_jacobi_cluster_cpt<8>(L, p, a, x, rhs);
    return launch_jacobi_cluster_cpt<16>(L, p, a, x, rhs);
}

}  // namespace cfdb
