// no_alloc.cu -- an entry point that allocates NOTHING: the L-infinity residual max |lap x - rhs| over the interior
// cells (the reference's poisson_solver_compute_residual, linear_solver.c:304-346) on caller-owned DEVICE arrays.
//
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It is a variant of
// k_linf_residual (poisson_kernels.cu): same expression, same operation order (ediv, no FMA contraction), hence the
// same bits as the residual every stationary solve reports -- but where that kernel reduces through a session's
// cudaMalloc'ed RedBuf (per-block partials + ticket), this one reduces through a file-scope device array:
//
//   g_noalloc_scratch   32 doubles of header (the ticket lives in its first word) + one partial maximum per block,
//                       CFD_B200_NOALLOC_MAX_CELLS / 256 = 65 536 blocks at most  =>  524 544 bytes on EVERY GPU that
//                       loads the library, held until the process exits.
//
// In host code the array's NAME is a host-side stand-in, not a device address (a kernel handed `g_noalloc_scratch`
// from the host would fault or, where the GPU can reach host memory, silently use the wrong memory): the device
// address comes from cudaGetSymbolAddress, once per device, with that device current.
//
// Every call on a device shares the array and finds the previous call's contents in it: the ticket is cleared on the
// caller's stream before each launch, the partials need no clearing (every block writes its own before the last
// block reads them).  Two calls that OVERLAP on one GPU (different streams, or two host threads without ordering)
// race on the scratch and may both return garbage -- order them on one stream or with events.
// The entry point calls no cudaMalloc / cudaFree / allocating routine; the memset and the launch are capturable.
#include "session.h"

#include <mutex>

namespace cfdb {

namespace {

constexpr int NA_THREADS = 256;
constexpr size_t NA_MAX_CELLS = CFD_B200_NOALLOC_MAX_CELLS;
constexpr size_t NA_MAX_BLOCKS = NA_MAX_CELLS / NA_THREADS;
constexpr size_t NA_HEADER = 32;  // doubles: 256 bytes, the ticket in the first word

static __device__ __align__(256) double g_noalloc_scratch[NA_HEADER + NA_MAX_BLOCKS];

// One thread per interior cell (linear over the interior, x fastest: coalesced centre / rhs loads, neighbours from
// L1 / L2), block maximum -> scratch, the block that draws the last ticket reduces the partials and writes *out.
template <bool D3>
__global__ void __launch_bounds__(NA_THREADS)
k_linf_residual_noalloc(Geom g, const double* __restrict__ x, const double* __restrict__ rhs, double* scratch,
                        double* __restrict__ out, long long interior) {
    __shared__ double sh[NA_THREADS / 32];
    __shared__ bool last;
    const long long t = (long long)blockIdx.x * NA_THREADS + threadIdx.x;
    double r = 0.0;
    if (t < interior) {
        const int nxi = g.nx - 2, nyi = g.ny - 2;
        const int i = 1 + (int)(t % nxi);
        const long long row = t / nxi;
        const int j = 1 + (int)(row % nyi), k = g.k0 + (int)(row / nyi);
        const long long n = (long long)k * g.sz + (long long)j * g.sy + i;
        const double zc = x[n];
        const double zm = D3 ? x[n - g.sz] : 0.0, zp = D3 ? x[n + g.sz] : 0.0;
        const double lap = ediv(x[n + 1] - 2.0 * zc + x[n - 1], g.dx2, g.inv_dx2) +
                           ediv(x[n + g.sy] - 2.0 * zc + x[n - g.sy], g.dy2, g.inv_dy2) +
                           (zp + zm - 2.0 * zc) * g.inv_dz2;
        max_into(r, fabs(lap - __ldg(rhs + n)));
    }
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    r = warp_max(r);
    if (lane == 0) sh[wid] = r;
    __syncthreads();
    unsigned* ticket = reinterpret_cast<unsigned*>(scratch);
    double* partials = scratch + NA_HEADER;
    if (wid == 0) {
        double b = (lane < NA_THREADS / 32) ? sh[lane] : 0.0;
        b = warp_max(b);
        if (lane == 0) {
            partials[blockIdx.x] = b;
            __threadfence();
            last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
        }
    }
    __syncthreads();
    if (!last) return;
    __threadfence();
    double m = 0.0;
    for (unsigned b = threadIdx.x; b < gridDim.x; b += NA_THREADS) max_into(m, __ldcg(partials + b));
    m = warp_max(m);
    __syncthreads();  // sh reuse
    if (lane == 0) sh[wid] = m;
    __syncthreads();
    if (wid == 0) {
        double b = (lane < NA_THREADS / 32) ? sh[lane] : 0.0;
        b = warp_max(b);
        if (lane == 0) {
            out[0] = b;
            *ticket = 0u;  // (the next call clears it again: it may find the array after an aborted launch)
        }
    }
}

std::mutex g_na_mu;
double* g_na_ptr[64] = {};  // device address of g_noalloc_scratch per device ordinal

// Once per device, with that device current.
cfd_status_t scratch_address(double** out) {
    int dev = -1;
    CFDB_CUDA_OK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) {
        set_error(CFD_ERROR_INVALID, "device ordinal %d outside the supported range", dev);
        return CFD_ERROR_INVALID;
    }
    std::lock_guard<std::mutex> lk(g_na_mu);
    if (!g_na_ptr[dev]) {
        double* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) {
            cudaGetLastError();
            set_error(CFD_ERROR, "cudaGetSymbolAddress(g_noalloc_scratch) on device %d failed: %s", dev,
                      cudaGetErrorString(e));
            return CFD_ERROR;
        }
        g_na_ptr[dev] = p;
    }
    *out = g_na_ptr[dev];
    return CFD_SUCCESS;
}

}  // namespace

}  // namespace cfdb

using namespace cfdb;

extern "C" {

size_t cfd_b200_noalloc_max_cells(void) { return NA_MAX_CELLS; }

cfd_status_t cfd_b200_poisson_residual_linf_noalloc(const double* d_x, const double* d_rhs, size_t nx, size_t ny,
                                                    size_t nz, double dx, double dy, double dz, double* d_out,
                                                    cfd_b200_stream_t stream) {
    if (!d_x || !d_rhs || !d_out) {
        set_error(CFD_ERROR_INVALID, "cfd_b200_poisson_residual_linf_noalloc: NULL device pointer");
        return CFD_ERROR_INVALID;
    }
    if (nx < 3 || ny < 3 || nz == 0 || nz == 2 || !(dx > 0.0) || !(dy > 0.0) || (nz > 1 && !(dz > 0.0))) {
        set_error(CFD_ERROR_INVALID, "cfd_b200_poisson_residual_linf_noalloc: invalid extents or spacing");
        return CFD_ERROR_INVALID;
    }
    // (overflow-safe: every factor is checked against the limit before it is multiplied in)
    if (nx > NA_MAX_CELLS || ny > NA_MAX_CELLS / nx || nz > NA_MAX_CELLS / (nx * ny)) {
        set_error(CFD_ERROR_LIMIT_EXCEEDED,
                  "cfd_b200_poisson_residual_linf_noalloc: %zu x %zu x %zu cells exceed the static scratch "
                  "(maximum %zu cells); use a session (cfd_b200_session_poisson) for larger grids",
                  nx, ny, nz, (size_t)NA_MAX_CELLS);
        return CFD_ERROR_LIMIT_EXCEEDED;
    }
    double* scratch = nullptr;
    cfd_status_t rc = scratch_address(&scratch);
    if (rc != CFD_SUCCESS) return rc;
    Geom g;
    memset(&g, 0, sizeof(g));
    g.nx = (int)nx;
    g.ny = (int)ny;
    g.nz = (int)nz;
    g.dim3 = nz > 1;
    g.sy = (long long)nx;
    g.sz = g.dim3 ? (long long)nx * (long long)ny : 0;
    g.k0 = g.dim3 ? 1 : 0;
    g.k1 = g.dim3 ? g.nz - 1 : 1;
    set_spacing(g, dx, dy, dz);
    const long long interior = (long long)(nx - 2) * (long long)(ny - 2) * (long long)(g.k1 - g.k0);
    const unsigned blocks = (unsigned)((interior + NA_THREADS - 1) / NA_THREADS);
    cudaStream_t st = (cudaStream_t)stream;
    // the ticket must start at zero whatever the previous call (or an aborted one) left behind
    CFDB_CUDA_OK(cudaMemsetAsync(scratch, 0, NA_HEADER * sizeof(double), st));
    if (g.dim3) k_linf_residual_noalloc<true><<<blocks, NA_THREADS, 0, st>>>(g, d_x, d_rhs, scratch, d_out, interior);
    else k_linf_residual_noalloc<false><<<blocks, NA_THREADS, 0, st>>>(g, d_x, d_rhs, scratch, d_out, interior);
    g_launches++;
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        set_error(CFD_ERROR, "cfd_b200_poisson_residual_linf_noalloc: launch failed: %s", cudaGetErrorString(e));
        return CFD_ERROR;
    }
    return CFD_SUCCESS;
}

}  // extern "C"
