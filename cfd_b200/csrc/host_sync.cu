// host_sync.cu -- boundary-face refresh of a device-resident slab.
//
// The reference's validation drivers edit only boundary cells on the host between steps
// (tests/validation/lid_driven_cavity_common.h:306-338: apply BCs, solver_step, repeat), while the reference's
// GPU backend re-uploads every field on every step (solver_projection_gpu.cu:478-521).  With the opt-in policy
// CFD_B200_SYNC_UPLOAD_BOUNDARY a step moves only the outermost layer of cells of each field:
//   z faces  two contiguous planes                        -> plain copies straight from the caller's arrays
//   y faces  one row of nx doubles per plane              -> one strided (2-D) copy per face
//   x faces  one double per row: packed on the host by a few threads into pinned memory, one copy,
//            scattered by a kernel (a DMA with 8-byte rows would crawl)
// At 1024^3 that is 6 x 8.4 MB per field instead of 8.6 GB.  Slabs upload only faces that are physical.
#include "runner.h"

#include <thread>

namespace cfdb {

namespace {

__global__ void k_scatter_xfaces(double* __restrict__ f, const double* __restrict__ packed, int nx, long long rows) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= rows) return;
    f[t * nx] = packed[t];
    f[t * nx + nx - 1] = packed[rows + t];
}

void pack_xfaces(const double* h, double* out, size_t nx, size_t rows) {
    unsigned hc = std::thread::hardware_concurrency();
    size_t nt = rows > (1u << 16) ? (hc > 16 ? 8 : (hc > 1 ? hc / 2 : 1)) : 1;
    if (nt < 1) nt = 1;
    auto work = [&](size_t t) {
        const size_t a = rows * t / nt, b = rows * (t + 1) / nt;
        for (size_t q = a; q < b; q++) {
            out[q] = h[q * nx];
            out[rows + q] = h[q * nx + nx - 1];
        }
    };
    std::vector<std::thread> th;
    for (size_t t = 1; t < nt; t++) th.emplace_back(work, t);
    work(0);
    for (std::thread& x : th) x.join();
}

}  // namespace

cfd_status_t upload_faces(Runner* r, int rank, const flow_field* f, int mask) {
    cfd_b200_session* s = r->s[rank];
    CFDB_CUDA_OK(cudaSetDevice(s->dev));
    const Geom& g = s->g;
    const size_t nx = g.nx, ny = g.ny, nz = g.nz, plane = nx * ny, rows = ny * nz;
    double* dv[6] = {s->u, s->v, s->w, s->p, s->rho, s->T};
    double* hv[6] = {f->u, f->v, f->w, f->p, f->rho, f->T};
    int nf = 0;
    for (int q = 0; q < 6; q++)
        if ((mask & (1 << q)) && dv[q] && hv[q]) nf++;
    if (!nf) return CFD_SUCCESS;
    const size_t need = (size_t)nf * 2 * rows;
    if (r->pack_cap[rank] < need || !r->h_pack[rank]) {
        if (r->h_pack[rank]) cudaFreeHost(r->h_pack[rank]);
        if (r->d_pack[rank]) cudaFree(r->d_pack[rank]);
        r->h_pack[rank] = r->d_pack[rank] = nullptr;
        r->pack_cap[rank] = 0;
        if (cudaHostAlloc((void**)&r->h_pack[rank], need * sizeof(double), cudaHostAllocDefault) != cudaSuccess ||
            cudaMalloc((void**)&r->d_pack[rank], need * sizeof(double)) != cudaSuccess) {
            cudaGetLastError();
            set_error(CFD_ERROR_NOMEM, "staging buffers for the boundary-face upload could not be allocated");
            return CFD_ERROR_NOMEM;
        }
        r->pack_cap[rank] = need;
    }
    double* hp = r->h_pack[rank];
    double* dp = r->d_pack[rank];
    int k = 0;
    for (int q = 0; q < 6; q++) {
        if (!(mask & (1 << q)) || !dv[q] || !hv[q]) continue;
        // x faces: every local layer (halo layers of a slab included: the neighbour uploads the same values)
        pack_xfaces(hv[q], hp + (size_t)k * 2 * rows, nx, rows);
        CFDB_CUDA_OK(cudaMemcpyAsync(dp + (size_t)k * 2 * rows, hp + (size_t)k * 2 * rows, 2 * rows * sizeof(double),
                                     cudaMemcpyHostToDevice, s->st));
        const int nb = (int)((rows + 255) / 256);
        t_prof = &s->prof;
        CFDB_LAUNCH(KK_BOUNDARY, s->st, k_scatter_xfaces<<<nb, 256, 0, s->st>>>(dv[q], dp + (size_t)k * 2 * rows, (int)nx,
                                                                              (long long)rows));
        k++;
        if (g.dim3) {
            // y faces: rows j = 0 and j = ny-1 of every plane
            CFDB_CUDA_OK(cudaMemcpy2DAsync(dv[q], plane * sizeof(double), hv[q], plane * sizeof(double),
                                           nx * sizeof(double), nz, cudaMemcpyHostToDevice, s->st));
            CFDB_CUDA_OK(cudaMemcpy2DAsync(dv[q] + (ny - 1) * nx, plane * sizeof(double), hv[q] + (ny - 1) * nx,
                                           plane * sizeof(double), nx * sizeof(double), nz, cudaMemcpyHostToDevice, s->st));
            // z faces where this slab holds a physical one
            if (g.glo[2])
                CFDB_CUDA_OK(cudaMemcpyAsync(dv[q], hv[q], plane * sizeof(double), cudaMemcpyHostToDevice, s->st));
            if (g.ghi[2])
                CFDB_CUDA_OK(cudaMemcpyAsync(dv[q] + (nz - 1) * plane, hv[q] + (nz - 1) * plane, plane * sizeof(double),
                                             cudaMemcpyHostToDevice, s->st));
        } else {
            if (g.glo[1]) CFDB_CUDA_OK(cudaMemcpyAsync(dv[q], hv[q], nx * sizeof(double), cudaMemcpyHostToDevice, s->st));
            if (g.ghi[1])
                CFDB_CUDA_OK(cudaMemcpyAsync(dv[q] + (ny - 1) * nx, hv[q] + (ny - 1) * nx, nx * sizeof(double),
                                             cudaMemcpyHostToDevice, s->st));
        }
    }
    CFDB_CUDA_OK(cudaStreamSynchronize(s->st));
    return CFD_SUCCESS;
}

}  // namespace cfdb
