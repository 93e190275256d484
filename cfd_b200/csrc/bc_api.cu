// bc_api.cu -- the reference's GPU boundary-condition API on DEVICE pointers:
// cfd/boundary/boundary_conditions_gpu.cuh:32-151.  Same names, arguments and error behaviour.
//
// Semantics are those of the reference's CPU implementation (boundary_conditions_core_impl.h:
// x faces, then y faces, then z faces, each over the FULL extent of the other axes), evaluated in
// closed form with one thread per boundary cell and sources in the interior only:
//   Neumann   ghost = nearest interior cell (clamped indices)
//   periodic  ghost = wrapped interior cell
//   Dirichlet the face applied LAST wins on edges and corners (z over y over x)
// so the result is race-free and identical to the sequential CPU passes.  (The reference's own 3-D
// GPU kernels partition edges between face groups and read ghost cells another group is writing,
// boundary_conditions_gpu.cu:279-330; they match the CPU only away from edges.)
// Types the reference has not implemented (NOSLIP / INLET / OUTLET / SYMMETRY passed to the generic
// entry points) fall back to Neumann exactly as in the reference (:199-207).
//
// Inlet profiles are evaluated on the HOST per edge cell with the reference's formulas
// (boundary_conditions_inlet_common.h:133-215; cos/sin from the same libm as the CPU backends) and
// scattered by one small kernel, so custom profile callbacks work too (the reference GPU path
// returns CFD_ERROR_UNSUPPORTED for them, boundary_conditions_inlet_gpu.cu:200-203).
#include "kernels.h"
#include "common.cuh"

#include <math.h>
#include <vector>

namespace cfdb {
namespace {

struct BcFields {
    double* f[3];
    int n;
};
struct BcValues {  // per field: left, right, bottom, top, back (k = 0), front (k = nz-1)
    double v[3][6];
};
enum BcMode { BCM_NEUMANN = 0, BCM_PERIODIC = 1, BCM_DIRICHLET = 2 };

__global__ void k_bc_faces(Geom g, BcFields F, int mode, BcValues V) {
    const long long n = face_cell(g, (long long)blockIdx.x * blockDim.x + threadIdx.x);
    if (n < 0) return;
    const long long nxy = (long long)g.nx * g.ny;
    const int i = (int)(n % g.nx), j = (int)((n / g.nx) % g.ny), k = (int)(n / nxy);
    if (mode == BCM_DIRICHLET) {
        int face;  // the pass that touches the cell last: z, else y, else x
        if (g.dim3 && (k == 0 || k == g.nz - 1)) face = (k == 0) ? 4 : 5;
        else if (j == 0 || j == g.ny - 1) face = (j == 0) ? 2 : 3;
        else face = (i == 0) ? 0 : 1;
        for (int q = 0; q < F.n; q++) F.f[q][n] = V.v[q][face];
        return;
    }
    int ci = i, cj = j, ck = k;
    if (mode == BCM_NEUMANN) {
        ci = min(max(i, 1), g.nx - 2);
        cj = min(max(j, 1), g.ny - 2);
        if (g.dim3) ck = min(max(k, 1), g.nz - 2);
    } else {
        if (i == 0) ci = g.nx - 2;
        if (i == g.nx - 1) ci = 1;
        if (j == 0) cj = g.ny - 2;
        if (j == g.ny - 1) cj = 1;
        if (g.dim3) {
            if (k == 0) ck = g.nz - 2;
            if (k == g.nz - 1) ck = 1;
        }
    }
    const long long src = (long long)ck * nxy + (long long)cj * g.nx + ci;
    for (int q = 0; q < F.n; q++) F.f[q][n] = F.f[q][src];
}

// values along one x/y edge of a 2-D field: idx(t) = t * stride + offset
__global__ void k_bc_edge_scatter(double* u, double* v, const double* uv, int count, long long stride, long long off) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= count) return;
    u[off + t * stride] = uv[t];
    v[off + t * stride] = uv[count + t];
}

Geom bc_geom(size_t nx, size_t ny, size_t nz) {
    Geom g;
    memset(&g, 0, sizeof(g));
    g.nx = (int)nx;
    g.ny = (int)ny;
    g.nz = (int)nz;
    g.dim3 = nz > 1;
    for (int a = 0; a < 3; a++) g.glo[a] = g.ghi[a] = 1;
    return g;
}

void launch_faces(double* a, double* b, double* c, size_t nx, size_t ny, size_t nz, int mode, const BcValues* vals,
                  cudaStream_t st) {
    const Geom g = bc_geom(nx, ny, nz);
    BcFields F;
    F.n = 0;
    double* in[3] = {a, b, c};
    for (int q = 0; q < 3; q++) {
        F.f[q] = nullptr;
        if (in[q]) F.f[F.n++] = in[q];
    }
    BcValues V;
    if (vals) V = *vals; else memset(&V, 0, sizeof(V));
    const int nb = (int)((face_cells(g) + 255) / 256);
    k_bc_faces<<<nb, 256, 0, st>>>(g, F, mode, V);
    g_launches++;
}

int mode_of(bc_type_t type) {  // boundary_conditions_gpu.cu:190-207: everything but PERIODIC is zero-gradient
    return type == BC_TYPE_PERIODIC ? BCM_PERIODIC : BCM_NEUMANN;
}

}  // namespace
}  // namespace cfdb

using namespace cfdb;

extern "C" {

void bc_apply_neumann_gpu(double* d_field, size_t nx, size_t ny, cudaStream_t stream) {
    if (!d_field || nx < 3 || ny < 3) return;
    launch_faces(d_field, nullptr, nullptr, nx, ny, 1, BCM_NEUMANN, nullptr, stream);
}

void bc_apply_scalar_gpu(double* d_field, size_t nx, size_t ny, bc_type_t type, cudaStream_t stream) {
    if (!d_field || nx < 3 || ny < 3) return;
    launch_faces(d_field, nullptr, nullptr, nx, ny, 1, mode_of(type), nullptr, stream);
}

void bc_apply_velocity_gpu(double* d_u, double* d_v, size_t nx, size_t ny, bc_type_t type, cudaStream_t stream) {
    if (!d_u || !d_v || nx < 3 || ny < 3) return;
    launch_faces(d_u, d_v, nullptr, nx, ny, 1, mode_of(type), nullptr, stream);
}

void bc_apply_dirichlet_scalar_gpu(double* d_field, size_t nx, size_t ny, const bc_dirichlet_values_t* values,
                                   cudaStream_t stream) {
    if (!d_field || !values || nx < 3 || ny < 3) return;
    BcValues V;
    memset(&V, 0, sizeof(V));
    const double v[6] = {values->left, values->right, values->bottom, values->top, values->back, values->front};
    for (int f = 0; f < 6; f++) V.v[0][f] = v[f];
    launch_faces(d_field, nullptr, nullptr, nx, ny, 1, BCM_DIRICHLET, &V, stream);
}

void bc_apply_dirichlet_velocity_gpu(double* d_u, double* d_v, size_t nx, size_t ny,
                                     const bc_dirichlet_values_t* u_values, const bc_dirichlet_values_t* v_values,
                                     cudaStream_t stream) {
    if (!d_u || !d_v || !u_values || !v_values || nx < 3 || ny < 3) return;
    BcValues V;
    memset(&V, 0, sizeof(V));
    const bc_dirichlet_values_t* in[2] = {u_values, v_values};
    for (int q = 0; q < 2; q++) {
        const double v[6] = {in[q]->left, in[q]->right, in[q]->bottom, in[q]->top, in[q]->back, in[q]->front};
        for (int f = 0; f < 6; f++) V.v[q][f] = v[f];
    }
    launch_faces(d_u, d_v, nullptr, nx, ny, 1, BCM_DIRICHLET, &V, stream);
}

void bc_apply_scalar_3d_gpu(double* d_field, size_t nx, size_t ny, size_t nz, bc_type_t type, cudaStream_t stream) {
    if (!d_field || nx < 3 || ny < 3) return;
    if (nz == 1) {
        bc_apply_scalar_gpu(d_field, nx, ny, type, stream);
        return;
    }
    if (nz < 3) return;
    launch_faces(d_field, nullptr, nullptr, nx, ny, nz, mode_of(type), nullptr, stream);
}

void bc_apply_velocity_3d_gpu(double* d_u, double* d_v, double* d_w, size_t nx, size_t ny, size_t nz, bc_type_t type,
                              cudaStream_t stream) {
    if (!d_u || !d_v || nx < 3 || ny < 3) return;
    if (nz == 1) {
        bc_apply_velocity_gpu(d_u, d_v, nx, ny, type, stream);  // w is ignored, as in the reference (:137)
        return;
    }
    if (nz < 3) return;
    launch_faces(d_u, d_v, d_w, nx, ny, nz, mode_of(type), nullptr, stream);
}

cfd_status_t bc_apply_inlet_gpu(double* d_u, double* d_v, size_t nx, size_t ny, const bc_inlet_config_t* config,
                                cudaStream_t stream) {
    if (!d_u || !d_v || !config || nx < 3 || ny < 3) return CFD_ERROR_INVALID;
    if (config->edge != BC_EDGE_LEFT && config->edge != BC_EDGE_RIGHT && config->edge != BC_EDGE_BOTTOM &&
        config->edge != BC_EDGE_TOP)
        return CFD_ERROR_INVALID;
    // base velocity: boundary_conditions_inlet_common.h:133-176 (+ the GPU wrapper's stricter mass-flow check)
    double ub = 0.0, vb = 0.0;
    switch (config->spec_type) {
        case BC_INLET_SPEC_VELOCITY:
            ub = config->spec.velocity.u;
            vb = config->spec.velocity.v;
            break;
        case BC_INLET_SPEC_MAGNITUDE_DIR:
            ub = config->spec.magnitude_dir.magnitude * cos(config->spec.magnitude_dir.direction);
            vb = config->spec.magnitude_dir.magnitude * sin(config->spec.magnitude_dir.direction);
            break;
        case BC_INLET_SPEC_MASS_FLOW: {
            const double rho_L = config->spec.mass_flow.density * config->spec.mass_flow.inlet_length;
            if (rho_L <= 0.0) return CFD_ERROR_INVALID;  // boundary_conditions_inlet_gpu.cu:229-232
            const double avg = config->spec.mass_flow.mass_flow_rate / rho_L;
            const int us = config->edge == BC_EDGE_LEFT ? 1 : (config->edge == BC_EDGE_RIGHT ? -1 : 0);
            const int vs = config->edge == BC_EDGE_BOTTOM ? 1 : (config->edge == BC_EDGE_TOP ? -1 : 0);
            ub = avg * us;
            vb = avg * vs;
            break;
        }
        default:
            return CFD_ERROR_INVALID;
    }
    const bool along_y = (config->edge == BC_EDGE_LEFT || config->edge == BC_EDGE_RIGHT);
    const size_t count = along_y ? ny : nx;
    const double denom = (count > 1) ? (double)(count - 1) : 1.0;
    std::vector<double> uv(2 * count);
    for (size_t t = 0; t < count; t++) {  // boundary_conditions_inlet_scalar.c:62-68, inlet_common.h:179-215
        const double pos = (count > 1) ? (double)t / denom : 0.5;
        double uo = ub, vo = vb;
        if (config->profile == BC_INLET_PROFILE_PARABOLIC) {
            const double pf = 4.0 * pos * (1.0 - pos);
            uo = ub * pf;
            vo = vb * pf;
        } else if (config->profile == BC_INLET_PROFILE_CUSTOM && config->custom_profile) {
            config->custom_profile(pos, &uo, &vo, config->custom_profile_user_data);
        }
        uv[t] = uo;
        uv[count + t] = vo;
    }
    double* d_uv = nullptr;
    if (cudaMallocAsync((void**)&d_uv, 2 * count * sizeof(double), stream) != cudaSuccess) {
        cudaGetLastError();
        return CFD_ERROR_NOMEM;
    }
    // pageable source: the copy is staged before the call returns, so the vector may go out of scope
    if (cudaMemcpyAsync(d_uv, uv.data(), 2 * count * sizeof(double), cudaMemcpyHostToDevice, stream) != cudaSuccess) {
        cudaFreeAsync(d_uv, stream);
        cudaGetLastError();
        return CFD_ERROR;
    }
    long long stride, off;
    switch (config->edge) {
        case BC_EDGE_LEFT: stride = (long long)nx; off = 0; break;
        case BC_EDGE_RIGHT: stride = (long long)nx; off = (long long)nx - 1; break;
        case BC_EDGE_BOTTOM: stride = 1; off = 0; break;
        default: stride = 1; off = (long long)(ny - 1) * (long long)nx; break;
    }
    k_bc_edge_scatter<<<(int)((count + 255) / 256), 256, 0, stream>>>(d_u, d_v, d_uv, (int)count, stride, off);
    g_launches++;
    cudaFreeAsync(d_uv, stream);
    return cudaGetLastError() == cudaSuccess ? CFD_SUCCESS : CFD_ERROR;
}

}  // extern "C"
