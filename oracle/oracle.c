/*
 * oracle.c -- CPU restatement of the shaia/CFD projection-method hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product path (cfd_b200/, libcfd_b200.so) may
 * link, import or call this file; only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs use it, and only as the checker.
 *
 * Parity status: PINNED.  tests/test_oracle.py checks every function here bit-for-bit
 * against the compiled reference (oracle/_ref/libcfd_ref.so, built by oracle/Makefile from
 * /root/reference where it lies) and against the reference's golden L2 norms
 * (tests/solvers/navier_stokes/cpu/test_ns_solver_3d.c:343-367, tolerance 1e-12).
 *
 * Every routine follows the operation ORDER of the reference line it cites so that
 * gcc -O2 -ffp-contract=off reproduces the reference's IEEE-754 results exactly.
 * The one thing this file adds is a choice of Poisson method inside the projection
 * step (the reference hard-codes CG, solver_projection.c:217-218).
 */
#include "../include/cfd_b200.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

#define ORC_API __attribute__((visibility("default")))

static double clampd(double lo, double hi, double v) { return fmax(lo, fmin(hi, v)); }

/* ---- geometry shared by all loops (solver_projection.c:72-84) ---- */
typedef struct {
    size_t nx, ny, nz, plane, total, sz, k0, k1;
    double dx, dy, dz, inv_2dz, inv_dz2;
} geo_t;

static geo_t make_geo(size_t nx, size_t ny, size_t nz, double dx, double dy, double dz) {
    geo_t g;
    g.nx = nx; g.ny = ny; g.nz = nz;
    g.plane = nx * ny;
    g.total = g.plane * nz;
    g.sz = (nz > 1) ? g.plane : 0;
    g.k0 = (nz > 1) ? 1 : 0;
    g.k1 = (nz > 1) ? nz - 1 : 1;
    g.dx = dx; g.dy = dy; g.dz = dz;
    g.inv_2dz = (nz > 1 && dz > 0.0) ? 1.0 / (2.0 * dz) : 0.0;
    g.inv_dz2 = (nz > 1 && dz > 0.0) ? 1.0 / (dz * dz) : 0.0;
    return g;
}

#define FOR_INTERIOR(g, i, j, k, idx)                                   \
    for (size_t k = (g).k0; k < (g).k1; k++)                            \
        for (size_t j = 1; j < (g).ny - 1; j++)                         \
            for (size_t i = 1, idx = k * (g).sz + j * (g).nx + 1; i < (g).nx - 1; i++, idx++)

/* Loops without cross-cell dependencies (predictor, divergence, corrector, energy, RK RHS) use
 * FOR_INTERIOR_PAR: identical to FOR_INTERIOR in the scalar build (liboracle.so, no -fopenmp);
 * in the timed multi-threaded baseline build (_ref/liboracle_ref.so) it is the reference's own
 * OpenMP structure (solver_projection_omp.c:93-160: parallel for over the outer loops). */
#ifdef _OPENMP
#define FOR_INTERIOR_PAR(g, i, j, k, idx)                               \
    _Pragma("omp parallel for collapse(2) schedule(static)")            \
    for (size_t k = (g).k0; k < (g).k1; k++)                            \
        for (size_t j = 1; j < (g).ny - 1; j++)                         \
            for (size_t i = 1, idx = k * (g).sz + j * (g).nx + 1; i < (g).nx - 1; i++, idx++)
#else
#define FOR_INTERIOR_PAR FOR_INTERIOR
#endif

/* ======================================================================
 * Boundary helpers
 * ====================================================================== */

/* Homogeneous Neumann on all faces -- linear_solver.c:348-392 followed by
 * boundary_conditions_core_impl.h:41-69 per plane.  Order: z planes, then per plane
 * x faces, then y faces. */
ORC_API void orc_neumann_scalar(double* x, size_t nx, size_t ny, size_t nz) {
    size_t plane = nx * ny;
    if (nz > 1) {
        memcpy(x, x + plane, plane * sizeof(double));
        memcpy(x + (nz - 1) * plane, x + (nz - 2) * plane, plane * sizeof(double));
    }
    for (size_t k = 0; k < nz; k++) {
        double* f = x + k * plane;
        for (size_t j = 0; j < ny; j++) {
            f[j * nx] = f[j * nx + 1];
            f[j * nx + nx - 1] = f[j * nx + nx - 2];
        }
        for (size_t i = 0; i < nx; i++) {
            f[i] = f[nx + i];
            f[(ny - 1) * nx + i] = f[(ny - 2) * nx + i];
        }
    }
}

/* Periodic wrap of one array: x faces, y faces (per plane), z faces --
 * cpu/solver_explicit_euler.c:231-314 (applied there to u,v,w,p,rho,T together). */
static void periodic_one(double* f, size_t nx, size_t ny, size_t nz) {
    size_t plane = nx * ny;
    for (size_t k = 0; k < nz; k++) {
        double* q = f + k * plane;
        for (size_t j = 0; j < ny; j++) {
            q[j * nx] = q[j * nx + nx - 2];
            q[j * nx + nx - 1] = q[j * nx + 1];
        }
        for (size_t i = 0; i < nx; i++) {
            q[i] = q[(ny - 2) * nx + i];
            q[(ny - 1) * nx + i] = q[nx + i];
        }
    }
    if (nz > 1) {
        for (size_t n = 0; n < plane; n++) {
            f[n] = f[(nz - 2) * plane + n];
            f[(nz - 1) * plane + n] = f[plane + n];
        }
    }
}

ORC_API void orc_apply_periodic(flow_field* f) {
    periodic_one(f->u, f->nx, f->ny, f->nz);
    periodic_one(f->v, f->nx, f->ny, f->nz);
    periodic_one(f->w, f->nx, f->ny, f->nz);
    periodic_one(f->p, f->nx, f->ny, f->nz);
    periodic_one(f->rho, f->nx, f->ny, f->nz);
    periodic_one(f->T, f->nx, f->ny, f->nz);
}

/* Velocity boundary carry-over -- boundary_copy_utils.h:93-148 (w only when nz > 1). */
static void copy_vel_boundary(double* du, double* dv, double* dw, const double* su,
                              const double* sv, const double* sw, size_t nx, size_t ny, size_t nz) {
    size_t plane = nx * ny;
    for (size_t k = 0; k < nz; k++)
        for (size_t j = 0; j < ny; j++)
            for (size_t i = 0; i < nx; i++) {
                int on_xy = (i == 0 || i == nx - 1 || j == 0 || j == ny - 1);
                int on_z = (nz > 1) && (k == 0 || k == nz - 1);
                if (!on_xy && !on_z) continue;
                size_t n = k * plane + j * nx + i;
                du[n] = su[n];
                dv[n] = sv[n];
                if (nz > 1) dw[n] = sw[n];
            }
}

/* Per-face thermal BCs in the order left,right,bottom,top,back,front --
 * energy/cpu/energy_solver.c:204-334. */
ORC_API int orc_thermal_bcs(flow_field* f, const ns_solver_params_t* prm) {
    if (!f || !prm || !f->T) return CFD_ERROR_INVALID;
    if (prm->alpha <= 0.0) return CFD_SUCCESS;
    const ns_thermal_bc_config_t* b = &prm->thermal_bc;
    size_t nx = f->nx, ny = f->ny, nz = f->nz, plane = nx * ny;
    bc_type_t faces[6] = {b->left, b->right, b->bottom, b->top, b->back, b->front};
    for (int q = 0; q < (nz > 1 ? 6 : 4); q++)
        if (faces[q] != BC_TYPE_PERIODIC && faces[q] != BC_TYPE_NEUMANN &&
            faces[q] != BC_TYPE_DIRICHLET)
            return CFD_ERROR_INVALID;
    double* T = f->T;
    for (size_t k = 0; k < nz; k++)
        for (size_t j = 0; j < ny; j++) {
            size_t n = k * plane + j * nx;
            if (b->left == BC_TYPE_DIRICHLET) T[n] = b->dirichlet_values.left;
            else if (b->left == BC_TYPE_NEUMANN) T[n] = T[n + 1];
            else T[n] = T[n + nx - 2];
        }
    for (size_t k = 0; k < nz; k++)
        for (size_t j = 0; j < ny; j++) {
            size_t n = k * plane + j * nx + nx - 1;
            if (b->right == BC_TYPE_DIRICHLET) T[n] = b->dirichlet_values.right;
            else if (b->right == BC_TYPE_NEUMANN) T[n] = T[n - 1];
            else T[n] = T[k * plane + j * nx + 1];
        }
    for (size_t k = 0; k < nz; k++)
        for (size_t i = 0; i < nx; i++) {
            size_t n = k * plane + i;
            if (b->bottom == BC_TYPE_DIRICHLET) T[n] = b->dirichlet_values.bottom;
            else if (b->bottom == BC_TYPE_NEUMANN) T[n] = T[n + nx];
            else T[n] = T[k * plane + (ny - 2) * nx + i];
        }
    for (size_t k = 0; k < nz; k++)
        for (size_t i = 0; i < nx; i++) {
            size_t n = k * plane + (ny - 1) * nx + i;
            if (b->top == BC_TYPE_DIRICHLET) T[n] = b->dirichlet_values.top;
            else if (b->top == BC_TYPE_NEUMANN) T[n] = T[n - nx];
            else T[n] = T[k * plane + nx + i];
        }
    if (nz > 1) {
        for (size_t n = 0; n < plane; n++) {
            if (b->back == BC_TYPE_DIRICHLET) T[n] = b->dirichlet_values.back;
            else if (b->back == BC_TYPE_NEUMANN) T[n] = T[plane + n];
            else T[n] = T[(nz - 2) * plane + n];
        }
        for (size_t n = 0; n < plane; n++) {
            size_t m = (nz - 1) * plane + n;
            if (b->front == BC_TYPE_DIRICHLET) T[m] = b->dirichlet_values.front;
            else if (b->front == BC_TYPE_NEUMANN) T[m] = T[(nz - 2) * plane + n];
            else T[m] = T[plane + n];
        }
    }
    return CFD_SUCCESS;
}

/* ======================================================================
 * Source terms and energy
 * ====================================================================== */

/* cpu/solver_explicit_euler.c:317-333 (default branch) + energy_solver.c:185-196. */
static void sources(const ns_solver_params_t* prm, double x, double y, int iter, double dt,
                    double T_local, double* su, double* sv, double* sw) {
    *su = prm->source_amplitude_u * sin(M_PI * y) * exp(-prm->source_decay_rate * iter * dt);
    *sv = prm->source_amplitude_v * sin(2.0 * M_PI * x) * exp(-prm->source_decay_rate * iter * dt);
    *sw = 0.0;
    if (prm->beta != 0.0) {
        double dT = T_local - prm->T_ref;
        *su += -prm->beta * dT * prm->gravity[0];
        *sv += -prm->beta * dT * prm->gravity[1];
        *sw += -prm->beta * dT * prm->gravity[2];
    }
}

/* Explicit-Euler temperature step -- energy/cpu/energy_solver.c:21-176.
 * ws must hold g.total doubles. */
static int energy_step(flow_field* f, const geo_t* g, const ns_solver_params_t* prm, double dt,
                       double* ws) {
    if (prm->alpha <= 0.0) return CFD_SUCCESS;
    double inv_2dx = 1.0 / (2.0 * g->dx), inv_2dy = 1.0 / (2.0 * g->dy);
    double inv_dx2 = 1.0 / (g->dx * g->dx), inv_dy2 = 1.0 / (g->dy * g->dy);
    const double* T = f->T;
    size_t nx = g->nx, sz = g->sz;
    memcpy(ws, T, g->total * sizeof(double));
    FOR_INTERIOR_PAR(*g, i, j, k, n) {
        double Tc = T[n];
        double dTdx = (T[n + 1] - T[n - 1]) * inv_2dx;
        double dTdy = (T[n + nx] - T[n - nx]) * inv_2dy;
        double dTdz = (T[n + sz] - T[n - sz]) * g->inv_2dz;
        double adv = f->u[n] * dTdx + f->v[n] * dTdy + f->w[n] * dTdz;
        double dxx = (T[n + 1] - 2.0 * Tc + T[n - 1]) * inv_dx2;
        double dyy = (T[n + nx] - 2.0 * Tc + T[n - nx]) * inv_dy2;
        double dzz = (T[n + sz] - 2.0 * Tc + T[n - sz]) * g->inv_dz2;
        double dif = prm->alpha * (dxx + dyy + dzz);
        double Q = 0.0;
        double dT = dt * (-adv + dif + Q);
        ws[n] = Tc + dT;
    }
    for (size_t n = 0; n < g->total; n++)
        if (!isfinite(ws[n])) return CFD_ERROR_DIVERGED;
    memcpy(f->T, ws, g->total * sizeof(double));
    return CFD_SUCCESS;
}

/* ======================================================================
 * Poisson solvers (scalar backend)
 * ====================================================================== */

/* L-infinity residual of (lap x - rhs) -- linear_solver.c:304-346. */
static double linf_residual(const geo_t* g, const double* x, const double* rhs) {
    double dx2 = g->dx * g->dx, dy2 = g->dy * g->dy, m = 0.0;
    size_t nx = g->nx, sz = g->sz;
    FOR_INTERIOR(*g, i, j, k, n) {
        double lap = (x[n + 1] - 2.0 * x[n] + x[n - 1]) / dx2 +
                     (x[n + nx] - 2.0 * x[n] + x[n - nx]) / dy2 +
                     (x[n + sz] + x[n - sz] - 2.0 * x[n]) * g->inv_dz2;
        double r = fabs(lap - rhs[n]);
        if (r > m) m = r;
    }
    return m;
}

/* Optimal SOR omega -- linear_solver_internal.h:184-220. */
ORC_API double orc_optimal_omega(size_t nx, size_t ny, size_t nz, double dx, double dy, double dz) {
    double ix = 1.0 / (dx * dx), iy = 1.0 / (dy * dy);
    double iz = (dz > 0.0) ? 1.0 / (dz * dz) : 0.0;
    double num = cos(M_PI / (double)(nx - 1)) * ix + cos(M_PI / (double)(ny - 1)) * iy;
    double den = ix + iy;
    if (nz > 1 && iz > 0.0) {
        num += cos(M_PI / (double)(nz - 1)) * iz;
        den += iz;
    }
    double rj = num / den;
    return 2.0 / (1.0 + sqrt(1.0 - (rj * rj)));
}

/* One Jacobi iteration -- cpu/linear_solver_jacobi.c:76-129. */
static void jacobi_iterate(const geo_t* g, double* x, double* xt, const double* rhs, double inv_factor) {
    double dx2 = g->dx * g->dx, dy2 = g->dy * g->dy;
    size_t nx = g->nx, sz = g->sz;
    FOR_INTERIOR(*g, i, j, k, n) {
        xt[n] = -(rhs[n] - (x[n + 1] + x[n - 1]) / dx2 - (x[n + nx] + x[n - nx]) / dy2 -
                  (x[n + sz] + x[n - sz]) * g->inv_dz2) * inv_factor;
    }
    memcpy(x, xt, g->total * sizeof(double));
    orc_neumann_scalar(x, g->nx, g->ny, g->nz);
}

/* One red-black SOR iteration -- cpu/linear_solver_redblack.c:79-146. */
static void redblack_iterate(const geo_t* g, double* x, const double* rhs, double inv_factor,
                             double omega) {
    double dx2 = g->dx * g->dx, dy2 = g->dy * g->dy;
    size_t nx = g->nx, sz = g->sz;
    for (int colour = 0; colour < 2; colour++)
        for (size_t k = g->k0; k < g->k1; k++)
            for (size_t j = 1; j < g->ny - 1; j++) {
                size_t i0 = (((j + k) % 2 == 0) == (colour == 0)) ? 1 : 2;
                for (size_t i = i0; i < nx - 1; i += 2) {
                    size_t n = k * sz + j * nx + i;
                    double pn = -(rhs[n] - (x[n + 1] + x[n - 1]) / dx2 -
                                  (x[n + nx] + x[n - nx]) / dy2 -
                                  (x[n + sz] + x[n - sz]) * g->inv_dz2) * inv_factor;
                    x[n] = x[n] + omega * (pn - x[n]);
                }
            }
    orc_neumann_scalar(x, g->nx, g->ny, g->nz);
}

/* Driver for the stationary methods -- linear_solver.c:397-485. */
static int stationary_solve(const geo_t* g, poisson_solver_method_t method,
                            const poisson_solver_params_t* prm, double* x, double* xt,
                            const double* rhs, poisson_solver_stats_t* st) {
    double dx2 = g->dx * g->dx, dy2 = g->dy * g->dy;
    double factor = 2.0 * (1.0 / dx2 + 1.0 / dy2 + g->inv_dz2);
    double inv_factor = 1.0 / factor;
    double omega = (prm->omega <= 0.0)
                       ? orc_optimal_omega(g->nx, g->ny, g->nz, g->dx, g->dy, g->dz)
                       : prm->omega;
    double r0 = linf_residual(g, x, rhs);
    double tol = prm->tolerance * r0;
    if (tol < prm->absolute_tolerance) tol = prm->absolute_tolerance;
    st->initial_residual = r0;
    if (r0 < prm->absolute_tolerance) {
        st->status = POISSON_CONVERGED;
        st->iterations = 0;
        st->final_residual = r0;
        return CFD_SUCCESS;
    }
    int converged = 0, iter;
    double res = r0;
    for (iter = 0; iter < prm->max_iterations; iter++) {
        if (method == POISSON_METHOD_JACOBI) jacobi_iterate(g, x, xt, rhs, inv_factor);
        else redblack_iterate(g, x, rhs, inv_factor, omega);
        if (iter % prm->check_interval == 0) {
            res = linf_residual(g, x, rhs);
            if (res < tol || res < prm->absolute_tolerance) {
                converged = 1;
                break;
            }
        }
    }
    st->iterations = iter + 1;
    st->final_residual = res;
    st->status = converged ? POISSON_CONVERGED : POISSON_MAX_ITER;
    return converged ? CFD_SUCCESS : CFD_ERROR_MAX_ITER;
}

/* Krylov building blocks on interior cells -- cpu/linear_solver_cg.c:67-193. */
static double dot(const geo_t* g, const double* a, const double* b) {
#ifdef ORC_DOT_LONGDOUBLE
    /* Sensitivity probe only (liboracle_ld.so, tests/test_oracle.py::test_bicgstab_is_sensitive...):
     * the same sum accumulated in 80-bit precision, i.e. a legal re-rounding of the reduction.
     * Shows how far the reference's own Krylov iterates move when a dot product changes in the
     * last bit -- which is what any parallel (re-ordered) reduction does. */
    long double sl = 0.0L;
    FOR_INTERIOR(*g, i, j, k, n) sl += (long double)a[n] * (long double)b[n];
    return (double)sl;
#endif
    double s = 0.0;
    FOR_INTERIOR(*g, i, j, k, n) s += a[n] * b[n];
    return s;
}
static void axpy(const geo_t* g, double alpha, const double* x, double* y) {
    FOR_INTERIOR(*g, i, j, k, n) y[n] += alpha * x[n];
}
static double lap_at(const geo_t* g, const double* p, size_t n, double ix, double iy) {
    size_t nx = g->nx, sz = g->sz;
    return ((p[n + 1] - (2.0 * p[n]) + p[n - 1]) * ix) +
           ((p[n + nx] - (2.0 * p[n]) + p[n - nx]) * iy) +
           ((p[n + sz] + p[n - sz] - (2.0 * p[n])) * g->inv_dz2);
}
static void neg_laplacian(const geo_t* g, const double* p, double* Ap) {
    double ix = 1.0 / (g->dx * g->dx), iy = 1.0 / (g->dy * g->dy);
    FOR_INTERIOR(*g, i, j, k, n) Ap[n] = -lap_at(g, p, n, ix, iy);
}
static void krylov_residual(const geo_t* g, const double* x, const double* rhs, double* r) {
    double ix = 1.0 / (g->dx * g->dx), iy = 1.0 / (g->dy * g->dy);
    FOR_INTERIOR(*g, i, j, k, n) r[n] = -rhs[n] + lap_at(g, x, n, ix, iy);
}

/* CG / Jacobi-PCG -- cpu/linear_solver_cg.c:290-461. */
static int cg_solve(const geo_t* g, const poisson_solver_params_t* prm, double* x,
                    const double* rhs, poisson_solver_stats_t* st) {
    int pre = (prm->preconditioner == POISSON_PRECOND_JACOBI);
    double dx2 = g->dx * g->dx, dy2 = g->dy * g->dy;
    double dinv = 1.0 / (2.0 / dx2 + 2.0 / dy2 + 2.0 * g->inv_dz2);
    double* r = calloc(g->total, sizeof(double));
    double* p = calloc(g->total, sizeof(double));
    double* Ap = calloc(g->total, sizeof(double));
    double* z = pre ? calloc(g->total, sizeof(double)) : NULL;
    int rc = CFD_ERROR_MAX_ITER;
    orc_neumann_scalar(x, g->nx, g->ny, g->nz);
    krylov_residual(g, x, rhs, r);
    double rho;
    if (pre) {
        FOR_INTERIOR(*g, i, j, k, n) z[n] = dinv * r[n];
        FOR_INTERIOR(*g, i, j, k, n) p[n] = z[n];
        rho = dot(g, r, z);
    } else {
        FOR_INTERIOR(*g, i, j, k, n) p[n] = r[n];
        rho = dot(g, r, r);
    }
    double r0 = sqrt(dot(g, r, r));
    st->initial_residual = r0;
    double tol = prm->tolerance * r0;
    if (tol < prm->absolute_tolerance) tol = prm->absolute_tolerance;
    if (r0 < prm->absolute_tolerance) {
        st->status = POISSON_CONVERGED;
        st->iterations = 0;
        st->final_residual = r0;
        rc = CFD_SUCCESS;
        goto done; /* reference returns before the final BC (linear_solver_cg.c:353-361) */
    }
    int converged = 0, iter;
    double res = r0;
    for (iter = 0; iter < prm->max_iterations; iter++) {
        neg_laplacian(g, p, Ap);
        double pAp = dot(g, p, Ap);
        if (fabs(pAp) < 1e-30) {
            st->status = POISSON_STAGNATED;
            st->iterations = iter + 1;
            st->final_residual = res;
            goto done;
        }
        double alpha = rho / pAp;
        axpy(g, alpha, p, x);
        axpy(g, -alpha, Ap, r);
        double rho_new;
        if (pre) {
            FOR_INTERIOR(*g, i, j, k, n) z[n] = dinv * r[n];
            rho_new = dot(g, r, z);
        } else {
            rho_new = dot(g, r, r);
        }
        res = sqrt(dot(g, r, r));
        if (iter % prm->check_interval == 0 && (res < tol || res < prm->absolute_tolerance)) {
            converged = 1;
            break;
        }
        if (fabs(rho) < 1e-30) {
            st->status = POISSON_STAGNATED;
            st->iterations = iter + 1;
            st->final_residual = res;
            goto done;
        }
        double beta = rho_new / rho;
        const double* d = pre ? z : r;
        FOR_INTERIOR(*g, i, j, k, n) p[n] = d[n] + beta * p[n];
        rho = rho_new;
    }
    if (!converged && (res < tol || res < prm->absolute_tolerance)) converged = 1;
    orc_neumann_scalar(x, g->nx, g->ny, g->nz);
    st->iterations = (iter < prm->max_iterations) ? iter + 1 : iter;
    st->final_residual = res;
    st->status = converged ? POISSON_CONVERGED : POISSON_MAX_ITER;
    rc = converged ? CFD_SUCCESS : CFD_ERROR_MAX_ITER;
done:
    free(r); free(p); free(Ap); free(z);
    return rc;
}

/* BiCGSTAB -- cpu/linear_solver_bicgstab.c:265-495. */
static int bicgstab_solve(const geo_t* g, const poisson_solver_params_t* prm, double* x,
                          const double* rhs, poisson_solver_stats_t* st) {
    double* r = calloc(g->total, sizeof(double));
    double* rh = calloc(g->total, sizeof(double));
    double* p = calloc(g->total, sizeof(double));
    double* v = calloc(g->total, sizeof(double));
    double* s = calloc(g->total, sizeof(double));
    double* t = calloc(g->total, sizeof(double));
    int rc = CFD_ERROR_MAX_ITER;
    orc_neumann_scalar(x, g->nx, g->ny, g->nz);
    krylov_residual(g, x, rhs, r);
    FOR_INTERIOR(*g, i, j, k, n) rh[n] = r[n];
    double rho = 1.0, alpha = 1.0, omega = 1.0;
    double r0 = sqrt(dot(g, r, r));
    st->initial_residual = r0;
    double tol = prm->tolerance * r0;
    if (tol < prm->absolute_tolerance) tol = prm->absolute_tolerance;
    if (r0 < prm->absolute_tolerance) {
        st->status = POISSON_CONVERGED;
        st->iterations = 0;
        st->final_residual = r0;
        rc = CFD_SUCCESS;
        goto done;
    }
    int converged = 0, iter;
    double res = r0;
    for (iter = 0; iter < prm->max_iterations; iter++) {
        double rho_new = dot(g, rh, r);
        if (fabs(rho_new) < 1e-30) goto stagnated;
        double beta = (rho_new / rho) * (alpha / omega);
        FOR_INTERIOR(*g, i, j, k, n) p[n] = r[n] + beta * (p[n] - omega * v[n]);
        neg_laplacian(g, p, v);
        double rhv = dot(g, rh, v);
        if (fabs(rhv) < 1e-30) goto stagnated;
        alpha = rho_new / rhv;
        FOR_INTERIOR(*g, i, j, k, n) s[n] = r[n] - alpha * v[n];
        double sn = sqrt(dot(g, s, s));
        if (sn < tol || sn < prm->absolute_tolerance) {
            axpy(g, alpha, p, x);
            res = sn;
            converged = 1;
            break;
        }
        neg_laplacian(g, s, t);
        double ts = dot(g, t, s);
        double tt = dot(g, t, t);
        if (fabs(tt) < 1e-30) {
            axpy(g, alpha, p, x);
            res = sn;
            goto stagnated;
        }
        omega = ts / tt;
        FOR_INTERIOR(*g, i, j, k, n) x[n] += alpha * p[n] + omega * s[n];
        FOR_INTERIOR(*g, i, j, k, n) r[n] = s[n] - omega * t[n];
        rho = rho_new;
        res = sqrt(dot(g, r, r));
        if (iter % prm->check_interval == 0 && (res < tol || res < prm->absolute_tolerance)) {
            converged = 1;
            break;
        }
        if (fabs(omega) < 1e-30) goto stagnated;
    }
    if (!converged && (res < tol || res < prm->absolute_tolerance)) converged = 1;
    orc_neumann_scalar(x, g->nx, g->ny, g->nz);
    st->iterations = (iter < prm->max_iterations) ? iter + 1 : iter;
    st->final_residual = res;
    st->status = converged ? POISSON_CONVERGED : POISSON_MAX_ITER;
    rc = converged ? CFD_SUCCESS : CFD_ERROR_MAX_ITER;
    goto done;
stagnated:
    st->status = POISSON_STAGNATED;
    st->iterations = iter + 1;
    st->final_residual = res;
done:
    free(r); free(rh); free(p); free(v); free(s); free(t);
    return rc;
}

/* poisson_solver_solve() equivalent for the scalar backend -- linear_solver.c:487-509.
 * x_temp may be NULL unless method is Jacobi. */
#ifdef ORC_REF_POISSON
/* Timed-baseline build only (_ref/liboracle_ref.so, linked against _ref/libcfd_ref_fast.so):
 * the pressure solve is the REFERENCE's own solver object, created through its public API
 * (poisson_solver.h) with the backend chosen by orc_set_ref_backend(); the stencil loops above
 * stay the restated ones.  Lets bench.py time "projection step with method M" on the CPU
 * although the reference hard-codes CG inside solve_projection_method. */
extern poisson_solver_t* poisson_solver_create(poisson_solver_method_t, poisson_solver_backend_t);
extern cfd_status_t poisson_solver_init(poisson_solver_t*, size_t, size_t, size_t, double, double,
                                        double, const poisson_solver_params_t*);
extern cfd_status_t poisson_solver_solve(poisson_solver_t*, double*, double*, const double*,
                                         poisson_solver_stats_t*);
extern void poisson_solver_destroy(poisson_solver_t*);
static int g_ref_backend = POISSON_BACKEND_SCALAR;
static poisson_solver_t* g_ref_solver = NULL;
static int g_ref_key_method = -1, g_ref_key_backend = -1;
static poisson_solver_params_t g_ref_key_params;
ORC_API void orc_set_ref_backend(int backend) { g_ref_backend = backend; }
ORC_API int orc_ref_backend_available(int method, int backend) {
    poisson_solver_t* s = poisson_solver_create((poisson_solver_method_t)method,
                                                (poisson_solver_backend_t)backend);
    if (!s) return 0;
    poisson_solver_destroy(s);
    return 1;
}
#endif

#include "oracle_mg.inc"

ORC_API int orc_poisson_solve(int method, size_t nx, size_t ny, size_t nz, double dx, double dy,
                              double dz, const poisson_solver_params_t* prm, double* x,
                              double* x_temp, const double* rhs, poisson_solver_stats_t* stats) {
#ifdef ORC_REF_POISSON
    if (method != POISSON_METHOD_MULTIGRID)  /* the reference has no multigrid: the statement below runs */
    {
    if (!g_ref_solver || g_ref_key_method != method || g_ref_key_backend != g_ref_backend ||
        g_ref_solver->nx != nx || g_ref_solver->ny != ny || g_ref_solver->nz != nz ||
        memcmp(&g_ref_key_params, prm, sizeof(*prm)) != 0) {
        if (g_ref_solver) poisson_solver_destroy(g_ref_solver);
        g_ref_solver = poisson_solver_create((poisson_solver_method_t)method,
                                             (poisson_solver_backend_t)g_ref_backend);
        if (!g_ref_solver) return CFD_ERROR_UNSUPPORTED;
        cfd_status_t irc = poisson_solver_init(g_ref_solver, nx, ny, nz, dx, dy, dz, prm);
        if (irc != CFD_SUCCESS) return irc;
        g_ref_key_method = method;
        g_ref_key_backend = g_ref_backend;
        g_ref_key_params = *prm;
    }
    return poisson_solver_solve(g_ref_solver, x, x_temp, rhs, stats);
    }
#endif
    poisson_solver_stats_t local;
    poisson_solver_stats_t* st = stats ? stats : &local;
    st->status = POISSON_ERROR;
    st->iterations = 0;
    st->initial_residual = st->final_residual = st->elapsed_time_ms = 0.0;
    if (nx < 3 || ny < 3 || (nz > 1 && nz < 3)) return CFD_ERROR_INVALID;
    geo_t g = make_geo(nx, ny, nz, dx, dy, dz);
    switch (method) {
        case POISSON_METHOD_JACOBI:
            if (!x_temp) return CFD_ERROR_INVALID;
            return stationary_solve(&g, POISSON_METHOD_JACOBI, prm, x, x_temp, rhs, st);
        case POISSON_METHOD_REDBLACK_SOR:
            return stationary_solve(&g, POISSON_METHOD_REDBLACK_SOR, prm, x, x_temp, rhs, st);
        case POISSON_METHOD_CG: return cg_solve(&g, prm, x, rhs, st);
        case POISSON_METHOD_BICGSTAB: return bicgstab_solve(&g, prm, x, rhs, st);
        case POISSON_METHOD_MULTIGRID: return mg_solve(&g, prm, x, rhs, st);
        default: return CFD_ERROR_UNSUPPORTED;
    }
}

/* ======================================================================
 * Projection step -- cpu/solver_projection.c:46-297
 * ====================================================================== */
ORC_API int orc_projection(flow_field* f, const grid* gr, const ns_solver_params_t* prm,
                           const cfd_b200_poisson_config_t* pc, int nsteps,
                           poisson_solver_stats_t* last_stats) {
    if (!f || !gr || !prm || !pc) return CFD_ERROR_INVALID;
    if (f->nx < 3 || f->ny < 3 || (f->nz > 1 && f->nz < 3)) return CFD_ERROR_INVALID;
    if (prm->source_func || prm->heat_source_func) return CFD_ERROR_UNSUPPORTED;
    double dzv = (f->nz > 1 && gr->dz) ? gr->dz[0] : 0.0;
    geo_t g = make_geo(f->nx, f->ny, f->nz, gr->dx[0], gr->dy[0], dzv);
    size_t nx = g.nx, sz = g.sz, bytes = g.total * sizeof(double);
    double dx = g.dx, dy = g.dy, dt = prm->dt, nu = prm->mu;
    double* us = calloc(g.total, sizeof(double));
    double* vs = calloc(g.total, sizeof(double));
    double* ws = calloc(g.total, sizeof(double));
    double* pn = calloc(g.total, sizeof(double));
    double* pt = calloc(g.total, sizeof(double));
    double* rhs = calloc(g.total, sizeof(double));
    double* Tw = calloc(g.total, sizeof(double));
    int rc = CFD_SUCCESS;
    memcpy(us, f->u, bytes);
    memcpy(vs, f->v, bytes);
    memcpy(ws, f->w, bytes);
    memcpy(pn, f->p, bytes);
    const double *u = f->u, *v = f->v, *w = f->w;
    for (int iter = 0; iter < nsteps; iter++) {
        /* predictor :116-185 */
        FOR_INTERIOR_PAR(g, i, j, k, n) {
            double uc = u[n], vc = v[n], wc = w[n];
            double du_dx = (u[n + 1] - u[n - 1]) / (2.0 * dx);
            double du_dy = (u[n + nx] - u[n - nx]) / (2.0 * dy);
            double du_dz = (u[n + sz] - u[n - sz]) * g.inv_2dz;
            double dv_dx = (v[n + 1] - v[n - 1]) / (2.0 * dx);
            double dv_dy = (v[n + nx] - v[n - nx]) / (2.0 * dy);
            double dv_dz = (v[n + sz] - v[n - sz]) * g.inv_2dz;
            double dw_dx = (w[n + 1] - w[n - 1]) / (2.0 * dx);
            double dw_dy = (w[n + nx] - w[n - nx]) / (2.0 * dy);
            double dw_dz = (w[n + sz] - w[n - sz]) * g.inv_2dz;
            double cu = uc * du_dx + vc * du_dy + wc * du_dz;
            double cv = uc * dv_dx + vc * dv_dy + wc * dv_dz;
            double cw = uc * dw_dx + vc * dw_dy + wc * dw_dz;
            double uxx = (u[n + 1] - 2.0 * uc + u[n - 1]) / (dx * dx);
            double uyy = (u[n + nx] - 2.0 * uc + u[n - nx]) / (dy * dy);
            double uzz = (u[n + sz] - 2.0 * uc + u[n - sz]) * g.inv_dz2;
            double vxx = (v[n + 1] - 2.0 * vc + v[n - 1]) / (dx * dx);
            double vyy = (v[n + nx] - 2.0 * vc + v[n - nx]) / (dy * dy);
            double vzz = (v[n + sz] - 2.0 * vc + v[n - sz]) * g.inv_dz2;
            double wxx = (w[n + 1] - 2.0 * wc + w[n - 1]) / (dx * dx);
            double wyy = (w[n + nx] - 2.0 * wc + w[n - nx]) / (dy * dy);
            double wzz = (w[n + sz] - 2.0 * wc + w[n - sz]) * g.inv_dz2;
            double visc_u = nu * (uxx + uyy + uzz);
            double visc_v = nu * (vxx + vyy + vzz);
            double visc_w = nu * (wxx + wyy + wzz);
            double su, sv, sw;
            sources(prm, gr->x[i], gr->y[j], iter, dt, f->T[n], &su, &sv, &sw);
            us[n] = clampd(-100.0, 100.0, uc + dt * (-cu + visc_u + su));
            vs[n] = clampd(-100.0, 100.0, vc + dt * (-cv + visc_v + sv));
            ws[n] = clampd(-100.0, 100.0, wc + dt * (-cw + visc_w + sw));
        }
        copy_vel_boundary(us, vs, ws, f->u, f->v, f->w, g.nx, g.ny, g.nz); /* :188 */
        /* divergence -> rhs :195-214 */
        double rho = f->rho[0];
        if (rho < 1e-10) rho = 1.0;
        FOR_INTERIOR_PAR(g, i, j, k, n) {
            double a = (us[n + 1] - us[n - 1]) / (2.0 * dx);
            double b = (vs[n + nx] - vs[n - nx]) / (2.0 * dy);
            double c = (ws[n + sz] - ws[n - sz]) * g.inv_2dz;
            double div = a + b + c;
            rhs[n] = (rho / dt) * div;
        }
        /* pressure solve :217-224 */
        poisson_solver_stats_t st;
        int prc = orc_poisson_solve(pc->method, g.nx, g.ny, g.nz, dx, dy, dzv, &pc->params, pn, pt,
                                    rhs, &st);
        if (last_stats) *last_stats = st;
        int ok = (prc == CFD_SUCCESS && st.status == POISSON_CONVERGED);
        if (!ok && pc->on_nonconvergence == CFD_B200_NONCONV_ERROR) {
            rc = CFD_ERROR_MAX_ITER;
            break;
        }
        /* corrector :230-250 */
        double dtr = dt / rho;
        FOR_INTERIOR_PAR(g, i, j, k, n) {
            double px = (pn[n + 1] - pn[n - 1]) / (2.0 * dx);
            double py = (pn[n + nx] - pn[n - nx]) / (2.0 * dy);
            double pz = (pn[n + sz] - pn[n - sz]) * g.inv_2dz;
            f->u[n] = clampd(-100.0, 100.0, us[n] - dtr * px);
            f->v[n] = clampd(-100.0, 100.0, vs[n] - dtr * py);
            f->w[n] = clampd(-100.0, 100.0, ws[n] - dtr * pz);
        }
        memcpy(f->p, pn, bytes); /* :253 */
        rc = energy_step(f, &g, prm, dt, Tw); /* :257 */
        if (rc != CFD_SUCCESS) break;
        rc = orc_thermal_bcs(f, prm); /* :268 */
        if (rc != CFD_SUCCESS) break;
        copy_vel_boundary(f->u, f->v, f->w, us, vs, ws, g.nx, g.ny, g.nz); /* :277 */
        for (size_t n = 0; n < g.total; n++) /* :281-289 */
            if (!isfinite(f->u[n]) || !isfinite(f->v[n]) || !isfinite(f->w[n]) ||
                !isfinite(f->p[n])) {
                rc = CFD_ERROR_DIVERGED;
                break;
            }
        if (rc != CFD_SUCCESS) break;
    }
    free(us); free(vs); free(ws); free(pn); free(pt); free(rhs); free(Tw);
    return rc;
}

/* ======================================================================
 * RK2 / RK4 -- momentum_rhs/ns_momentum_rhs_scalar.h:49-190, cpu/solver_rk{2,4}.c
 * ====================================================================== */
static void rk_rhs(const flow_field* f, const grid* gr, const geo_t* g,
                   const ns_solver_params_t* prm, int iter, double dt, double* ku, double* kv,
                   double* kw, double* kp) {
    const double *u = f->u, *v = f->v, *w = f->w, *p = f->p, *rho = f->rho, *T = f->T;
    size_t nx = g->nx, ny = g->ny, nz = g->nz, sz = g->sz;
    FOR_INTERIOR_PAR(*g, i, j, k, n) {
        if (rho[n] <= 1e-10 || fabs(gr->dx[i]) < 1e-10 || fabs(gr->dy[j]) < 1e-10) {
            ku[n] = kv[n] = kw[n] = kp[n] = 0.0;
            continue;
        }
        size_t row = k * sz + j * nx, col = k * sz + i;
        size_t il = (i > 1) ? n - 1 : row + nx - 2;
        size_t ir = (i < nx - 2) ? n + 1 : row + 1;
        size_t jd = (j > 1) ? n - nx : col + (ny - 2) * nx;
        size_t ju = (j < ny - 2) ? n + nx : col + nx;
        size_t kd = (k > 1) ? n - sz : (nz - 2) * sz + j * nx + i;
        size_t kk = (k < nz - 2) ? n + sz : sz + j * nx + i;
        double hx = gr->dx[i], hy = gr->dy[j];
        double du_dx = (u[ir] - u[il]) / (2.0 * hx), du_dy = (u[ju] - u[jd]) / (2.0 * hy);
        double du_dz = (u[kk] - u[kd]) * g->inv_2dz;
        double dv_dx = (v[ir] - v[il]) / (2.0 * hx), dv_dy = (v[ju] - v[jd]) / (2.0 * hy);
        double dv_dz = (v[kk] - v[kd]) * g->inv_2dz;
        double dw_dx = (w[ir] - w[il]) / (2.0 * hx), dw_dy = (w[ju] - w[jd]) / (2.0 * hy);
        double dw_dz = (w[kk] - w[kd]) * g->inv_2dz;
        double dp_dx = (p[ir] - p[il]) / (2.0 * hx), dp_dy = (p[ju] - p[jd]) / (2.0 * hy);
        double dp_dz = (p[kk] - p[kd]) * g->inv_2dz;
        double uxx = (u[ir] - 2.0 * u[n] + u[il]) / (hx * hx);
        double uyy = (u[ju] - 2.0 * u[n] + u[jd]) / (hy * hy);
        double uzz = (u[kk] - 2.0 * u[n] + u[kd]) * g->inv_dz2;
        double vxx = (v[ir] - 2.0 * v[n] + v[il]) / (hx * hx);
        double vyy = (v[ju] - 2.0 * v[n] + v[jd]) / (hy * hy);
        double vzz = (v[kk] - 2.0 * v[n] + v[kd]) * g->inv_dz2;
        double wxx = (w[ir] - 2.0 * w[n] + w[il]) / (hx * hx);
        double wyy = (w[ju] - 2.0 * w[n] + w[jd]) / (hy * hy);
        double wzz = (w[kk] - 2.0 * w[n] + w[kd]) * g->inv_dz2;
        double nu = fmin(prm->mu / fmax(rho[n], 1e-10), 1.0);
#define C1(a) a = clampd(-100.0, 100.0, a)
#define C2(a) a = clampd(-1000.0, 1000.0, a)
        C1(du_dx); C1(du_dy); C1(du_dz); C1(dv_dx); C1(dv_dy); C1(dv_dz);
        C1(dw_dx); C1(dw_dy); C1(dw_dz); C1(dp_dx); C1(dp_dy); C1(dp_dz);
        C2(uxx); C2(uyy); C2(uzz); C2(vxx); C2(vyy); C2(vzz); C2(wxx); C2(wyy); C2(wzz);
#undef C1
#undef C2
        double su, sv, sw;
        sources(prm, gr->x[i], gr->y[j], iter, dt, T ? T[n] : prm->T_ref, &su, &sv, &sw);
        ku[n] = -u[n] * du_dx - v[n] * du_dy - w[n] * du_dz - dp_dx / rho[n] +
                nu * (uxx + uyy + uzz) + su;
        kv[n] = -u[n] * dv_dx - v[n] * dv_dy - w[n] * dv_dz - dp_dy / rho[n] +
                nu * (vxx + vyy + vzz) + sv;
        kw[n] = -u[n] * dw_dx - v[n] * dw_dy - w[n] * dw_dz - dp_dz / rho[n] +
                nu * (wxx + wyy + wzz) + sw;
        double div = clampd(-10.0, 10.0, du_dx + dv_dy + dw_dz);
        kp[n] = -0.1 * rho[n] * div;
    }
}

/* order 2 = Heun (solver_rk2.c:47-228), order 4 = classical RK4 (solver_rk4.c:69-259). */
ORC_API int orc_rk(flow_field* f, const grid* gr, const ns_solver_params_t* prm, int order,
                   int nsteps) {
    if (!f || !gr || !prm) return CFD_ERROR_INVALID;
    if (order != 2 && order != 4) return CFD_ERROR_UNSUPPORTED;
    if (f->nx < 3 || f->ny < 3 || (f->nz > 1 && f->nz < 3)) return CFD_ERROR_INVALID;
    if (prm->source_func || prm->heat_source_func) return CFD_ERROR_UNSUPPORTED;
    double dzv = (f->nz > 1 && gr->dz) ? gr->dz[0] : 0.0;
    geo_t g = make_geo(f->nx, f->ny, f->nz, gr->dx[0], gr->dy[0], dzv);
    size_t N = g.total, bytes = N * sizeof(double);
    double* K = calloc(N * 4 * (size_t)order, sizeof(double)); /* k[stage][comp] */
    double* Q0 = calloc(N * 4, sizeof(double));
    double* Tw = calloc(N, sizeof(double));
    double* q[4] = {f->u, f->v, f->w, f->p};
    double dt = prm->dt;
    int rc = CFD_SUCCESS;
#define KS(s, c) (K + ((size_t)(s) * 4 + (c)) * N)
    for (int iter = 0; iter < nsteps; iter++) {
        for (int c = 0; c < 4; c++) memcpy(Q0 + c * N, q[c], bytes);
        for (int s = 0; s < order; s++) {
            if (s > 0) { /* stage state = Q0 + f*k_{s-1}, velocities clamped, all cells */
                double fac = (order == 2) ? dt : ((s == 3) ? dt : 0.5 * dt);
                for (int c = 0; c < 4; c++)
                    for (size_t n = 0; n < N; n++) {
                        double val = Q0[c * N + n] + fac * KS(s - 1, c)[n];
                        q[c][n] = (c < 3) ? clampd(-100.0, 100.0, val) : val;
                    }
            }
            for (int c = 0; c < 4; c++) memset(KS(s, c), 0, bytes);
            rk_rhs(f, gr, &g, prm, iter, dt, KS(s, 0), KS(s, 1), KS(s, 2), KS(s, 3));
        }
        for (int c = 0; c < 4; c++)
            for (size_t n = 0; n < N; n++) {
                double val;
                if (order == 2)
                    val = Q0[c * N + n] + (0.5 * dt) * (KS(0, c)[n] + KS(1, c)[n]);
                else
                    val = Q0[c * N + n] + (dt / 6.0) * (KS(0, c)[n] + 2.0 * KS(1, c)[n] +
                                                        2.0 * KS(2, c)[n] + KS(3, c)[n]);
                q[c][n] = (c < 3) ? clampd(-100.0, 100.0, val) : val;
            }
        rc = energy_step(f, &g, prm, dt, Tw);
        if (rc != CFD_SUCCESS) break;
        orc_apply_periodic(f);
        rc = orc_thermal_bcs(f, prm);
        if (rc != CFD_SUCCESS) break;
        for (size_t n = 0; n < N; n++)
            if (!isfinite(f->u[n]) || !isfinite(f->v[n]) || !isfinite(f->w[n]) ||
                !isfinite(f->p[n])) {
                rc = CFD_ERROR_DIVERGED;
                break;
            }
        if (rc != CFD_SUCCESS) break;
    }
#undef KS
    free(K); free(Q0); free(Tw);
    return rc;
}

/* ======================================================================
 * Explicit (forward) Euler -- cpu/solver_explicit_euler.c:337-582.  Unlike the RK right-hand
 * side it reads the PLAIN neighbours (boundary cells included, no index wrap), limits every
 * increment to +-1 (UPDATE_LIMIT) and never steps further than 1e-4 (DT_CONSERVATIVE_LIMIT).
 * Boundary handling per step (:527-556): the new interior goes back into the field, the energy
 * equation advances T, every field is wrapped periodically, then the caller's boundary
 * velocities (u, v, and w when nz > 1) are put back, the thermal BCs applied and the state
 * scanned for NaN/Inf.
 * ====================================================================== */
ORC_API int orc_explicit_euler(flow_field* f, const grid* gr, const ns_solver_params_t* prm,
                               int nsteps) {
    if (!f || !gr || !prm) return CFD_ERROR_INVALID;
    if (f->nx < 3 || f->ny < 3 || (f->nz > 1 && f->nz < 3)) return CFD_ERROR_INVALID;
    if (prm->source_func || prm->heat_source_func) return CFD_ERROR_UNSUPPORTED;
    if (f->nz > 1 && gr->dz)
        for (size_t k = 1; k < f->nz - 1; k++)
            if (fabs(gr->dz[k] - gr->dz[0]) > 1e-14) return CFD_ERROR_INVALID;
    double dzv = (f->nz > 1 && gr->dz) ? gr->dz[0] : 0.0;
    geo_t g = make_geo(f->nx, f->ny, f->nz, gr->dx[0], gr->dy[0], dzv);
    const size_t N = g.total, bytes = N * sizeof(double), nx = g.nx, sz = g.sz;
    double* nw[5]; /* next u, v, w, p, rho: start as copies, only interior cells are rewritten */
    double* cur[5] = {f->u, f->v, f->w, f->p, f->rho};
    double* Tw = calloc(N, sizeof(double));
    for (int c = 0; c < 5; c++) {
        nw[c] = malloc(bytes);
        memcpy(nw[c], cur[c], bytes);
    }
    const double dt = fmin(prm->dt, 1e-4);
    int rc = CFD_SUCCESS;
    for (int iter = 0; iter < nsteps && rc == CFD_SUCCESS; iter++) {
        const double *u = f->u, *v = f->v, *w = f->w, *p = f->p, *rho = f->rho;
        FOR_INTERIOR_PAR(g, i, j, k, n) {
            if (rho[n] <= 1e-10 || fabs(gr->dx[i]) < 1e-10 || fabs(gr->dy[j]) < 1e-10) continue;
            const double hx = gr->dx[i], hy = gr->dy[j];
            double d1[4][3], d2[3][3]; /* d1[field][axis], field 3 = p; d2 for u, v, w */
            const double* q[4] = {u, v, w, p};
            for (int c = 0; c < 4; c++) {
                const double* a = q[c];
                d1[c][0] = clampd(-100.0, 100.0, (a[n + 1] - a[n - 1]) / (2.0 * hx));
                d1[c][1] = clampd(-100.0, 100.0, (a[n + nx] - a[n - nx]) / (2.0 * hy));
                d1[c][2] = clampd(-100.0, 100.0, (a[n + sz] - a[n - sz]) * g.inv_2dz);
                if (c == 3) break;
                d2[c][0] = clampd(-1000.0, 1000.0, (a[n + 1] - 2.0 * a[n] + a[n - 1]) / (hx * hx));
                d2[c][1] = clampd(-1000.0, 1000.0, (a[n + nx] - 2.0 * a[n] + a[n - nx]) / (hy * hy));
                d2[c][2] = clampd(-1000.0, 1000.0, (a[n + sz] - 2.0 * a[n] + a[n - sz]) * g.inv_dz2);
            }
            double nu = fmin(prm->mu / fmax(rho[n], 1e-10), 1.0);
            double src[3];
            sources(prm, gr->x[i], gr->y[j], iter, dt, f->T[n], &src[0], &src[1], &src[2]);
            for (int c = 0; c < 3; c++) {
                double inc = dt * (-u[n] * d1[c][0] - v[n] * d1[c][1] - w[n] * d1[c][2] -
                                   d1[3][c] / rho[n] + nu * (d2[c][0] + d2[c][1] + d2[c][2]) + src[c]);
                inc = clampd(-1.0, 1.0, inc);
                nw[c][n] = clampd(-100.0, 100.0, q[c][n] + inc);
            }
            double div = clampd(-10.0, 10.0, d1[0][0] + d1[1][1] + d1[2][2]);
            double dp = clampd(-1.0, 1.0, -0.1 * dt * rho[n] * div);
            nw[3][n] = p[n] + dp;
            nw[4][n] = rho[n];
        }
        for (int c = 0; c < 5; c++) memcpy(cur[c], nw[c], bytes);
        rc = energy_step(f, &g, prm, dt, Tw);
        if (rc != CFD_SUCCESS) break;
        orc_apply_periodic(f);
        copy_vel_boundary(f->u, f->v, f->w, nw[0], nw[1], nw[2], g.nx, g.ny, g.nz);
        rc = orc_thermal_bcs(f, prm);
        if (rc != CFD_SUCCESS) break;
        for (size_t n = 0; n < N; n++)
            if (!isfinite(f->u[n]) || !isfinite(f->v[n]) || !isfinite(f->w[n]) ||
                !isfinite(f->p[n])) {
                rc = CFD_ERROR_DIVERGED;
                break;
            }
    }
    for (int c = 0; c < 5; c++) free(nw[c]);
    free(Tw);
    return rc;
}

/* Stats as the registry wrappers compute them -- solver_registry.c:31-62. */
ORC_API void orc_field_stats(const flow_field* f, double* max_vel, double* max_p, double* max_T) {
    size_t N = f->nx * f->ny * f->nz;
    double mv = 0.0, mp = 0.0, mt = f->T[0];
    for (size_t n = 0; n < N; n++) {
        double vel = sqrt((f->u[n] * f->u[n]) + (f->v[n] * f->v[n]) + (f->w[n] * f->w[n]));
        if (vel > mv) mv = vel;
        double ap = fabs(f->p[n]);
        if (ap > mp) mp = ap;
        if (f->T[n] > mt) mt = f->T[n];
    }
    *max_vel = mv; *max_p = mp; *max_T = mt;
}

/* ======================================================================
 * Check of the device division sequence ediv() (cfd_b200/csrc/common.cuh): a/b from y = RN(1/b)
 * with q = a*y, r = fma(-b,q,a), q' = fma(r,y,q) must equal the IEEE quotient the reference
 * computes with '/'.  Returns the number of mismatches over `count` pseudo-random numerators
 * (xorshift64, exponents 2^-40 .. 2^10, both signs) for the divisor b.
 * ====================================================================== */
ORC_API long orc_ediv_mismatches(double b, unsigned long long seed, long count) {
    unsigned long long s = seed ? seed : 88172645463325252ULL;
    double y = 1.0 / b;
    long bad = 0;
    for (long t = 0; t < count; t++) {
        s ^= s << 13; s ^= s >> 7; s ^= s << 17;
        unsigned long long m = s;
        s ^= s << 13; s ^= s >> 7; s ^= s << 17;
        unsigned long long bits = (m & 0x800FFFFFFFFFFFFFULL) | ((unsigned long long)(1023 - 40 + (s % 51)) << 52);
        double a;
        memcpy(&a, &bits, 8);
        double q = a * y;
        double r = fma(-b, q, a);
        double q1 = fma(r, y, q);
        if (q1 != a / b) bad++;
    }
    return bad;
}
