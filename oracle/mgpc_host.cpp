// oracle/mgpc_host.cpp -- TEST INFRASTRUCTURE ONLY: the CPU restatement of the multigrid-preconditioned CG of
// cfd_b200/csrc/poisson_mgpc.cu.  It runs the SAME per-cell functions (cfd_b200/csrc/mgpc_cell.h: operator rows,
// restriction / prolongation weights, level plan) in plain host loops, with the PCG of the reference around them
// (linear_solver_cg.c:290-461: Neumann ghosts of x at start and end, r = -rhs + lap x, L2 stopping rule
// max(rel * r0, abs), breakdown tests, iterations = it + 1).  Nothing under cfd_b200/ links or calls this file.
//
//   g++ -O2 -ffp-contract=off -fPIC -shared -o libmgpc_host.so mgpc_host.cpp
#include "../cfd_b200/csrc/mgpc_cell.h"

#include <math.h>
#include <string.h>

#include <vector>

using namespace cfdb;

namespace {

struct Host {
    int nl;
    MgpcLv lv[MGPC_MAX_LEVELS];
    std::vector<double> x[MGPC_MAX_LEVELS], b[MGPC_MAX_LEVELS], t[MGPC_MAX_LEVELS];
};

template <class F>
void interior(const MgpcLv& L, F f) {
    const int k0 = L.act[2] ? 1 : 0, k1 = L.act[2] ? L.n[2] : 0;
    for (int k = k0; k <= k1; k++)
        for (int j = 1; j <= L.n[1]; j++)
            for (int i = 1; i <= L.n[0]; i++) f(i, j, k);
}

// x = omega D^-1 b (first sweep from zero)
void jacobi0(const MgpcLv& L, double* x, const double* b) {
    const double om = mgpc_omega(L);
    interior(L, [&](int i, int j, int k) { x[mgpc_idx(L, i, j, k)] = om * b[mgpc_idx(L, i, j, k)] / mgpc_diag(L, i, j, k); });
}
// xn = x + omega D^-1 (b - K x)
void jacobi(const MgpcLv& L, double* xn, const double* x, const double* b) {
    const double om = mgpc_omega(L);
    interior(L, [&](int i, int j, int k) {
        double dg;
        const double au = mgpc_apply(L, x, i, j, k, dg);
        const long long c = mgpc_idx(L, i, j, k);
        xn[c] = x[c] + om * (b[c] - au) / dg;
    });
}
void residual(const MgpcLv& L, double* r, const double* x, const double* b) {
    interior(L, [&](int i, int j, int k) {
        double dg;
        const double au = mgpc_apply(L, x, i, j, k, dg);
        const long long c = mgpc_idx(L, i, j, k);
        r[c] = b[c] - au;
    });
}

// one V(1,1) cycle: x[l] = M^-1 b[l]
void vcycle(Host& h, int l) {
    const MgpcLv& L = h.lv[l];
    double *x = h.x[l].data(), *b = h.b[l].data(), *t = h.t[l].data();
    if (l == h.nl - 1) {
        jacobi0(L, x, b);
        for (int s = 1; s < MGPC_COARSE_SWEEPS; s++) {
            jacobi(L, t, x, b);
            memcpy(x, t, sizeof(double) * (size_t)L.cells);
        }
        return;
    }
    const MgpcLv& Lc = h.lv[l + 1];
    jacobi0(L, x, b);
    residual(L, t, x, b);
    double* bc = h.b[l + 1].data();
    interior(Lc, [&](int I, int J, int K) { bc[mgpc_idx(Lc, I, J, K)] = mgpc_restrict(L, t, I, J, K); });
    vcycle(h, l + 1);
    const double* ec = h.x[l + 1].data();
    interior(L, [&](int i, int j, int k) { x[mgpc_idx(L, i, j, k)] += mgpc_prolong(L, Lc, ec, i, j, k); });
    jacobi(L, t, x, b);
    memcpy(x, t, sizeof(double) * (size_t)L.cells);
}

}  // namespace

extern "C" {

// One V-cycle applied to r (session layout, zero boundary cells): z = M^-1 r.  For unit tests of the cycle.
int mgpc_host_apply(int nx, int ny, int nz, double dx, double dy, double dz, const double* r, double* z) {
    Host h;
    h.nl = mgpc_plan(nx, ny, nz, dx, dy, dz, h.lv);
    for (int l = 0; l < h.nl; l++) {
        h.x[l].assign((size_t)h.lv[l].cells, 0.0);
        h.b[l].assign((size_t)h.lv[l].cells, 0.0);
        h.t[l].assign((size_t)h.lv[l].cells, 0.0);
    }
    const MgpcLv& L = h.lv[0];
    interior(L, [&](int i, int j, int k) { h.b[0][mgpc_idx(L, i, j, k)] = r[mgpc_idx(L, i, j, k)]; });
    vcycle(h, 0);
    memset(z, 0, sizeof(double) * (size_t)L.cells);
    interior(L, [&](int i, int j, int k) { z[mgpc_idx(L, i, j, k)] = h.x[0][mgpc_idx(L, i, j, k)]; });
    return h.nl;
}

// The whole solve.  x: nx*ny*nz doubles in/out (boundary cells = Neumann ghosts), rhs likewise.  Returns the
// reference's status convention: 0 converged, 1 max iterations, 2 stagnated; out[0..2] = iterations, r0, r.
int mgpc_host_solve(int nx, int ny, int nz, double dx, double dy, double dz, double* x, const double* rhs, double rel_tol,
                    double abs_tol, int max_iter, int check, double* out) {
    Host h;
    h.nl = mgpc_plan(nx, ny, nz, dx, dy, dz, h.lv);
    for (int l = 0; l < h.nl; l++) {
        h.x[l].assign((size_t)h.lv[l].cells, 0.0);
        h.b[l].assign((size_t)h.lv[l].cells, 0.0);
        h.t[l].assign((size_t)h.lv[l].cells, 0.0);
    }
    const MgpcLv& L = h.lv[0];
    const size_t N = (size_t)L.cells;
    const bool d3 = nz > 1;
    const double idx2 = 1.0 / (dx * dx), idy2 = 1.0 / (dy * dy), idz2 = d3 ? 1.0 / (dz * dz) : 0.0;
    auto neumann = [&]() {  // boundary cell = nearest interior cell (linear_solver.c:348-392, closed form)
        for (int k = 0; k < nz; k++)
            for (int j = 0; j < ny; j++)
                for (int i = 0; i < nx; i++) {
                    const int ci = i < 1 ? 1 : (i > nx - 2 ? nx - 2 : i), cj = j < 1 ? 1 : (j > ny - 2 ? ny - 2 : j);
                    const int ck = d3 ? (k < 1 ? 1 : (k > nz - 2 ? nz - 2 : k)) : 0;
                    if (ci != i || cj != j || ck != k) x[mgpc_idx(L, i, j, k)] = x[mgpc_idx(L, ci, cj, ck)];
                }
    };
    auto lap = [&](const double* u, int i, int j, int k) {
        const long long c = mgpc_idx(L, i, j, k);
        double s = (u[c + 1] - 2.0 * u[c] + u[c - 1]) * idx2 + (u[c + L.sy] - 2.0 * u[c] + u[c - L.sy]) * idy2;
        if (d3) s += (u[c + L.sz] - 2.0 * u[c] + u[c - L.sz]) * idz2;
        return s;
    };
    auto dot = [&](const std::vector<double>& a, const std::vector<double>& b) {
        double s = 0.0;
        interior(L, [&](int i, int j, int k) { s += a[mgpc_idx(L, i, j, k)] * b[mgpc_idx(L, i, j, k)]; });
        return s;
    };
    std::vector<double> r(N, 0.0), p(N, 0.0), Ap(N, 0.0), z(N, 0.0);
    auto precond = [&]() {
        interior(L, [&](int i, int j, int k) { h.b[0][mgpc_idx(L, i, j, k)] = r[mgpc_idx(L, i, j, k)]; });
        vcycle(h, 0);
        interior(L, [&](int i, int j, int k) { z[mgpc_idx(L, i, j, k)] = h.x[0][mgpc_idx(L, i, j, k)]; });
    };
    neumann();
    interior(L, [&](int i, int j, int k) { r[mgpc_idx(L, i, j, k)] = -rhs[mgpc_idx(L, i, j, k)] + lap(x, i, j, k); });
    const double r0 = sqrt(dot(r, r));
    double tol = rel_tol * r0;
    if (tol < abs_tol) tol = abs_tol;
    out[0] = 0.0;
    out[1] = out[2] = r0;
    if (r0 < abs_tol) {
        neumann();
        return 0;
    }
    precond();
    p = z;
    double rz = dot(r, z);
    int status = 1;
    if (check < 1) check = 1;
    for (int it = 0; it < max_iter; it++) {
        interior(L, [&](int i, int j, int k) { Ap[mgpc_idx(L, i, j, k)] = -lap(p.data(), i, j, k); });
        const double pAp = dot(p, Ap);
        if (fabs(pAp) < 1e-30) {
            status = 2;
            break;
        }
        const double alpha = rz / pAp;
        interior(L, [&](int i, int j, int k) {
            const long long c = mgpc_idx(L, i, j, k);
            x[c] += alpha * p[c];
            r[c] -= alpha * Ap[c];
        });
        const double rn = sqrt(dot(r, r));
        out[0] = it + 1;
        out[2] = rn;
        if (it % check == 0 && rn < tol) {
            status = 0;
            break;
        }
        precond();
        const double rz_new = dot(r, z);
        if (fabs(rz) < 1e-30) {
            status = 2;
            break;
        }
        const double beta = rz_new / rz;
        interior(L, [&](int i, int j, int k) {
            const long long c = mgpc_idx(L, i, j, k);
            p[c] = z[c] + beta * p[c];
        });
        rz = rz_new;
    }
    if (status == 1 && out[2] < tol) status = 0;  // (the reference's final check between check intervals)
    neumann();
    return status;
}

}  // extern "C"
