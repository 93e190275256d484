// mgpc_cell.h -- per-cell arithmetic and level planning of the multigrid PRECONDITIONER of the CG solver
// (poisson_mgpc.cu).  Plain C++ (host + device): the CUDA kernels are loops over these functions, and
// oracle/mgpc_host.cpp runs the very same functions on the host as the CPU restatement the tests compare with.
//
// The system.  The reference's CG works on vectors whose boundary cells stay zero (linear_solver_cg.c:103-123:
// r, p, Ap are calloc'ed and only interior cells are written), i.e. on the homogeneous-DIRICHLET Laplacian with the
// wall ON the boundary cell -- not on the mirrored-ghost system of the stationary solvers (poisson_mg.cu).  The
// hierarchy is built for that operator in its symmetric finite-volume form
//
//     K u (i) = sum over axes a of  w_a(i) * [ (u_i - u_{i-1}) / hl_a + (u_i - u_{i+1}) / hr_a ],
//     w_a = product over the OTHER axes b of m_b,   m_b = (hl_b + hr_b) / 2
//
// which on the uniform fine grid is (cell volume) x (the reference's -lap).  Coarsening keeps the even interior
// cells of an axis (spacing 2H); where the number of nodes is odd the last coarse node is H + d instead of 2H away
// from the wall, and that distance d is all a level has to remember per axis: hr of the last node.  Transfer is
// linear interpolation by node POSITION (wall value 0) and its transpose.  tools/mgpcg_dirichlet_prototype.py is the
// numpy experiment behind it (7-8 PCG iterations on every grid against hundreds to thousands for the reference's
// Jacobi-PCG).
#pragma once

#if defined(__CUDACC__)
#define MGPC_HD __host__ __device__ __forceinline__
#else
#define MGPC_HD inline
#endif

namespace cfdb {

constexpr int MGPC_MAX_LEVELS = 32;
constexpr double MGPC_COUPLING = 0.5;    // an axis is coarsened when its 1/H^2 is at least this share of the strongest
constexpr int MGPC_MIN_NODES = 3;        // an axis with <= 3 interior nodes is not coarsened further
constexpr int MGPC_COARSE_SWEEPS = 40;   // damped-Jacobi sweeps on the coarsest level (<= 27 unknowns)

struct MgpcLv {
    int n[3];          // interior nodes per axis (1 on the inactive z axis of a 2-D grid)
    int act[3];        // the axis is part of the operator
    int N[3];          // stored extents: n + 2 on an active axis (zero wall cells), 1 otherwise
    long long sy, sz;  // strides in doubles (sz = 0 in 2-D)
    double H[3];       // node spacing
    double d[3];       // distance of the LAST node to the high wall (0 < d <= H)
    int coarsen[3];    // the axis is halved on the way to the next level
    long long cells;   // N[0] * N[1] * N[2]
};

MGPC_HD long long mgpc_idx(const MgpcLv& L, int i, int j, int k) { return (long long)k * L.sz + (long long)j * L.sy + i; }

// 1/hl, 1/hr and the volume factor of node i (1..n) on axis a
MGPC_HD void mgpc_coef(const MgpcLv& L, int a, int i, double& kl, double& kr, double& m) {
    const double hl = L.H[a], hr = (i == L.n[a]) ? L.d[a] : L.H[a];
    kl = 1.0 / hl;
    kr = 1.0 / hr;
    m = 0.5 * (hl + hr);
}

// (K u)(i,j,k) and the diagonal entry of that row; u has zero wall cells
MGPC_HD double mgpc_apply(const MgpcLv& L, const double* u, int i, int j, int k, double& diag) {
    const int ijk[3] = {i, j, k};
    double kl[3] = {0.0, 0.0, 0.0}, kr[3] = {0.0, 0.0, 0.0}, m[3] = {1.0, 1.0, 1.0};
    for (int a = 0; a < 3; a++)
        if (L.act[a]) mgpc_coef(L, a, ijk[a], kl[a], kr[a], m[a]);
    const long long c = mgpc_idx(L, i, j, k);
    const long long st[3] = {1, L.sy, L.sz};
    const double uc = u[c];
    double s = 0.0;
    diag = 0.0;
    for (int a = 0; a < 3; a++) {
        if (!L.act[a]) continue;
        const double w = m[(a + 1) % 3] * m[(a + 2) % 3];
        s += w * (kl[a] * (uc - u[c - st[a]]) + kr[a] * (uc - u[c + st[a]]));
        diag += w * (kl[a] + kr[a]);
    }
    return s;
}

// the diagonal entry of row (i,j,k) alone
MGPC_HD double mgpc_diag(const MgpcLv& L, int i, int j, int k) {
    const int ijk[3] = {i, j, k};
    double kl[3] = {0.0, 0.0, 0.0}, kr[3] = {0.0, 0.0, 0.0}, m[3] = {1.0, 1.0, 1.0};
    for (int a = 0; a < 3; a++)
        if (L.act[a]) mgpc_coef(L, a, ijk[a], kl[a], kr[a], m[a]);
    double diag = 0.0;
    for (int a = 0; a < 3; a++)
        if (L.act[a]) diag += m[(a + 1) % 3] * m[(a + 2) % 3] * (kl[a] + kr[a]);
    return diag;
}

// restriction weights of coarse node I on axis a: fine centre index i0 and the weights of i0-1, i0, i0+1
MGPC_HD void mgpc_rweights(const MgpcLv& Lf, int a, int I, int& i0, double (&w)[3]) {
    if (!Lf.coarsen[a]) {
        i0 = I;
        w[0] = 0.0;
        w[1] = 1.0;
        w[2] = 0.0;
        return;
    }
    i0 = 2 * I;
    w[0] = 0.5;
    w[1] = 1.0;
    if (i0 + 1 > Lf.n[a]) w[2] = 0.0;
    else if (i0 + 1 == Lf.n[a]) w[2] = Lf.d[a] / (Lf.H[a] + Lf.d[a]);  // the last fine node sits between me and the wall
    else w[2] = 0.5;
}

// (P^T r)(I,J,K)
MGPC_HD double mgpc_restrict(const MgpcLv& Lf, const double* rf, int I, int J, int K) {
    int i0, j0, k0;
    double wx[3], wy[3], wz[3];
    mgpc_rweights(Lf, 0, I, i0, wx);
    mgpc_rweights(Lf, 1, J, j0, wy);
    mgpc_rweights(Lf, 2, K, k0, wz);
    double s = 0.0;
    for (int dk = 0; dk < 3; dk++) {
        if (wz[dk] == 0.0) continue;
        for (int dj = 0; dj < 3; dj++) {
            if (wy[dj] == 0.0) continue;
            const long long row = mgpc_idx(Lf, 0, j0 + dj - 1, Lf.act[2] ? k0 + dk - 1 : 0);
            double t = 0.0;
            for (int di = 0; di < 3; di++)
                if (wx[di] != 0.0) t += wx[di] * rf[row + i0 + di - 1];
            s += wz[dk] * wy[dj] * t;
        }
    }
    return s;
}

// prolongation of axis a at fine node i: the two coarse nodes and their weights (a coarse index beyond the last node
// is the zero wall cell of the coarse array)
MGPC_HD void mgpc_pweights(const MgpcLv& Lf, int a, int i, int& I0, int& I1, double& w0, double& w1) {
    if (!Lf.coarsen[a]) {
        I0 = I1 = i;
        w0 = 1.0;
        w1 = 0.0;
    } else if ((i & 1) == 0) {
        I0 = I1 = i / 2;
        w0 = 1.0;
        w1 = 0.0;
    } else {
        I0 = (i - 1) / 2;
        I1 = (i + 1) / 2;
        w0 = (i == Lf.n[a]) ? Lf.d[a] / (Lf.H[a] + Lf.d[a]) : 0.5;
        w1 = 0.5;  // (times the zero wall cell when I1 is beyond the last coarse node)
    }
}

// (P e)(i,j,k)
MGPC_HD double mgpc_prolong(const MgpcLv& Lf, const MgpcLv& Lc, const double* ec, int i, int j, int k) {
    int I[2], J[2], K[2];
    double wx[2], wy[2], wz[2];
    mgpc_pweights(Lf, 0, i, I[0], I[1], wx[0], wx[1]);
    mgpc_pweights(Lf, 1, j, J[0], J[1], wy[0], wy[1]);
    if (Lf.act[2]) mgpc_pweights(Lf, 2, k, K[0], K[1], wz[0], wz[1]);
    else {
        K[0] = K[1] = 0;
        wz[0] = 1.0;
        wz[1] = 0.0;
    }
    double s = 0.0;
    for (int c = 0; c < 2; c++) {
        if (wz[c] == 0.0) continue;
        for (int b = 0; b < 2; b++) {
            if (wy[b] == 0.0) continue;
            for (int a = 0; a < 2; a++)
                if (wx[a] != 0.0) s += wz[c] * wy[b] * wx[a] * ec[mgpc_idx(Lc, I[a], J[b], K[c])];
        }
    }
    return s;
}

// Levels of a grid of nx x ny x nz cells (boundary cells included, nz = 1: 2-D) with spacings dx, dy, dz; returns
// the level count.  Level 0 has the session's own layout.
inline int mgpc_plan(int nx, int ny, int nz, double dx, double dy, double dz, MgpcLv* lv) {
    const int nn[3] = {nx, ny, nz};
    const double hh[3] = {dx, dy, dz};
    MgpcLv L;
    for (int a = 0; a < 3; a++) {
        L.act[a] = (a < 2) || (nz > 1);
        L.n[a] = L.act[a] ? nn[a] - 2 : 1;
        L.H[a] = L.act[a] ? hh[a] : 1.0;
        L.d[a] = L.H[a];
    }
    int nl = 0;
    for (;;) {
        bool any = false;
        for (int a = 0; a < 3; a++) {
            L.N[a] = L.act[a] ? L.n[a] + 2 : 1;
            L.coarsen[a] = L.act[a] && L.n[a] > MGPC_MIN_NODES;
        }
        {
            // semi-coarsening: only the axes whose coupling 1 / H^2 is within a factor 2 of the strongest one are
            // halved -- a point smoother only damps the error that oscillates along strongly coupled axes.  With
            // it anisotropic spacings take the 7-8 iterations of isotropic ones (514 x 130: 19 -> 7,
            // 130 x 34 x 83: 20 -> 8); the strongest coarsenable axis always qualifies, so the plan terminates.
            double cmax = 0.0;
            for (int a = 0; a < 3; a++)
                if (L.coarsen[a] && 1.0 / (L.H[a] * L.H[a]) > cmax) cmax = 1.0 / (L.H[a] * L.H[a]);
            for (int a = 0; a < 3; a++)
                if (L.coarsen[a] && 1.0 / (L.H[a] * L.H[a]) < MGPC_COUPLING * cmax) L.coarsen[a] = 0;
            for (int a = 0; a < 3; a++) any = any || L.coarsen[a];
        }
        L.sy = L.N[0];
        L.sz = L.act[2] ? (long long)L.N[0] * L.N[1] : 0;
        L.cells = (long long)L.N[0] * L.N[1] * L.N[2];
        if (!any || nl == MGPC_MAX_LEVELS - 1)
            for (int a = 0; a < 3; a++) L.coarsen[a] = 0;
        lv[nl++] = L;
        if (!any || nl == MGPC_MAX_LEVELS) break;
        for (int a = 0; a < 3; a++) {
            if (!L.coarsen[a]) continue;
            const int n = L.n[a];
            L.d[a] = (n & 1) ? L.H[a] + L.d[a] : L.d[a];  // odd: the last node is dropped, the wall moves away
            L.H[a] *= 2.0;
            L.n[a] = n / 2;
        }
    }
    return nl;
}

MGPC_HD double mgpc_omega(const MgpcLv& L) { return L.act[2] ? 0.85 : 0.8; }

}  // namespace cfdb
