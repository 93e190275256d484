// ns_kernels.cu -- Navier-Stokes step kernels: predictor, divergence/RHS, corrector (+energy,
// +NaN flag), thermal / periodic boundary conditions, RK2/RK4 stages, field statistics.
//
// Reference semantics followed (file:line under /root/reference/lib/src):
//   solvers/navier_stokes/cpu/solver_projection.c:116-289   projection step
//   solvers/navier_stokes/boundary_copy_utils.h:93-148      velocity boundary carry-over
//   solvers/energy/cpu/energy_solver.c:21-176, 185-196, 204-334
//   solvers/navier_stokes/momentum_rhs/ns_momentum_rhs_scalar.h:49-190, cpu/solver_rk{2,4}.c
//   solvers/navier_stokes/cpu/solver_explicit_euler.c:231-314 (periodic BC), 317-333 (source)
//   api/solver_registry.c:31-62 (stats)
//
// Same 2.5-D structure as poisson_kernels.cu: x-y tile per block, z marched with the
// z-1/z/z+1 values in registers, in-plane neighbours through L1.
#include "common.cuh"
#include "kernels.h"
#include "stencil_tma.cuh"

#include <limits.h>

namespace cfdb {
CFDB_ARITH_BEGIN

constexpr int TX = 32, TY = 8;

__device__ __forceinline__ bool ns_column(const Geom& g, int zchunk, int& i, int& j, int& kb, int& ke) {
    i = blockIdx.x * TX + threadIdx.x;
    j = blockIdx.y * TY + threadIdx.y;
    kb = g.k0 + blockIdx.z * zchunk;
    ke = min(kb + zchunk, g.k1);
    return i < g.nx && j < g.ny;
}
__device__ __forceinline__ bool interior_xy(const Geom& g, int i, int j) {
    return i >= 1 && i < g.nx - 1 && j >= 1 && j < g.ny - 1;
}

// One stencil operand marched along z.
struct March {
    double m, c, p;
};

// ---------------------------------------------------------------------------------------
// Predictor: u* = clamp(u + dt(-(u.grad)u + nu lap u + S))  (solver_projection.c:116-185),
// S = default sinusoidal source (tables) + Boussinesq buoyancy.  Cells on the x-y boundary
// and on physical z boundary planes carry u,v,w over unchanged (boundary_copy_utils.h:93-148).
// 48 B/cell (56 with T).
// ---------------------------------------------------------------------------------------
template <bool D3>
__global__ void __launch_bounds__(TX* TY)
k_predictor(Geom g, NsConst c, int zchunk, const double* __restrict__ u, const double* __restrict__ v,
            const double* __restrict__ w, const double* __restrict__ T,
            const double* __restrict__ su_tab, const double* __restrict__ sv_tab,
            double* __restrict__ us, double* __restrict__ vs, double* __restrict__ ws) {
    int i, j, kb, ke;
    if (!ns_column(g, zchunk, i, j, kb, ke)) return;
    const long long col = (long long)j * g.sy + i;
    if (!interior_xy(g, i, j)) {  // boundary column: carry over on every plane of this chunk
        const int k_lo = (D3 && kb == g.k0) ? kb - 1 : kb, k_hi = (D3 && ke == g.k1) ? ke + 1 : ke;
        for (int k = k_lo; k < k_hi; k++) {
            const long long n = (long long)k * g.sz + col;
            us[n] = u[n];
            vs[n] = v[n];
            ws[n] = w[n];
        }
        return;
    }
    long long n = (long long)kb * g.sz + col;
    if (D3 && kb == g.k0 && g.glo[2]) {  // physical low z face
        us[n - g.sz] = u[n - g.sz];
        vs[n - g.sz] = v[n - g.sz];
        ws[n - g.sz] = w[n - g.sz];
    }
    const double su = c.src_on ? su_tab[j] * c.src_decay : 0.0, sv = c.src_on ? sv_tab[i] * c.src_decay : 0.0;
    March U, V, W;
    U.m = D3 ? u[n - g.sz] : u[n]; U.c = u[n]; U.p = u[n];  // nz = 1: z neighbours = the cell itself
    V.m = D3 ? v[n - g.sz] : v[n]; V.c = v[n]; V.p = v[n];
    W.m = D3 ? w[n - g.sz] : w[n]; W.c = w[n]; W.p = w[n];
    for (int k = kb; k < ke; k++, n += g.sz) {
        if (D3) {
            U.p = u[n + g.sz];
            V.p = v[n + g.sz];
            W.p = w[n + g.sz];
        }
        const double ue = u[n + 1], uw = u[n - 1], un = u[n + g.sy], usth = u[n - g.sy];
        const double ve = v[n + 1], vw = v[n - 1], vn = v[n + g.sy], vsth = v[n - g.sy];
        const double we = w[n + 1], ww = w[n - 1], wn = w[n + g.sy], wsth = w[n - g.sy];
        const double uc = U.c, vc = V.c, wc = W.c;
        // operation order and the /(2.0*dx), /(dx*dx) divisions of solver_projection.c:120-171
        const double conv_u = uc * (ediv(ue - uw, g.two_dx, g.inv_2dx)) + vc * (ediv(un - usth, g.two_dy, g.inv_2dy)) +
                              wc * ((U.p - U.m) * g.inv_2dz);
        const double conv_v = uc * (ediv(ve - vw, g.two_dx, g.inv_2dx)) + vc * (ediv(vn - vsth, g.two_dy, g.inv_2dy)) +
                              wc * ((V.p - V.m) * g.inv_2dz);
        const double conv_w = uc * (ediv(we - ww, g.two_dx, g.inv_2dx)) + vc * (ediv(wn - wsth, g.two_dy, g.inv_2dy)) +
                              wc * ((W.p - W.m) * g.inv_2dz);
        const double visc_u = c.nu * (ediv(ue - 2.0 * uc + uw, g.dx2, g.inv_dx2) + ediv(un - 2.0 * uc + usth, g.dy2, g.inv_dy2) +
                                      (U.p - 2.0 * uc + U.m) * g.inv_dz2);
        const double visc_v = c.nu * (ediv(ve - 2.0 * vc + vw, g.dx2, g.inv_dx2) + ediv(vn - 2.0 * vc + vsth, g.dy2, g.inv_dy2) +
                                      (V.p - 2.0 * vc + V.m) * g.inv_dz2);
        const double visc_w = c.nu * (ediv(we - 2.0 * wc + ww, g.dx2, g.inv_dx2) + ediv(wn - 2.0 * wc + wsth, g.dy2, g.inv_dy2) +
                                      (W.p - 2.0 * wc + W.m) * g.inv_dz2);
        double fu = su, fv = sv, fw = 0.0;
        if (c.beta != 0.0) {
            const double dT = __ldg(T + n) - c.T_ref;
            fu += -c.beta * dT * c.gx;
            fv += -c.beta * dT * c.gy;
            fw += -c.beta * dT * c.gz;
        }
        us[n] = clamp100(uc + c.dt * (-conv_u + visc_u + fu));
        vs[n] = clamp100(vc + c.dt * (-conv_v + visc_v + fv));
        ws[n] = clamp100(wc + c.dt * (-conv_w + visc_w + fw));
        U.m = U.c; U.c = U.p;
        V.m = V.c; V.c = V.p;
        W.m = W.c; W.c = W.p;
    }
    if (D3 && ke == g.k1 && g.ghi[2]) {  // physical high z face (n is plane nz-1 now)
        us[n] = u[n];
        vs[n] = v[n];
        ws[n] = w[n];
    }
}

// rhs = (rho/dt) div(u*)  (solver_projection.c:195-214); boundary cells of rhs stay 0.  32 B/cell.
template <bool D3>
__global__ void __launch_bounds__(TX* TY)
k_divergence(Geom g, NsConst c, int zchunk, const double* __restrict__ us, const double* __restrict__ vs,
             const double* __restrict__ ws, double* __restrict__ rhs) {
    int i, j, kb, ke;
    if (!ns_column(g, zchunk, i, j, kb, ke) || !interior_xy(g, i, j)) return;
    long long n = (long long)kb * g.sz + (long long)j * g.sy + i;
    double wm = D3 ? ws[n - g.sz] : 0.0, wc = D3 ? ws[n] : 0.0, wp = 0.0;
    for (int k = kb; k < ke; k++, n += g.sz) {
        if (D3) wp = ws[n + g.sz];
        const double div = ediv(us[n + 1] - us[n - 1], g.two_dx, g.inv_2dx) + ediv(vs[n + g.sy] - vs[n - g.sy], g.two_dy, g.inv_2dy) +
                           (wp - wm) * g.inv_2dz;
        rhs[n] = c.rho_over_dt * div;
        wm = wc;
        wc = wp;
    }
}

// Corrector u = clamp(u* - (dt/rho) grad p) (solver_projection.c:230-250), fused with the
// explicit-Euler temperature step (energy_solver.c:120-160, uses the corrected velocity at the
// cell centre) and the NaN/Inf scan of u,v,w,p (:281-289) and of T_new (energy_solver.c:163-170).
// 56 B/cell (72 with energy).
template <bool D3, bool ENERGY>
__global__ void __launch_bounds__(TX* TY)
k_corrector(Geom g, NsConst c, int zchunk, const double* __restrict__ us, const double* __restrict__ vs,
            const double* __restrict__ ws, const double* __restrict__ p, double* __restrict__ u,
            double* __restrict__ v, double* __restrict__ w, const double* __restrict__ T,
            double* __restrict__ Tn, Ctl* ctl) {
    int i, j, kb, ke;
    if (!ns_column(g, zchunk, i, j, kb, ke)) return;
    const long long col = (long long)j * g.sy + i;
    bool bad = false;
    if (!interior_xy(g, i, j)) {
        // boundary column: u,v,w keep the caller's values; scan them and p; carry T over.
        const int k_lo = (D3 && kb == g.k0) ? kb - 1 : kb, k_hi = (D3 && ke == g.k1) ? ke + 1 : ke;
        for (int k = k_lo; k < k_hi; k++) {
            const long long n = (long long)k * g.sz + col;
            bad |= !(isfinite(u[n]) && isfinite(v[n]) && isfinite(w[n]) && isfinite(p[n]));
            if (ENERGY) {
                const double t = T[n];
                Tn[n] = t;
                bad |= !isfinite(t);
            }
        }
    } else {
        long long n = (long long)kb * g.sz + col;
        if (D3 && kb == g.k0) {
            const long long b = n - g.sz;
            bad |= !(isfinite(u[b]) && isfinite(v[b]) && isfinite(w[b]) && isfinite(p[b]));
            if (ENERGY) { Tn[b] = T[b]; bad |= !isfinite(T[b]); }
        }
        March P, Q;
        P.c = p[n]; P.m = D3 ? p[n - g.sz] : P.c; P.p = P.c;  // nz = 1: z neighbours = the cell itself
        Q.m = Q.c = Q.p = 0.0;
        if (ENERGY) { Q.c = T[n]; Q.m = D3 ? T[n - g.sz] : Q.c; Q.p = Q.c; }
        for (int k = kb; k < ke; k++, n += g.sz) {
            if (D3) {
                P.p = p[n + g.sz];
                if (ENERGY) Q.p = T[n + g.sz];
            }
            const double un = clamp100(us[n] - c.dt_over_rho * (ediv(p[n + 1] - p[n - 1], g.two_dx, g.inv_2dx)));
            const double vn = clamp100(vs[n] - c.dt_over_rho * (ediv(p[n + g.sy] - p[n - g.sy], g.two_dy, g.inv_2dy)));
            const double wn = clamp100(ws[n] - c.dt_over_rho * ((P.p - P.m) * g.inv_2dz));
            u[n] = un;
            v[n] = vn;
            w[n] = wn;
            bad |= !(isfinite(un) && isfinite(vn) && isfinite(wn) && isfinite(P.c));
            if (ENERGY) {
                const double Te = T[n + 1], Tw = T[n - 1], Tnn = T[n + g.sy], Ts = T[n - g.sy], Tc = Q.c;
                const double adv = un * ((Te - Tw) * g.inv_2dx) + vn * ((Tnn - Ts) * g.inv_2dy) +
                                   wn * ((Q.p - Q.m) * g.inv_2dz);
                const double dif = c.alpha * ((Te - 2.0 * Tc + Tw) * g.inv_dx2 + (Tnn - 2.0 * Tc + Ts) * g.inv_dy2 +
                                              (Q.p - 2.0 * Tc + Q.m) * g.inv_dz2);
                const double tn = Tc + c.dt * (-adv + dif);
                Tn[n] = tn;
                bad |= !isfinite(tn);
                Q.m = Q.c; Q.c = Q.p;
            }
            P.m = P.c; P.c = P.p;
        }
        if (D3 && ke == g.k1) {  // n is plane nz-1
            bad |= !(isfinite(u[n]) && isfinite(v[n]) && isfinite(w[n]) && isfinite(p[n]));
            if (ENERGY) { Tn[n] = T[n]; bad |= !isfinite(T[n]); }
        }
    }
    if (bad) ctl->nonfinite = 1;
}

// Stand-alone temperature step (RK steppers: after the combine, solver_rk4.c:222-230).
template <bool D3>
__global__ void __launch_bounds__(TX* TY)
k_energy(Geom g, NsConst c, int zchunk, const double* __restrict__ u, const double* __restrict__ v,
         const double* __restrict__ w, const double* __restrict__ T, double* __restrict__ Tn, Ctl* ctl) {
    int i, j, kb, ke;
    if (!ns_column(g, zchunk, i, j, kb, ke)) return;
    const long long col = (long long)j * g.sy + i;
    bool bad = false;
    if (!interior_xy(g, i, j)) {
        const int k_lo = (D3 && kb == g.k0) ? kb - 1 : kb, k_hi = (D3 && ke == g.k1) ? ke + 1 : ke;
        for (int k = k_lo; k < k_hi; k++) {
            const long long n = (long long)k * g.sz + col;
            Tn[n] = T[n];
            bad |= !isfinite(T[n]);
        }
    } else {
        long long n = (long long)kb * g.sz + col;
        if (D3 && kb == g.k0) { Tn[n - g.sz] = T[n - g.sz]; bad |= !isfinite(T[n - g.sz]); }
        March Q;
        Q.c = T[n]; Q.m = D3 ? T[n - g.sz] : Q.c; Q.p = Q.c;
        for (int k = kb; k < ke; k++, n += g.sz) {
            if (D3) Q.p = T[n + g.sz];
            const double Te = T[n + 1], Tw = T[n - 1], Tnn = T[n + g.sy], Ts = T[n - g.sy], Tc = Q.c;
            const double adv = u[n] * ((Te - Tw) * g.inv_2dx) + v[n] * ((Tnn - Ts) * g.inv_2dy) +
                               w[n] * ((Q.p - Q.m) * g.inv_2dz);
            const double dif = c.alpha * ((Te - 2.0 * Tc + Tw) * g.inv_dx2 + (Tnn - 2.0 * Tc + Ts) * g.inv_dy2 +
                                          (Q.p - 2.0 * Tc + Q.m) * g.inv_dz2);
            const double tn = Tc + c.dt * (-adv + dif);
            Tn[n] = tn;
            bad |= !isfinite(tn);
            Q.m = Q.c; Q.c = Q.p;
        }
        if (D3 && ke == g.k1) { Tn[n] = T[n]; bad |= !isfinite(T[n]); }
    }
    if (bad) ctl->nonfinite = 1;
}


#include "ns_tma.inc"

// ---- host launchers ----
static inline dim3 blk2() { return dim3(TX, TY); }
static inline dim3 tma_grid2(const LaunchCtx& L) {
    const int nzi = L.g.k1 - L.g.k0, ty = (L.g.ny - 2 + Tile::BY - 1) / Tile::BY;
    if (!L.g.dim3)  // 2-D: zchunk_tma counts tiles along y
        return dim3(Tile::GX(L.g.nx), (ty + L.zchunk_tma - 1) / L.zchunk_tma, 1);
    return dim3(Tile::GX(L.g.nx), ty, (nzi + L.zchunk_tma - 1) / L.zchunk_tma);
}
template <int N>
static inline bool tma_maps2(const LaunchCtx& L, const double* const (&ptr)[N], TmaSet<N>& out) {
    if (!L.tc || !L.tc->usable) return false;
    for (int q = 0; q < N; q++) {
        const CUtensorMap* m = L.tc->get(ptr[q]);
        if (!m) return false;
        out.m[q] = *m;
    }
    return true;
}
static inline int face_blocks(const Geom& g) { return (int)((face_cells(g) + 255) / 256); }
void launch_predictor(const LaunchCtx& L, const NsConst& c, const double* u, const double* v, const double* w,
                      const double* T, const double* su_tab, const double* sv_tab, double* us, double* vs,
                      double* ws) {
    {
        TmaSet<3> tm;
        const double* const ops[3] = {u, v, w};
        if (tma_maps2(L, ops, tm)) {
            constexpr size_t SM = DirectPipe<3>::SMEM;
            CFDB_ONCE_PER_DEVICE(tma_optin(k_predictor_tma<true>, SM), tma_optin(k_predictor_tma<false>, SM));
            LaunchCtx L1 = L;  // one block per SM (93 registers x 544 threads): its own chunk length
            L1.zchunk_tma = L.zchunk_tma1;
            if (L.g.dim3) CFDB_LAUNCH(KK_PREDICTOR, L.st, (k_predictor_tma<true><<<tma_grid2(L1), Tile::THREADS_PW, SM, L.st>>>(tm, L.g, c, L1.zchunk_tma, T, su_tab, sv_tab, us, vs, ws)));
            else CFDB_LAUNCH(KK_PREDICTOR, L.st, (k_predictor_tma<false><<<tma_grid2(L1), Tile::THREADS_PW, SM, L.st>>>(tm, L.g, c, L1.zchunk_tma, T, su_tab, sv_tab, us, vs, ws)));
            CFDB_LAUNCH(KK_BOUNDARY, L.st, (k_boundary_carry3<<<face_blocks(L.g), 256, 0, L.st>>>(L.g, u, v, w, us, vs, ws)));
            return;
        }
    }
    if (L.g.dim3) CFDB_LAUNCH(KK_PREDICTOR, L.st, k_predictor<true><<<L.grid(), blk2(), 0, L.st>>>(L.g, c, L.zchunk, u, v, w, T, su_tab, sv_tab, us, vs, ws));
    else CFDB_LAUNCH(KK_PREDICTOR, L.st, k_predictor<false><<<L.grid(), blk2(), 0, L.st>>>(L.g, c, L.zchunk, u, v, w, T, su_tab, sv_tab, us, vs, ws));
}
void launch_boundary_carry3(const LaunchCtx& L, const double* u, const double* v, const double* w, double* us,
                            double* vs, double* ws) {
    CFDB_LAUNCH(KK_BOUNDARY, L.st, (k_boundary_carry3<<<face_blocks(L.g), 256, 0, L.st>>>(L.g, u, v, w, us, vs, ws)));
}
void launch_boundary_scan(const LaunchCtx& L, const double* u, const double* v, const double* w, const double* p) {
    CFDB_LAUNCH(KK_BOUNDARY, L.st, (k_boundary_scan<<<face_blocks(L.g), 256, 0, L.st>>>(L.g, u, v, w, p, nullptr, nullptr, L.ctl)));
}
void launch_divergence(const LaunchCtx& L, const NsConst& c, const double* us, const double* vs,
                       const double* ws, double* rhs) {
    {
        TmaSet<3> tm;
        const double* const ops[3] = {us, vs, ws};
        if (tma_maps2(L, ops, tm)) {
            constexpr size_t SM = DirectPipe<3>::SMEM;
            CFDB_ONCE_PER_DEVICE(tma_optin(k_divergence_tma<true>, SM), tma_optin(k_divergence_tma<false>, SM));
            if (L.g.dim3) CFDB_LAUNCH(KK_DIVERGENCE, L.st, (k_divergence_tma<true><<<tma_grid2(L), Tile::THREADS_PW, SM, L.st>>>(tm, L.g, c, L.zchunk_tma, rhs)));
            else CFDB_LAUNCH(KK_DIVERGENCE, L.st, (k_divergence_tma<false><<<tma_grid2(L), Tile::THREADS_PW, SM, L.st>>>(tm, L.g, c, L.zchunk_tma, rhs)));
            return;
        }
    }
    if (L.g.dim3) CFDB_LAUNCH(KK_DIVERGENCE, L.st, k_divergence<true><<<L.grid(), blk2(), 0, L.st>>>(L.g, c, L.zchunk, us, vs, ws, rhs));
    else CFDB_LAUNCH(KK_DIVERGENCE, L.st, k_divergence<false><<<L.grid(), blk2(), 0, L.st>>>(L.g, c, L.zchunk, us, vs, ws, rhs));
}
void launch_corrector(const LaunchCtx& L, const NsConst& c, const double* us, const double* vs,
                      const double* ws, const double* p, double* u, double* v, double* w, const double* T,
                      double* Tn) {
    const bool en = c.alpha > 0.0;
    if (en) {
        TmaSet<2> tm;
        const double* const ops[2] = {p, T};
        if (tma_maps2(L, ops, tm)) {
            constexpr size_t SM = DirectPipe<2>::SMEM;
            CFDB_ONCE_PER_DEVICE(tma_optin(k_corrector_tma<true, true>, SM), tma_optin(k_corrector_tma<false, true>, SM));
            if (L.g.dim3) CFDB_LAUNCH(KK_CORRECTOR, L.st, (k_corrector_tma<true, true><<<tma_grid2(L), Tile::THREADS_PW, SM, L.st>>>(tm, L.g, c, L.zchunk_tma, us, vs, ws, u, v, w, Tn, L.ctl)));
            else CFDB_LAUNCH(KK_CORRECTOR, L.st, (k_corrector_tma<false, true><<<tma_grid2(L), Tile::THREADS_PW, SM, L.st>>>(tm, L.g, c, L.zchunk_tma, us, vs, ws, u, v, w, Tn, L.ctl)));
            CFDB_LAUNCH(KK_BOUNDARY, L.st, (k_boundary_scan<<<face_blocks(L.g), 256, 0, L.st>>>(L.g, u, v, w, p, T, Tn, L.ctl)));
            return;
        }
    } else {
        TmaSet<1> tm;
        const double* const ops[1] = {p};
        if (tma_maps2(L, ops, tm)) {
            constexpr size_t SM = DirectPipe<1>::SMEM;
            CFDB_ONCE_PER_DEVICE(tma_optin(k_corrector_tma<true, false>, SM), tma_optin(k_corrector_tma<false, false>, SM));
            if (L.g.dim3) CFDB_LAUNCH(KK_CORRECTOR, L.st, (k_corrector_tma<true, false><<<tma_grid2(L), Tile::THREADS_PW, SM, L.st>>>(tm, L.g, c, L.zchunk_tma, us, vs, ws, u, v, w, Tn, L.ctl)));
            else CFDB_LAUNCH(KK_CORRECTOR, L.st, (k_corrector_tma<false, false><<<tma_grid2(L), Tile::THREADS_PW, SM, L.st>>>(tm, L.g, c, L.zchunk_tma, us, vs, ws, u, v, w, Tn, L.ctl)));
            CFDB_LAUNCH(KK_BOUNDARY, L.st, (k_boundary_scan<<<face_blocks(L.g), 256, 0, L.st>>>(L.g, u, v, w, p, nullptr, nullptr, L.ctl)));
            return;
        }
    }
    if (L.g.dim3) {
        if (en) CFDB_LAUNCH(KK_CORRECTOR, L.st, k_corrector<true, true><<<L.grid(), blk2(), 0, L.st>>>(L.g, c, L.zchunk, us, vs, ws, p, u, v, w, T, Tn, L.ctl));
        else CFDB_LAUNCH(KK_CORRECTOR, L.st, k_corrector<true, false><<<L.grid(), blk2(), 0, L.st>>>(L.g, c, L.zchunk, us, vs, ws, p, u, v, w, T, Tn, L.ctl));
    } else {
        if (en) CFDB_LAUNCH(KK_CORRECTOR, L.st, k_corrector<false, true><<<L.grid(), blk2(), 0, L.st>>>(L.g, c, L.zchunk, us, vs, ws, p, u, v, w, T, Tn, L.ctl));
        else CFDB_LAUNCH(KK_CORRECTOR, L.st, k_corrector<false, false><<<L.grid(), blk2(), 0, L.st>>>(L.g, c, L.zchunk, us, vs, ws, p, u, v, w, T, Tn, L.ctl));
    }
}
void launch_energy(const LaunchCtx& L, const NsConst& c, const double* u, const double* v, const double* w,
                   const double* T, double* Tn) {
    {
        TmaSet<1> tm;
        const double* const ops[1] = {T};
        if (tma_maps2(L, ops, tm)) {
            constexpr size_t SM = DirectPipe<1>::SMEM;
            CFDB_ONCE_PER_DEVICE(tma_optin(k_energy_tma<true>, SM), tma_optin(k_energy_tma<false>, SM));
            if (L.g.dim3) CFDB_LAUNCH(KK_CORRECTOR, L.st, (k_energy_tma<true><<<tma_grid2(L), Tile::THREADS_PW, SM, L.st>>>(tm, L.g, c, L.zchunk_tma, u, v, w, Tn, L.ctl)));
            else CFDB_LAUNCH(KK_CORRECTOR, L.st, (k_energy_tma<false><<<tma_grid2(L), Tile::THREADS_PW, SM, L.st>>>(tm, L.g, c, L.zchunk_tma, u, v, w, Tn, L.ctl)));
            // T carry-over and finiteness on boundary cells (u stands in for p: scanned twice, harmless)
            CFDB_LAUNCH(KK_BOUNDARY, L.st, (k_boundary_scan<<<face_blocks(L.g), 256, 0, L.st>>>(L.g, u, v, w, u, T, Tn, L.ctl)));
            return;
        }
    }
    if (L.g.dim3) CFDB_LAUNCH(KK_CORRECTOR, L.st, k_energy<true><<<L.grid(), blk2(), 0, L.st>>>(L.g, c, L.zchunk, u, v, w, T, Tn, L.ctl));
    else CFDB_LAUNCH(KK_CORRECTOR, L.st, k_energy<false><<<L.grid(), blk2(), 0, L.st>>>(L.g, c, L.zchunk, u, v, w, T, Tn, L.ctl));
}

CFDB_ARITH_END
}  // namespace cfdb
