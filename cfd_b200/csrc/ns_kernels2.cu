// ns_kernels2.cu -- RK2/RK4 stage kernel, periodic / thermal boundary conditions, field
// statistics, synthetic initial state, small utilities.  See ns_kernels.cu for the citations.
#include "common.cuh"
#include "kernels.h"
#include "ns_args.h"
#include "stencil_tma.cuh"

#include <limits.h>

namespace cfdb {
CFDB_ARITH_BEGIN

constexpr int TX = 32, TY = 8;

struct March2 {
    double m, c, p;
};


// ---------------------------------------------------------------------------------------
// RK2 / RK4 stage, fused: k = RHS(Q_s) (ns_momentum_rhs_scalar.h:49-190, periodic-wrapped
// neighbours, clamps), running weighted sum acc (same association order as the reference's
// k1 + 2 k2 + 2 k3 + k4, solver_rk4.c:211-214), next stage state Q0 + f k or the final
// combine (velocity clamp +-100).  Boundary cells are never read by the RHS (wrapped indices)
// and are overwritten by the periodic BC after the step, so only interior cells are touched.
// ---------------------------------------------------------------------------------------
template <bool D3>
__global__ void __launch_bounds__(TX* TY)
k_rk_stage(Geom g, NsConst c, RkArgs ra, int zchunk, P4 qin, P4 q0, W4 acc, W4 qout,
           const double* __restrict__ rho, const double* __restrict__ T,
           const double* __restrict__ su_tab, const double* __restrict__ sv_tab, Ctl* ctl) {
    const int i = blockIdx.x * TX + threadIdx.x, j = blockIdx.y * TY + threadIdx.y;
    const int kb = g.k0 + blockIdx.z * zchunk, ke = min(kb + zchunk, g.k1);
    if (!(i >= 1 && i < g.nx - 1 && j >= 1 && j < g.ny - 1)) return;
    long long n = (long long)kb * g.sz + (long long)j * g.sy + i;
    const long long oxm = (i > 1) ? -1 : (g.nx - 3), oxp = (i < g.nx - 2) ? 1 : -(g.nx - 3);
    const long long oym = (j > 1 || !ra.wrap_y) ? -g.sy : (g.ny - 3) * g.sy;
    const long long oyp = (j < g.ny - 2 || !ra.wrap_y) ? g.sy : -(g.ny - 3) * g.sy;
    const double su = c.src_on ? su_tab[j] * ra.src_decay : 0.0;
    const double sv = c.src_on ? sv_tab[i] * ra.src_decay : 0.0;
    March2 F[4];
#pragma unroll
    for (int q = 0; q < 4; q++) {
        const long long zm = (kb > 1 || !ra.wrap_z) ? n - g.sz : n + (long long)(g.nz - 3) * g.sz;
        F[q].c = qin.a[q][n];
        F[q].m = D3 ? qin.a[q][zm] : F[q].c;  // nz = 1: the reference's k wraps onto k itself
        F[q].p = F[q].c;
    }
    bool bad = false;
    for (int k = kb; k < ke; k++, n += g.sz) {
        double d1[12], d2[9];  // d1[3q+axis]: first derivatives of u,v,w,p; d2[3q+axis]: second derivatives of u,v,w
        const long long zp = (k < g.nz - 2 || !ra.wrap_z) ? n + g.sz : n - (long long)(g.nz - 3) * g.sz;
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const double* f = qin.a[q];
            if (D3) F[q].p = f[zp];
            const double e = f[n + oxp], w_ = f[n + oxm], nn = f[n + oyp], s_ = f[n + oym], cc = F[q].c;
            d1[3 * q + 0] = ediv(e - w_, g.two_dx, g.inv_2dx);
            d1[3 * q + 1] = ediv(nn - s_, g.two_dy, g.inv_2dy);
            d1[3 * q + 2] = (F[q].p - F[q].m) * g.inv_2dz;
            if (q < 3) {
                d2[3 * q + 0] = ediv(e - 2.0 * cc + w_, g.dx2, g.inv_dx2);
                d2[3 * q + 1] = ediv(nn - 2.0 * cc + s_, g.dy2, g.inv_dy2);
                d2[3 * q + 2] = (F[q].p - 2.0 * cc + F[q].m) * g.inv_dz2;
            }
        }
        clamp_group<HI_100>(d1);
        clamp_group<HI_1000>(d2);
        const double rh = __ldg(rho + n);
        double kq[4] = {0.0, 0.0, 0.0, 0.0};
        if (rh > 1e-10) {
            const RhoDiv over_rho(rh);
            const double nu = min_one(over_rho(c.mu));  // fmin(mu / fmax(rho, 1e-10), 1): rho > 1e-10 here
            double fu = su, fv = sv, fw = 0.0;
            if (c.beta != 0.0) {
                const double dT = __ldg(T + n) - c.T_ref;
                fu += -c.beta * dT * c.gx;
                fv += -c.beta * dT * c.gy;
                fw += -c.beta * dT * c.gz;
            }
            const double uc = F[0].c, vc = F[1].c, wc = F[2].c;
            const double src[3] = {fu, fv, fw};
#pragma unroll
            for (int q = 0; q < 3; q++)
                kq[q] = -uc * d1[3 * q] - vc * d1[3 * q + 1] - wc * d1[3 * q + 2] - over_rho(d1[9 + q]) +
                        nu * (d2[3 * q] + d2[3 * q + 1] + d2[3 * q + 2]) + src[q];
            const double div = clampb<HI_10>(d1[0] + d1[4] + d1[8]);
            kq[3] = -0.1 * rh * div;
        }
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const double base = ra.first ? F[q].c : __ldg(q0.a[q] + n);
            double out;
            if (ra.last) {
                const double a_ = ra.first ? 0.0 : acc.a[q][n];
                out = base + ra.w_final * (a_ + kq[q]);
            } else {
                acc.a[q][n] = ra.first ? kq[q] : acc.a[q][n] + ra.w_k * kq[q];
                out = base + ra.fac_next * kq[q];
            }
            if (q < 3) out = clamp100(out);
            qout.a[q][n] = out;
            bad |= !isfinite(out);
            F[q].m = F[q].c;
            F[q].c = F[q].p;
        }
    }
    if (bad && ra.last) ctl->nonfinite = 1;
}


// ---------------------------------------------------------------------------------------
// TMA plane-pipeline version of the RK stage.  The reference RHS reads periodic-WRAPPED neighbour
// indices (ns_momentum_rhs_scalar.h:80-90), i.e. never the boundary cells; here the boundary cells
// of every stage input are first set to their periodic images (k_periodic_faces, O(N^(2/3))), so
// the plain halo of the TMA box delivers exactly the wrapped values.  Same arithmetic and
// operation order as k_rk_stage above.
// ---------------------------------------------------------------------------------------
// STG = 3-D: Q0 and the running sum of the NEXT plane are staged in shared memory with 8-byte
// asynchronous copies (no destination registers), issued when the current plane's values have
// been consumed -- with 128 registers per thread the compiler otherwise issues these 8 loads per
// cell right before their use (or spills them on arrival) and every warp of the block sits on
// DRAM latency at the same time (measured: 40% of the stall samples).
struct RkStage {
    static constexpr int PER_CELL = 8;                                   // Q0[4], acc[4]
    static constexpr size_t OFFSET = (DirectPipe<4>::SMEM + 15) & ~(size_t)15;
    static constexpr size_t BYTES = (size_t)2 * PER_CELL * Tile::THREADS * 8;  // one slot per own cell
    static constexpr size_t SMEM = OFFSET + BYTES;
};

template <bool STG>
struct RkOp {
    static constexpr bool PRODUCER_WARP = false;  // needs > 96 registers: 512-thread variant of the pipeline
    static constexpr bool STAGED = STG;
    static constexpr int NF = 4, NPW = 2;  // prefetched pointwise operands: rho, T
    const Geom& g;
    const NsConst& c;
    const RkArgs& ra;
    const P4& q0;
    const W4& acc;
    const W4& qout;
    const double* __restrict__ rho;
    const double* __restrict__ T;
    const double* __restrict__ su_tab;
    const double* __restrict__ sv_tab;
    double* stage;  // this thread's column of the staging area: [cell][8][THREADS]
    int ke, y0;     // end of the block's z-chunk, first row of its tile
    bool both;      // both own cells of the thread are active
    bool bad;
    __device__ __forceinline__ void pw_load(long long n, double (&pw)[2]) const {
        pw[0] = __ldg(rho + n);
        pw[1] = (c.beta != 0.0) ? __ldg(T + n) : 0.0;
    }
    __device__ __forceinline__ void issue(long long n, int cs) const {
        double* st = stage + cs * RkStage::PER_CELL * Tile::THREADS;
#pragma unroll
        for (int q = 0; q < 4; q++) {
            cp_async8(st + q * Tile::THREADS, q0.a[q] + n);
            cp_async8(st + (4 + q) * Tile::THREADS, acc.a[q] + n);
        }
        cp_async_commit();
    }
    __device__ __forceinline__ void prime(long long n, int cs) const {
        if (!ra.first) issue(n, cs);
    }
    __device__ __forceinline__ void cell(const St<4>& s, const double (&pw)[2], long long n, int i, int j, int k) {
        double b0[4] = {0.0, 0.0, 0.0, 0.0}, a0[4] = {0.0, 0.0, 0.0, 0.0};  // Q0 and the running sum
        if (!STG) {
#pragma unroll
            for (int q = 0; q < 4; q++) {
                b0[q] = ra.first ? 0.0 : __ldg(q0.a[q] + n);
                a0[q] = ra.first ? 0.0 : acc.a[q][n];
            }
        }
        double d1[12], d2[9];
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const double cc = s.c(q);
            d1[3 * q + 0] = ediv(s.e(q) - s.w(q), g.two_dx, g.inv_2dx);
            d1[3 * q + 1] = ediv(s.n(q) - s.s(q), g.two_dy, g.inv_2dy);
            d1[3 * q + 2] = (s.p(q) - s.m(q)) * g.inv_2dz;
            if (q < 3) {
                d2[3 * q + 0] = ediv(s.e(q) - 2.0 * cc + s.w(q), g.dx2, g.inv_dx2);
                d2[3 * q + 1] = ediv(s.n(q) - 2.0 * cc + s.s(q), g.dy2, g.inv_dy2);
                d2[3 * q + 2] = (s.p(q) - 2.0 * cc + s.m(q)) * g.inv_dz2;
            }
        }
        clamp_group<HI_100>(d1);
        clamp_group<HI_1000>(d2);
        const double rh = pw[0];
        double kq[4] = {0.0, 0.0, 0.0, 0.0};
        if (rh > 1e-10) {
            const RhoDiv over_rho(rh);
            const double nu = min_one(over_rho(c.mu));
            double fu = c.src_on ? su_tab[j] * ra.src_decay : 0.0, fv = c.src_on ? sv_tab[i] * ra.src_decay : 0.0, fw = 0.0;
            if (c.beta != 0.0) {
                const double dT = pw[1] - c.T_ref;
                fu += -c.beta * dT * c.gx;
                fv += -c.beta * dT * c.gy;
                fw += -c.beta * dT * c.gz;
            }
            const double uc = s.c(0), vc = s.c(1), wc = s.c(2);
            const double src[3] = {fu, fv, fw};
#pragma unroll
            for (int q = 0; q < 3; q++)
                kq[q] = -uc * d1[3 * q] - vc * d1[3 * q + 1] - wc * d1[3 * q + 2] - over_rho(d1[9 + q]) +
                        nu * (d2[3 * q] + d2[3 * q + 1] + d2[3 * q + 2]) + src[q];
            const double div = clampb<HI_10>(d1[0] + d1[4] + d1[8]);
            kq[3] = -0.1 * rh * div;
        }
        const int cs = (j - 1 - y0) >> 3;  // which of the thread's two cells (rows ly and ly + 8)
        const bool has_next = (k + 1 < ke);
        if (STG && !ra.first) {
            // groups complete in issue order; the most recent one belongs to the thread's other cell
            if (has_next && both) cp_async_wait<1>(); else cp_async_wait<0>();
            const double* st = stage + cs * RkStage::PER_CELL * Tile::THREADS;
#pragma unroll
            for (int q = 0; q < 4; q++) {
                b0[q] = st[q * Tile::THREADS];
                a0[q] = st[(4 + q) * Tile::THREADS];
            }
        }
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const double base = ra.first ? s.c(q) : b0[q];
            double out;
            if (ra.last) {
                const double a_ = ra.first ? 0.0 : a0[q];
                out = base + ra.w_final * (a_ + kq[q]);
            } else {
                acc.a[q][n] = ra.first ? kq[q] : a0[q] + ra.w_k * kq[q];
                out = base + ra.fac_next * kq[q];
            }
            if (q < 3) out = clamp100(out);  // (always finite afterwards)
            else bad |= (abs_hi(out) >= HI_INF);
            qout.a[q][n] = out;
        }
        // the slot's values are in registers (consumed above): refill it with the next plane's
        if (STG && !ra.first && has_next) issue(n + g.sz, cs);
    }
};

template <bool D3>
__global__ void __launch_bounds__(Tile::THREADS, 1)
k_rk_stage_tma(const __grid_constant__ TmaSet<4> tm, Geom g, NsConst c, RkArgs ra, int zchunk, P4 q0, W4 acc,
               W4 qout, const double* __restrict__ rho, const double* __restrict__ T,
               const double* __restrict__ su_tab, const double* __restrict__ sv_tab, Ctl* ctl) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int tid = threadIdx.x;
    const int i = Tile::X0(blockIdx.x) + Tile::XO + (tid & (Tile::BX - 1)), y0 = blockIdx.y * Tile::BY;
    const int j0 = y0 + 1 + (tid >> 6);
    RkOp<D3> op{g, c, ra, q0, acc, qout, rho, T, su_tab, sv_tab,
                reinterpret_cast<double*>(smem_raw + RkStage::OFFSET) + tid,
                min(g.k0 + ((int)blockIdx.z + 1) * zchunk, g.k1), y0,
                (i >= 1) && (i < g.nx - 1) && (j0 + 8 < g.ny - 1), false};
    tma_direct_sweep<D3>(tm, g, zchunk, op);
    if (op.bad && ra.last) ctl->nonfinite = 1;
}

// ---------------------------------------------------------------------------------------
// Explicit (forward) Euler step, cpu/solver_explicit_euler.c:396-524, one pass: the right-hand
// side has the RK stage's terms and clamps but reads the PLAIN neighbours (boundary cells
// included, no index wrap), every increment is limited to +-1 (UPDATE_LIMIT), the velocity to
// +-100, and the pressure moves by clamp(-0.1 dt rho div, +-1).  Cells with rho <= 1e-10 keep
// their state (:399-401).  Interior cells only; out of place.
// ---------------------------------------------------------------------------------------
struct EulerCell {
    // d1[3q+axis]: raw first derivatives of u,v,w,p; d2[3q+axis]: raw second derivatives of u,v,w (clamped here)
    __device__ __forceinline__ static bool update(const NsConst& c, const EulerArgs& ea, const double (&cc)[4],
                                                  double (&d1)[12], double (&d2)[9], double rh,
                                                  double Tc, double su, double sv, double (&out)[4]) {
        clamp_group<HI_100>(d1);
        clamp_group<HI_1000>(d2);
        if (rh > 1e-10) {
            const RhoDiv over_rho(rh);
            const double nu = min_one(over_rho(c.mu));  // fmin(mu / fmax(rho, 1e-10), 1): rho > 1e-10 here
            double fu = su, fv = sv, fw = 0.0;
            if (c.beta != 0.0) {
                const double dT = Tc - c.T_ref;
                fu += -c.beta * dT * c.gx;
                fv += -c.beta * dT * c.gy;
                fw += -c.beta * dT * c.gz;
            }
            const double src[3] = {fu, fv, fw};
#pragma unroll
            for (int q = 0; q < 3; q++) {
                const double inc = ea.dt * (-cc[0] * d1[3 * q] - cc[1] * d1[3 * q + 1] - cc[2] * d1[3 * q + 2] - over_rho(d1[9 + q]) +
                                            nu * (d2[3 * q] + d2[3 * q + 1] + d2[3 * q + 2]) + src[q]);
                out[q] = clamp100(cc[q] + clampb<HI_1>(inc));
            }
            const double div = clampb<HI_10>(d1[0] + d1[4] + d1[8]);
            out[3] = cc[3] + clampb<HI_1>(ea.p_factor * rh * div);
        } else {
#pragma unroll
            for (int q = 0; q < 4; q++) out[q] = cc[q];
        }
        return (abs_hi(out[0]) >= HI_INF) | (abs_hi(out[1]) >= HI_INF) | (abs_hi(out[2]) >= HI_INF) |
               (abs_hi(out[3]) >= HI_INF);
    }
};

template <bool D3>
__global__ void __launch_bounds__(TX* TY)
k_euler(Geom g, NsConst c, EulerArgs ea, int zchunk, P4 qin, W4 qout, const double* __restrict__ rho,
        const double* __restrict__ T, const double* __restrict__ su_tab, const double* __restrict__ sv_tab, Ctl* ctl) {
    const int i = blockIdx.x * TX + threadIdx.x, j = blockIdx.y * TY + threadIdx.y;
    const int kb = g.k0 + blockIdx.z * zchunk, ke = min(kb + zchunk, g.k1);
    if (!(i >= 1 && i < g.nx - 1 && j >= 1 && j < g.ny - 1)) return;
    long long n = (long long)kb * g.sz + (long long)j * g.sy + i;
    const double su = c.src_on ? su_tab[j] * ea.src_decay : 0.0;
    const double sv = c.src_on ? sv_tab[i] * ea.src_decay : 0.0;
    March2 F[4];
#pragma unroll
    for (int q = 0; q < 4; q++) {
        F[q].c = qin.a[q][n];
        F[q].m = D3 ? qin.a[q][n - g.sz] : F[q].c;  // nz = 1: stride_z = 0, the z neighbours are the cell itself
        F[q].p = F[q].c;
    }
    bool bad = false;
    for (int k = kb; k < ke; k++, n += g.sz) {
        double d1[12], d2[9], cc[4], out[4];
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const double* f = qin.a[q];
            if (D3) F[q].p = f[n + g.sz];
            const double e = f[n + 1], w_ = f[n - 1], nn = f[n + g.sy], s_ = f[n - g.sy];
            cc[q] = F[q].c;
            d1[3 * q + 0] = ediv(e - w_, g.two_dx, g.inv_2dx);
            d1[3 * q + 1] = ediv(nn - s_, g.two_dy, g.inv_2dy);
            d1[3 * q + 2] = (F[q].p - F[q].m) * g.inv_2dz;
            if (q < 3) {
                d2[3 * q + 0] = ediv(e - 2.0 * cc[q] + w_, g.dx2, g.inv_dx2);
                d2[3 * q + 1] = ediv(nn - 2.0 * cc[q] + s_, g.dy2, g.inv_dy2);
                d2[3 * q + 2] = (F[q].p - 2.0 * cc[q] + F[q].m) * g.inv_dz2;
            }
        }
        bad |= EulerCell::update(c, ea, cc, d1, d2, __ldg(rho + n), (c.beta != 0.0) ? __ldg(T + n) : 0.0, su, sv, out);
#pragma unroll
        for (int q = 0; q < 4; q++) {
            qout.a[q][n] = out[q];
            F[q].m = F[q].c;
            F[q].c = F[q].p;
        }
    }
    if (bad) ctl->nonfinite = 1;
}

struct EulerOp {
    static constexpr bool PRODUCER_WARP = false, STAGED = false;  // 4 stencil fields: the 512-thread pipeline
    static constexpr int NF = 4, NPW = 2;                         // pointwise: rho, T
    const Geom& g;
    const NsConst& c;
    const EulerArgs& ea;
    const W4& qout;
    const double* __restrict__ rho;
    const double* __restrict__ T;
    const double* __restrict__ su_tab;
    const double* __restrict__ sv_tab;
    bool bad;
    __device__ __forceinline__ void pw_load(long long n, double (&pw)[2]) const {
        pw[0] = __ldg(rho + n);
        pw[1] = (c.beta != 0.0) ? __ldg(T + n) : 0.0;
    }
    __device__ __forceinline__ void cell(const St<4>& s, const double (&pw)[2], long long n, int i, int j, int) {
        double d1[12], d2[9], cc[4], out[4];
#pragma unroll
        for (int q = 0; q < 4; q++) {
            cc[q] = s.c(q);
            d1[3 * q + 0] = ediv(s.e(q) - s.w(q), g.two_dx, g.inv_2dx);
            d1[3 * q + 1] = ediv(s.n(q) - s.s(q), g.two_dy, g.inv_2dy);
            d1[3 * q + 2] = (s.p(q) - s.m(q)) * g.inv_2dz;
            if (q < 3) {
                d2[3 * q + 0] = ediv(s.e(q) - 2.0 * cc[q] + s.w(q), g.dx2, g.inv_dx2);
                d2[3 * q + 1] = ediv(s.n(q) - 2.0 * cc[q] + s.s(q), g.dy2, g.inv_dy2);
                d2[3 * q + 2] = (s.p(q) - 2.0 * cc[q] + s.m(q)) * g.inv_dz2;
            }
        }
        const double su = c.src_on ? su_tab[j] * ea.src_decay : 0.0, sv = c.src_on ? sv_tab[i] * ea.src_decay : 0.0;
        bad |= EulerCell::update(c, ea, cc, d1, d2, pw[0], pw[1], su, sv, out);
#pragma unroll
        for (int q = 0; q < 4; q++) qout.a[q][n] = out[q];
    }
};

// The halo of the TMA boxes delivers the boundary cells themselves -- exactly the plain neighbours
// the reference reads (no periodic pre-pass, unlike the RK stage).
template <bool D3>
__global__ void __launch_bounds__(Tile::THREADS, 1)
k_euler_tma(const __grid_constant__ TmaSet<4> tm, Geom g, NsConst c, EulerArgs ea, int zchunk, W4 qout,
            const double* __restrict__ rho, const double* __restrict__ T, const double* __restrict__ su_tab,
            const double* __restrict__ sv_tab, Ctl* ctl) {
    EulerOp op{g, c, ea, qout, rho, T, su_tab, sv_tab, false};
    tma_direct_sweep<D3>(tm, g, zchunk, op);
    if (op.bad) ctl->nonfinite = 1;
}

// Periodic images on the boundary cells of up to 6 fields, one thread per face cell (closed form of
// the sequential x, y, z wrap of solver_explicit_euler.c:231-314: boundary = wrapped interior).
struct F6 {
    double* a[6];
    int n;
};
__global__ void k_periodic_faces(Geom g, F6 f, int wrap_y, int wrap_z) {
    const long long n = face_cell(g, (long long)blockIdx.x * blockDim.x + threadIdx.x);
    if (n < 0) return;
    const long long nxy = (long long)g.nx * g.ny;
    const int i = (int)(n % g.nx), j = (int)((n / g.nx) % g.ny), k = (int)(n / nxy);
    int ci = i, cj = j, ck = k;
    if (i == 0) ci = g.nx - 2;
    if (i == g.nx - 1) ci = 1;
    if (wrap_y) {
        if (j == 0) cj = g.ny - 2;
        if (j == g.ny - 1) cj = 1;
    }
    if (g.dim3 && wrap_z) {
        if (k == 0) ck = g.nz - 2;
        if (k == g.nz - 1) ck = 1;
    }
    if (ci == i && cj == j && ck == k) return;
    const long long src = (long long)ck * nxy + (long long)cj * g.nx + ci;
    for (int q = 0; q < f.n; q++) f.a[q][n] = f.a[q][src];
}

// Periodic wrap of one field on every physical boundary (solver_explicit_euler.c:231-314):
// x faces, then y faces, then z faces  ==>  closed form: boundary cell = wrapped interior cell.
__global__ void k_periodic_bc(Geom g, double* f, int wrap_y, int wrap_z) {
    long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= g.ncell) return;
    const int i = (int)(t % g.nx), j = (int)((t / g.nx) % g.ny), k = (int)(t / ((long long)g.nx * g.ny));
    int ci = i, cj = j, ck = k;
    if (i == 0) ci = g.nx - 2;
    if (i == g.nx - 1) ci = 1;
    if (wrap_y) {
        if (j == 0) cj = g.ny - 2;
        if (j == g.ny - 1) cj = 1;
    }
    if (g.dim3 && wrap_z) {
        if (k == 0) ck = g.nz - 2;
        if (k == g.nz - 1) ck = 1;
    }
    if (ci == i && cj == j && ck == k) return;
    f[t] = f[(long long)ck * g.nx * g.ny + (long long)cj * g.nx + ci];
}

// Thermal BCs, one launch per axis, in the reference's order left/right, bottom/top, back/front
// (energy_solver.c:246-331): later faces read cells already finalised by earlier ones.
__global__ void k_thermal_bc(Geom g, double* T, int axis, int type_lo, int type_hi, double val_lo,
                             double val_hi, int do_lo, int do_hi) {
    // face extents: axis 0 -> (j,k), axis 1 -> (i,k), axis 2 -> (i,j)
    const long long na = (axis == 0) ? g.ny : g.nx;
    const long long nb = (axis == 2) ? g.ny : g.nz;
    long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= na * nb) return;
    const long long a = t % na, b = t / na;
    const long long plane = (long long)g.nx * g.ny;
    const int n_ax = (axis == 0) ? g.nx : (axis == 1 ? g.ny : g.nz);
    const long long st = (axis == 0) ? 1 : (axis == 1 ? g.nx : plane);
    long long base;
    if (axis == 0) base = b * plane + a * g.nx;
    else if (axis == 1) base = b * plane + a;
    else base = b * g.nx + a;
    if (do_lo) {
        double* c = T + base;
        if (type_lo == BC_TYPE_DIRICHLET) *c = val_lo;
        else if (type_lo == BC_TYPE_NEUMANN) *c = c[st];
        else *c = c[(long long)(n_ax - 2) * st];
    }
    if (do_hi) {
        double* c = T + base + (long long)(n_ax - 1) * st;
        if (type_hi == BC_TYPE_DIRICHLET) *c = val_hi;
        else if (type_hi == BC_TYPE_NEUMANN) *c = c[-st];
        else *c = T[base + st];
    }
}

// max sqrt(u^2+v^2+w^2), max |p|, max T over ALL cells (solver_registry.c:31-62).
__global__ void __launch_bounds__(256)
k_field_stats(Geom g, const double* __restrict__ u, const double* __restrict__ v,
              const double* __restrict__ w, const double* __restrict__ p, const double* __restrict__ T,
              Ctl* ctl, RedBuf rb) {
    double m[3] = {0.0, 0.0, -INFINITY};
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < g.ncell; t += stride) {
        const double a = u[t], b = v[t], c = w[t];
        m[0] = fmax(m[0], sqrt(a * a + b * b + c * c));
        m[1] = fmax(m[1], fabs(p[t]));
        if (T) m[2] = fmax(m[2], T[t]);
    }
    block_reduce_finish<3, true>(m, rb, &ctl->stat[0]);
}

__device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
// Synthetic state of SURVEY 8(d) C5: u = 0.1 sin2pix cos2piy cos2piz, v = -0.1 cos2pix sin2piy cos2piz,
// w = 0, each + 1e-3 * splitmix64(seed, global idx) in [-1,1); p = 0, rho = 1, T = 300.
__global__ void k_synth_init(Geom g, uint64_t seed, long long gk0, long long gnz, double* u, double* v,
                             double* w, double* p, double* T, double* rho) {
    long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= g.ncell) return;
    const int i = (int)(t % g.nx), j = (int)((t / g.nx) % g.ny);
    const long long k = t / ((long long)g.nx * g.ny) + gk0;
    const double tp = 6.283185307179586;
    const double x = (double)i / (g.nx - 1), y = (double)j / (g.ny - 1);
    const double z = gnz > 1 ? (double)k / (double)(gnz - 1) : 0.0;
    const uint64_t gid = (uint64_t)((k * g.ny + j) * (long long)g.nx + i);
    auto rnd = [&](uint64_t c) {
        const uint64_t h = splitmix64(seed ^ (gid * 3ull + c));
        return ((double)(h >> 11) * (1.0 / 9007199254740992.0)) * 2.0 - 1.0;
    };
    const double cz = g.dim3 ? cos(tp * z) : 1.0;
    u[t] = 0.1 * sin(tp * x) * cos(tp * y) * cz + 1e-3 * rnd(0);
    v[t] = -0.1 * cos(tp * x) * sin(tp * y) * cz + 1e-3 * rnd(1);
    w[t] = g.dim3 ? 1e-3 * rnd(2) : 0.0;
    p[t] = 0.0;
    if (T) T[t] = 300.0;
    if (rho) rho[t] = 1.0;
}

// (4 loads before 4 stores, 32 blocks per SM: 6.35 against 5.8 TB/s for the plain grid-stride loop on 8 GB arrays,
// tools/copy_probe.cu)
__global__ void k_copy(double* __restrict__ dst, const double* __restrict__ src, size_t n) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; t + 3 * stride < n; t += 4 * stride) {
        double v[4];
#pragma unroll
        for (int u = 0; u < 4; u++) v[u] = src[t + u * stride];
#pragma unroll
        for (int u = 0; u < 4; u++) dst[t + u * stride] = v[u];
    }
    for (; t < n; t += stride) dst[t] = src[t];
}

__global__ void k_fill(double* __restrict__ dst, double v, size_t n) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += stride) dst[t] = v;
}

__global__ void k_zero_ctl(Ctl* ctl) {
    const int t = threadIdx.x;
    if (t < 32) ctl->dot[t] = 0.0;
    if (t == 0) {
        ctl->tol = ctl->r0 = ctl->res = 0.0;
        ctl->done_seq = INT_MAX;
        ctl->status = POISSON_ERROR;
        ctl->iterations = 0;
        ctl->result_buf = 0;
        ctl->skip_final_bc = 0;
    }
}

#include "ns_launch2.inc"

CFDB_ARITH_END
}  // namespace cfdb
