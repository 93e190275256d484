// poisson_mg.cu -- geometric multigrid for the pressure Poisson equation (POISSON_METHOD_MULTIGRID).
//
// The reference only reserves the name (lib/include/cfd/solvers/poisson_solver.h:60, "future";
// ROADMAP.md:116,194 "optional multigrid preconditioner for faster convergence"): its projection step
// spends 70-80 % of its time in hundreds to thousands of CG / Jacobi iterations per step
// (docs/reference/solvers.md:372).  This is SURVEY 8(f) n4, second half.  There is no reference
// output to match; the algorithm is stated a second time on the CPU in oracle/oracle_mg.inc and the
// two agree bit for bit (every kernel below performs the same IEEE operations in the same order, and
// the only reduction is a maximum).
//
// Problem: the reference's 7-point Laplacian with mirrored boundary cells (linear_solver.c:304-392),
// lap(x) = rhs on the interior cells 1..n-2 of each axis -- a cell-centred finite-volume Neumann
// problem.  Hierarchy per axis: an even number of unknowns is halved by merging pairs of cells ("CC"),
// an odd one by keeping every second point ("VC"); in the second case the wall is no longer half a
// spacing away from the first point, so every level carries the control-volume width vL / vR of its
// first / last point (in units of its spacing) and the boundary rows are scaled by 1 / v.  Restriction =
// volume-weighted average over the coarse control volume, prolongation = linear, mirrored at the walls.
// Smoother: Red-Black Gauss-Seidel, V(2,2); 30 sweeps on the coarsest level (one block).  Neighbours
// are read with clamped indices, so no boundary pass is needed inside a cycle; the boundary cells of the
// solution are filled once at the end, as the other methods leave them.
//
// Single GPU only (a slab-decomposed hierarchy would need its own halo plan per level).
#include "session.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

namespace cfdb {

namespace {

constexpr int MG_MAX_LEVELS = 32, MG_NU1 = 2, MG_NU2 = 2, MG_NCOARSE = 30;
constexpr long long MG_ONE_BLOCK_CELLS = 4096;  // coarsest level up to this many cells: all sweeps in one block
enum { MG_NONE = 0, MG_CC = 1, MG_VC = 2 };

struct MgLv {
    int n[3], m[3], lo[3], hi[3], kind[3];
    double ih2[3], vL[3], vR[3], wL[3], wR[3];  // w = 1 / v (IEEE, computed once on the host)
    long long sy, sz, total;
    double *x, *b, *r;
    double* x2;  // ping-pong partner of x for the one-pass sweep (may be null: two-pass smoother)
};

__host__ __device__ inline double mg_vol(const MgLv& L, int d, int i) {
    if (i == 1) return L.vL[d];
    if (i == L.m[d]) return L.vR[d];
    return 1.0;
}

}  // namespace

struct MgState {
    int nl = 0;
    MgLv lv[MG_MAX_LEVELS];
    double h[MG_MAX_LEVELS][3], aL[MG_MAX_LEVELS][3], aR[MG_MAX_LEVELS][3];
    std::vector<double*> owned;
    // one V-cycle + the residual check, captured once as a CUDA graph (the cycle of a small grid is ~60 launches of
    // a few microseconds each: launch-bound); valid for the x / rhs pointers it was captured with
    cudaGraphExec_t exec = nullptr;
    const double *gx = nullptr, *gb = nullptr;
    bool graph_failed = false;
    unsigned long long graph_launches = 0, graph_kind[KK_COUNT] = {};
};

namespace {

#ifdef __CUDACC__

// c[d] = (1 / v_d(i)) * ih2[d]: 1 / 1 = 1 and 1 * ih2 = ih2 exactly, so interior cells pay no division
__device__ __forceinline__ void mg_coef(const MgLv& L, int i, int j, int k, double (&c)[3]) {
    const int ix[3] = {i, j, k};
#pragma unroll
    for (int d = 0; d < 3; d++) {
        const double w = (ix[d] == 1) ? L.wL[d] : ((ix[d] == L.m[d]) ? L.wR[d] : 1.0);
        c[d] = (L.n[d] > 1) ? w * L.ih2[d] : 0.0;
    }
}

__device__ __forceinline__ double mg_nbsum(const MgLv& L, const double* x, long long n, int i, int j,
                                           int k, const double (&c)[3]) {
    const double xe = x[(i == L.hi[0]) ? n : n + 1], xw = x[(i == L.lo[0]) ? n : n - 1];
    const double xn = x[(j == L.hi[1]) ? n : n + L.sy], xs = x[(j == L.lo[1]) ? n : n - L.sy];
    const double xu = x[(k == L.hi[2]) ? n : n + L.sz], xd = x[(k == L.lo[2]) ? n : n - L.sz];
    return (c[0] * (xe + xw) + c[1] * (xn + xs)) + c[2] * (xu + xd);
}

__device__ __forceinline__ void mg_relax_cell(const MgLv& L, double* x, const double* __restrict__ b,
                                              int i, int j, int k) {
    const long long n = (long long)k * L.sz + (long long)j * L.sy + i;
    double c[3];
    mg_coef(L, i, j, k, c);
    const double s = mg_nbsum(L, x, n, i, j, k, c);
    const double diag = 2.0 * ((c[0] + c[1]) + c[2]);
    x[n] = (s - b[n]) / diag;
}

// One colour of a Red-Black Gauss-Seidel sweep: a thread owns one cell of the colour.
__global__ void __launch_bounds__(256) k_mg_smooth(MgLv L, double* x, const double* __restrict__ b, int colour) {
    const int j = L.lo[1] + blockIdx.y * blockDim.y + threadIdx.y;
    const int k = L.lo[2] + blockIdx.z;
    if (j > L.hi[1]) return;
    const int i0 = L.lo[0] + (((L.lo[0] + j + k) & 1) != colour);  // first cell of the colour in this row
    const int i = i0 + 2 * (blockIdx.x * blockDim.x + threadIdx.x);
    if (i > L.hi[0]) return;
    mg_relax_cell(L, x, b, i, j, k);
}

// One whole Red-Black Gauss-Seidel sweep in ONE pass, out of place (xin -> xout).  The two-pass smoother above
// moves every sector of x and rhs once per colour; here a block loads its TX x TY x TZ tile plus a 2-cell frame
// into shared memory (indices clamped at the walls = the mirrored neighbours), relaxes the red cells of the tile
// plus a 1-cell frame in place, then the black cells of the tile, and writes the tile.  The frame's red cells are
// recomputed by the neighbouring blocks from the same inputs with the same operations, so the result is
// bit-identical to k_mg_smooth(colour 0) followed by k_mg_smooth(colour 1).
template <int TX, int TY, int TZ, bool D3>
__global__ void __launch_bounds__(256) k_mg_sweep2(MgLv L, const double* __restrict__ xin, double* __restrict__ xout,
                                                   const double* __restrict__ b) {
    constexpr int HZ = D3 ? 2 : 0;
    constexpr int SX = TX + 4, SY = TY + 4, SZ = TZ + 2 * HZ;
    __shared__ double sm[SZ][SY][SX];
    const int tid = threadIdx.x;
    const int i0 = L.lo[0] + blockIdx.x * TX, j0 = L.lo[1] + blockIdx.y * TY, k0 = D3 ? L.lo[2] + blockIdx.z * TZ : 0;
    for (int t = tid; t < SX * SY * SZ; t += 256) {
        const int a = t % SX, bb = (t / SX) % SY, c = t / (SX * SY);
        const int gi = min(max(i0 - 2 + a, L.lo[0]), L.hi[0]), gj = min(max(j0 - 2 + bb, L.lo[1]), L.hi[1]);
        const int gk = D3 ? min(max(k0 - 2 + c, L.lo[2]), L.hi[2]) : 0;
        sm[c][bb][a] = xin[(long long)gk * L.sz + (long long)gj * L.sy + gi];
    }
    __syncthreads();
    // phase 0: red cells of the tile + 1-cell frame; phase 1: black cells of the tile
#pragma unroll
    for (int ph = 0; ph < 2; ph++) {
        const int F = 1 - ph;  // frame width of this phase
        const int EX = TX + 2 * F, EY = TY + 2 * F, EZ = D3 ? TZ + 2 * F : 1;
        const int HXR = (EX + 1) / 2;
        for (int t = tid; t < HXR * EY * EZ; t += 256) {
            const int u = t % HXR, bb = 2 - F + (t / HXR) % EY, c = D3 ? 2 - F + t / (HXR * EY) : 0;
            const int gj = j0 - 2 + bb, gk = D3 ? k0 - 2 + c : 0;
            const int a0 = 2 - F;                                     // first column of the region
            const int par = ((i0 - 2 + a0 + gj + gk) & 1) != ph;      // skip one column if the first has the other colour
            const int a = a0 + 2 * u + par;
            const int gi = i0 - 2 + a;
            if (a >= a0 + EX || gi < L.lo[0] || gi > L.hi[0] || gj < L.lo[1] || gj > L.hi[1] ||
                (D3 && (gk < L.lo[2] || gk > L.hi[2])))
                continue;
            double cf[3];
            mg_coef(L, gi, gj, gk, cf);
            const double xc = sm[c][bb][a];
            const double xu = D3 ? sm[c + 1][bb][a] : xc, xd = D3 ? sm[c - 1][bb][a] : xc;
            const double sv = (cf[0] * (sm[c][bb][a + 1] + sm[c][bb][a - 1]) + cf[1] * (sm[c][bb + 1][a] + sm[c][bb - 1][a])) +
                              cf[2] * (xu + xd);
            const double diag = 2.0 * ((cf[0] + cf[1]) + cf[2]);
            sm[c][bb][a] = (sv - b[(long long)gk * L.sz + (long long)gj * L.sy + gi]) / diag;
        }
        __syncthreads();
    }
    for (int t = tid; t < TX * TY * TZ; t += 256) {
        const int a = t % TX, bb = (t / TX) % TY, c = t / (TX * TY);
        const int gi = i0 + a, gj = j0 + bb, gk = D3 ? k0 + c : 0;
        if (gi <= L.hi[0] && gj <= L.hi[1] && (!D3 || gk <= L.hi[2]))
            xout[(long long)gk * L.sz + (long long)gj * L.sy + gi] = sm[c + HZ][bb + 2][a + 2];
    }
}

// All sweeps of the coarsest level in one block (colours separated by block barriers).
__global__ void __launch_bounds__(256) k_mg_coarse(MgLv L, double* x, const double* __restrict__ b, int sweeps) {
    const int m0 = L.hi[0] - L.lo[0] + 1, m1 = L.hi[1] - L.lo[1] + 1, m2 = L.hi[2] - L.lo[2] + 1;
    const int cells = m0 * m1 * m2;
    for (int q = 0; q < 2 * sweeps; q++) {
        const int colour = q & 1;
        for (int c = threadIdx.x; c < cells; c += blockDim.x) {
            const int i = L.lo[0] + c % m0, j = L.lo[1] + (c / m0) % m1, k = L.lo[2] + c / (m0 * m1);
            if (((i + j + k) & 1) == colour) mg_relax_cell(L, x, b, i, j, k);
        }
        __syncthreads();
    }
}

// The same with the level held in shared memory (whole array, ghost layout included, so the indices are unchanged):
// a phase of the global-memory version pays an L2 round trip on top of the division's dependent chain (60 phases =
// 37 us per cycle); from shared memory the 60 phases take 29 us (profiles/r1x_mg_compare.jsonl) -- what is left is
// the barrier, the rhs load and the FP64 division latency of each phase.  Same operations, same order of the colours.
constexpr int MG_SMEM_DOUBLES = 4096;
__global__ void __launch_bounds__(256) k_mg_coarse_smem(MgLv L, double* x, const double* __restrict__ b, int sweeps) {
    __shared__ double xs[MG_SMEM_DOUBLES];
    const int total = (int)L.total;
    for (int t = threadIdx.x; t < total; t += blockDim.x) xs[t] = x[t];
    __syncthreads();
    const int m0 = L.hi[0] - L.lo[0] + 1, m1 = L.hi[1] - L.lo[1] + 1, m2 = L.hi[2] - L.lo[2] + 1;
    const int cells = m0 * m1 * m2;
    for (int q = 0; q < 2 * sweeps; q++) {
        const int colour = q & 1;
        for (int c = threadIdx.x; c < cells; c += blockDim.x) {
            const int i = L.lo[0] + c % m0, j = L.lo[1] + (c / m0) % m1, k = L.lo[2] + c / (m0 * m1);
            if (((i + j + k) & 1) == colour) mg_relax_cell(L, xs, b, i, j, k);
        }
        __syncthreads();
    }
    for (int t = threadIdx.x; t < total; t += blockDim.x) x[t] = xs[t];
}

// r = b - A x (r may be null); the maximum of |r| goes to *slot (bits of a non-negative double order like integers).
__global__ void __launch_bounds__(256) k_mg_residual(MgLv L, const double* __restrict__ x, const double* __restrict__ b,
                                                     double* __restrict__ r, unsigned long long* slot) {
    const int i = L.lo[0] + blockIdx.x * blockDim.x + threadIdx.x;
    const int j = L.lo[1] + blockIdx.y * blockDim.y + threadIdx.y;
    const int k = L.lo[2] + blockIdx.z;
    double av = 0.0;
    if (i <= L.hi[0] && j <= L.hi[1]) {
        const long long n = (long long)k * L.sz + (long long)j * L.sy + i;
        double c[3];
        mg_coef(L, i, j, k, c);
        const double s = mg_nbsum(L, x, n, i, j, k, c);
        const double diag = 2.0 * ((c[0] + c[1]) + c[2]);
        const double rv = b[n] - (s - diag * x[n]);
        if (r) r[n] = rv;
        const double a = fabs(rv);
        if (a > av) av = a;  // a NaN residual is ignored, as in the CPU statement
    }
    if (!slot) return;
    av = warp_max(av);
    __shared__ double wm[8];
    const int t = threadIdx.y * blockDim.x + threadIdx.x;
    if ((t & 31) == 0) wm[t >> 5] = av;
    __syncthreads();
    if (t == 0) {
        double m = wm[0];
        for (int q = 1; q < (int)(blockDim.x * blockDim.y + 31) / 32; q++) m = fmax(m, wm[q]);
        if (m > 0.0) atomicMax(slot, (unsigned long long)__double_as_longlong(m));
    }
}

constexpr int MG_ZC = 8;  // planes marched per thread by the 3-D prolongation

struct Taps3 {
    int n, f[3];
    double w[3], den;
};
__device__ __forceinline__ Taps3 mg_restrict_taps(const MgLv& F, const MgLv& C, int d, int I) {
    Taps3 t;
    if (F.kind[d] == MG_CC) {
        t.n = 2;
        t.f[0] = 2 * I - 1;
        t.f[1] = 2 * I;
        t.f[2] = 0;
        t.w[0] = mg_vol(F, d, t.f[0]);
        t.w[1] = mg_vol(F, d, t.f[1]);
        t.w[2] = 0.0;
        t.den = 2.0 * mg_vol(C, d, I);
    } else if (F.kind[d] == MG_VC) {
        t.n = 3;
        t.f[0] = 2 * I - 1;
        t.f[1] = 2 * I;
        t.f[2] = 2 * I + 1;
        t.w[0] = (I == 1) ? F.vL[d] : 0.5;
        t.w[1] = 1.0;
        t.w[2] = (I == C.m[d]) ? F.vR[d] : 0.5;
        t.den = 2.0 * mg_vol(C, d, I);
    } else {
        t.n = 1;
        t.f[0] = I;
        t.f[1] = t.f[2] = 0;
        t.w[0] = 1.0;
        t.w[1] = t.w[2] = 0.0;
        t.den = 1.0;
    }
    return t;
}

// b_C = R r_F, x_C = 0: a thread owns one coarse cell
__global__ void __launch_bounds__(256) k_mg_restrict(MgLv F, MgLv C) {
    const int I = C.lo[0] + blockIdx.x * blockDim.x + threadIdx.x;
    const int J = C.lo[1] + blockIdx.y * blockDim.y + threadIdx.y;
    const int K = C.lo[2] + blockIdx.z;
    if (I > C.hi[0] || J > C.hi[1]) return;
    const Taps3 tx = mg_restrict_taps(F, C, 0, I), ty = mg_restrict_taps(F, C, 1, J), tz = mg_restrict_taps(F, C, 2, K);
    double az = 0.0;  // (unrolled with guards: the tap arrays stay in registers)
#pragma unroll
    for (int c = 0; c < 3; c++) {
        if (c >= tz.n) continue;
        double ay = 0.0;
#pragma unroll
        for (int b = 0; b < 3; b++) {
            if (b >= ty.n) continue;
            double ax = 0.0;
            const double* row = F.r + (long long)tz.f[c] * F.sz + (long long)ty.f[b] * F.sy;
#pragma unroll
            for (int a = 0; a < 3; a++)
                if (a < tx.n) ax += tx.w[a] * row[tx.f[a]];
            ay += ty.w[b] * ax;
        }
        az += tz.w[c] * ay;
    }
    const long long n = (long long)K * C.sz + (long long)J * C.sy + I;
    C.b[n] = az / ((tx.den * ty.den) * tz.den);
    C.x[n] = 0.0;
}

struct Taps2 {
    int I[2];
    double w[2];
};
__device__ __forceinline__ Taps2 mg_prolong_taps(const MgLv& F, const MgLv& C, int d, int i) {
    Taps2 t;
    if (F.kind[d] == MG_CC) {
        if (i & 1) {
            t.I[0] = (i + 1) / 2;
            t.I[1] = t.I[0] - 1;
        } else {
            t.I[0] = i / 2;
            t.I[1] = t.I[0] + 1;
        }
        t.w[0] = 0.75;
        t.w[1] = 0.25;
    } else if (F.kind[d] == MG_VC) {
        if (i & 1) {
            t.I[0] = (i - 1) / 2;
            t.I[1] = (i + 1) / 2;
            t.w[0] = t.w[1] = 0.5;
        } else {
            t.I[0] = t.I[1] = i / 2;
            t.w[0] = 1.0;
            t.w[1] = 0.0;
        }
    } else {
        t.I[0] = t.I[1] = i;
        t.w[0] = 1.0;
        t.w[1] = 0.0;
        return t;
    }
#pragma unroll
    for (int q = 0; q < 2; q++) {
        if (t.I[q] < 1) t.I[q] = 1;
        if (t.I[q] > C.m[d]) t.I[q] = C.m[d];
    }
    return t;
}

// x_F += P x_C: a thread owns one fine cell
__global__ void __launch_bounds__(256) k_mg_prolong_add(MgLv F, MgLv C) {
    const int i = F.lo[0] + blockIdx.x * blockDim.x + threadIdx.x;
    const int j = F.lo[1] + blockIdx.y * blockDim.y + threadIdx.y;
    const int k = F.lo[2] + blockIdx.z;
    if (i > F.hi[0] || j > F.hi[1]) return;
    const Taps2 tx = mg_prolong_taps(F, C, 0, i), ty = mg_prolong_taps(F, C, 1, j), tz = mg_prolong_taps(F, C, 2, k);
    double vz[2];
#pragma unroll
    for (int c = 0; c < 2; c++) {
        double vy[2];
#pragma unroll
        for (int b = 0; b < 2; b++) {
            const double* row = C.x + (long long)tz.I[c] * C.sz + (long long)ty.I[b] * C.sy;
            vy[b] = tx.w[0] * row[tx.I[0]] + tx.w[1] * row[tx.I[1]];
        }
        vz[c] = ty.w[0] * vy[0] + ty.w[1] * vy[1];
    }
    F.x[(long long)k * F.sz + (long long)j * F.sy + i] += tz.w[0] * vz[0] + tz.w[1] * vz[1];
}

// 3-D levels: the same prolongation with a thread marching MG_ZC fine planes; the in-plane taps are set up once
// and the in-plane interpolation of a coarse plane is re-used by the (up to) two or three fine planes that need it.
__global__ void __launch_bounds__(256) k_mg_prolong_add_march(MgLv F, MgLv C) {
    const int i = F.lo[0] + blockIdx.x * blockDim.x + threadIdx.x;
    const int j = F.lo[1] + blockIdx.y * blockDim.y + threadIdx.y;
    const int kb = F.lo[2] + blockIdx.z * MG_ZC, ke = min(kb + MG_ZC - 1, F.hi[2]);
    if (i > F.hi[0] || j > F.hi[1]) return;
    const Taps2 tx = mg_prolong_taps(F, C, 0, i), ty = mg_prolong_taps(F, C, 1, j);
    const double* r0 = C.x + (long long)ty.I[0] * C.sy;
    const double* r1 = C.x + (long long)ty.I[1] * C.sy;
    auto plane = [&](int K) -> double {  // in-plane interpolation of coarse plane K at (i, j)
        const double* p0 = r0 + (long long)K * C.sz;
        const double* p1 = r1 + (long long)K * C.sz;
        const double v0 = tx.w[0] * p0[tx.I[0]] + tx.w[1] * p0[tx.I[1]];
        const double v1 = tx.w[0] * p1[tx.I[0]] + tx.w[1] * p1[tx.I[1]];
        return ty.w[0] * v0 + ty.w[1] * v1;
    };
    int KA = -1, KB = -1;  // cached coarse planes
    double vA = 0.0, vB = 0.0;
    long long n = (long long)kb * F.sz + (long long)j * F.sy + i;
    for (int k = kb; k <= ke; k++, n += F.sz) {
        const Taps2 tz = mg_prolong_taps(F, C, 2, k);
        double v[2];
#pragma unroll
        for (int q = 0; q < 2; q++) {
            const int K = tz.I[q];
            if (K == KA) v[q] = vA;
            else if (K == KB) v[q] = vB;
            else {
                v[q] = plane(K);
                // replace the older entry (planes are visited in non-decreasing order)
                if (KA <= KB) { KA = K; vA = v[q]; } else { KB = K; vB = v[q]; }
            }
        }
        F.x[n] += tz.w[0] * v[0] + tz.w[1] * v[1];
    }
}

#endif  // __CUDACC__

inline dim3 cell_grid(const MgLv& L, int cols, dim3 blk) {
    return dim3((unsigned)((cols + (int)blk.x - 1) / (int)blk.x),
                (unsigned)((L.hi[1] - L.lo[1] + 1 + (int)blk.y - 1) / (int)blk.y), (unsigned)(L.hi[2] - L.lo[2] + 1));
}

void smooth(const LaunchCtx& Lc, const MgLv& L, int sweeps) {
    // (environment read per call: a fresh session re-captures its graph)
    // Measured on B200 (profiles/r1s_mg_compare.jsonl): the one-pass sweep wins on 2-D levels (1025^2: 3.1 ms of
    // smoothing per solve against 4.5 ms) and loses on 3-D ones (512^3: 68 ms against 52 ms -- the 36 x 12 x 12 tile
    // re-reads 2.5 cells per cell through L2 and its phases are latency-bound), so 3-D levels keep one kernel per
    // colour unless CFD_B200_MG_FUSED3D=1.
    const bool fused = !getenv("CFD_B200_NO_MG_FUSED") && (L.n[2] == 1 || getenv("CFD_B200_MG_FUSED3D"));
    if (L.x2 && sweeps % 2 == 0 && fused) {  // whole sweeps out of place, x -> x2 -> x
        const int m0 = L.hi[0] - L.lo[0] + 1, m1 = L.hi[1] - L.lo[1] + 1, m2 = L.hi[2] - L.lo[2] + 1;
        for (int q = 0; q < sweeps; q++) {
            const double* in = (q & 1) ? L.x2 : L.x;
            double* out = (q & 1) ? L.x : L.x2;
            if (L.n[2] > 1) {
                const dim3 grd((m0 + 31) / 32, (m1 + 7) / 8, (m2 + 7) / 8);
                CFDB_LAUNCH(KK_MG_SWEEP2, Lc.st, (k_mg_sweep2<32, 8, 8, true><<<grd, 256, 0, Lc.st>>>(L, in, out, L.b)));
            } else {
                const dim3 grd((m0 + 63) / 64, (m1 + 15) / 16, 1);
                CFDB_LAUNCH(KK_MG_SWEEP2, Lc.st, (k_mg_sweep2<64, 16, 1, false><<<grd, 256, 0, Lc.st>>>(L, in, out, L.b)));
            }
        }
        return;
    }
    const dim3 blk(32, 8);
    const int m0 = L.hi[0] - L.lo[0] + 1;
    const dim3 grd = cell_grid(L, (m0 + 1) / 2, blk);
    for (int q = 0; q < 2 * sweeps; q++)
        CFDB_LAUNCH(KK_MG_SMOOTH, Lc.st, k_mg_smooth<<<grd, blk, 0, Lc.st>>>(L, L.x, L.b, q & 1));
}

void residual(const LaunchCtx& Lc, const MgLv& L, double* r, unsigned long long* slot) {
    const dim3 blk(32, 8);
    // (a z-marching variant of this kernel -- 6 loads per cell, reduction once per 8 cells -- was measured 25 % SLOWER
    // at 512^3, profiles/r1u_mg_compare.jsonl: the march serialises the loads a one-cell thread has in flight at once)
    CFDB_LAUNCH(KK_MG_RESIDUAL, Lc.st,
                k_mg_residual<<<cell_grid(L, L.hi[0] - L.lo[0] + 1, blk), blk, 0, Lc.st>>>(L, L.x, L.b, r, slot));
}

void vcycle(const LaunchCtx& Lc, MgState* st, int l) {
    const MgLv& L = st->lv[l];
    if (l == st->nl - 1) {
        const long long cells = (long long)(L.hi[0] - L.lo[0] + 1) * (L.hi[1] - L.lo[1] + 1) * (L.hi[2] - L.lo[2] + 1);
        if (cells <= MG_ONE_BLOCK_CELLS && L.total <= MG_SMEM_DOUBLES && !getenv("CFD_B200_NO_MG_COARSE_SMEM"))
            CFDB_LAUNCH(KK_MG_COARSE, Lc.st, k_mg_coarse_smem<<<1, 256, 0, Lc.st>>>(L, L.x, L.b, MG_NCOARSE));
        else if (cells <= MG_ONE_BLOCK_CELLS)
            CFDB_LAUNCH(KK_MG_COARSE, Lc.st, k_mg_coarse<<<1, 256, 0, Lc.st>>>(L, L.x, L.b, MG_NCOARSE));
        else
            smooth(Lc, L, MG_NCOARSE);
        return;
    }
    const MgLv& C = st->lv[l + 1];
    const dim3 blk(32, 8);
    smooth(Lc, L, MG_NU1);
    residual(Lc, L, L.r, nullptr);
    CFDB_LAUNCH(KK_MG_RESTRICT, Lc.st,
                k_mg_restrict<<<cell_grid(C, C.hi[0] - C.lo[0] + 1, blk), blk, 0, Lc.st>>>(L, C));
    vcycle(Lc, st, l + 1);
    if (L.n[2] > 1 && !getenv("CFD_B200_NO_MG_MARCH")) {
        dim3 grd = cell_grid(L, L.hi[0] - L.lo[0] + 1, blk);
        grd.z = (grd.z + MG_ZC - 1) / MG_ZC;
        CFDB_LAUNCH(KK_MG_PROLONG, Lc.st, k_mg_prolong_add_march<<<grd, blk, 0, Lc.st>>>(L, C));
    } else {
        CFDB_LAUNCH(KK_MG_PROLONG, Lc.st,
                    k_mg_prolong_add<<<cell_grid(L, L.hi[0] - L.lo[0] + 1, blk), blk, 0, Lc.st>>>(L, C));
    }
    smooth(Lc, L, MG_NU2);
}

void level_init(MgState* st, int l, const int n[3], const double h[3], const double aL[3], const double aR[3]) {
    MgLv& L = st->lv[l];
    for (int d = 0; d < 3; d++) {
        L.n[d] = n[d];
        L.m[d] = n[d] > 1 ? n[d] - 2 : 1;
        L.lo[d] = n[d] > 1 ? 1 : 0;
        L.hi[d] = n[d] > 1 ? n[d] - 2 : 0;
        L.ih2[d] = (n[d] > 1 && h[d] > 0.0) ? 1.0 / (h[d] * h[d]) : 0.0;
        L.vL[d] = aL[d] + 0.5;
        L.vR[d] = aR[d] + 0.5;
        L.wL[d] = 1.0 / L.vL[d];
        L.wR[d] = 1.0 / L.vR[d];
        L.kind[d] = MG_NONE;
        st->h[l][d] = h[d];
        st->aL[l][d] = aL[d];
        st->aR[l][d] = aR[d];
    }
    L.sy = n[0];
    L.sz = (long long)n[0] * n[1];
    L.total = L.sz * n[2];
    L.x = L.b = L.r = L.x2 = nullptr;
}

// how level l coarsens (same rule as oracle_mg.inc: every active axis must be able to halve)
bool plan_next(MgState* st, int l, int nn[3], double nh[3], double naL[3], double naR[3]) {
    MgLv& L = st->lv[l];
    bool any = false;
    for (int d = 0; d < 3; d++) {
        nn[d] = L.n[d];
        nh[d] = st->h[l][d];
        naL[d] = st->aL[l][d];
        naR[d] = st->aR[l][d];
        if (L.n[d] == 1) continue;
        const int m = L.m[d];
        if (m % 2 == 0 && m / 2 >= 2) {
            L.kind[d] = MG_CC;
            nn[d] = m / 2 + 2;
            naL[d] = (st->aL[l][d] + 0.5) * 0.5;
            naR[d] = (st->aR[l][d] + 0.5) * 0.5;
        } else if (m % 2 == 1 && (m - 1) / 2 >= 3) {
            L.kind[d] = MG_VC;
            nn[d] = (m - 1) / 2 + 2;
            naL[d] = (st->aL[l][d] + 1.0) * 0.5;
            naR[d] = (st->aR[l][d] + 1.0) * 0.5;
        } else {
            for (int e = 0; e < 3; e++) L.kind[e] = MG_NONE;
            return false;
        }
        nh[d] = st->h[l][d] * 2.0;
        any = true;
    }
    return any;
}

}  // namespace

void mg_release(cfd_b200_session* s) {
    if (!s->mg) return;
    if (s->mg->exec) cudaGraphExecDestroy(s->mg->exec);
    for (double* p : s->mg->owned) cudaFree(p);
    delete s->mg;
    s->mg = nullptr;
}

static cfd_status_t mg_prepare(cfd_b200_session* s) {
    if (s->mg) return CFD_SUCCESS;
    MgState* st = new MgState();
    s->mg = st;
    int n[3] = {s->g.nx, s->g.ny, s->g.nz};
    double h[3] = {s->desc.dx, s->desc.dy, s->g.nz > 1 ? s->desc.dz : 0.0};
    double aL[3] = {0.5, 0.5, 0.5}, aR[3] = {0.5, 0.5, 0.5};
    for (;;) {
        level_init(st, st->nl, n, h, aL, aR);
        MgLv& L = st->lv[st->nl];
        const int nbuf = st->nl == 0 ? 2 : 4;  // level 0 solves in the caller's x / rhs
        double* buf[4] = {nullptr, nullptr, nullptr, nullptr};
        for (int q = 0; q < nbuf; q++) {
            if (cudaMalloc(&buf[q], (size_t)L.total * sizeof(double)) != cudaSuccess) {
                cudaGetLastError();
                const int lvl = st->nl;
                mg_release(s);
                set_error(CFD_ERROR_NOMEM, "multigrid: out of device memory at level %d", lvl);
                return CFD_ERROR_NOMEM;
            }
            st->owned.push_back(buf[q]);
            s->bytes_allocated += (size_t)L.total * sizeof(double);
            cudaMemsetAsync(buf[q], 0, (size_t)L.total * sizeof(double), s->st);
        }
        L.r = buf[0];
        L.x2 = buf[1];
        L.x = buf[2];
        L.b = buf[3];
        st->nl++;
        int nn[3];
        double nh[3], naL[3], naR[3];
        if (st->nl == MG_MAX_LEVELS || !plan_next(st, st->nl - 1, nn, nh, naL, naR)) break;
        for (int d = 0; d < 3; d++) {
            n[d] = nn[d];
            h[d] = nh[d];
            aL[d] = naL[d];
            aR[d] = naR[d];
        }
    }
    return CFD_SUCCESS;
}

int mg_level_count(cfd_b200_session* s) { return s->mg ? s->mg->nl : 0; }

// V-cycles until the L-infinity residual passes the reference's stationary-solver rule
// (tol = max(rel * r0, abs), linear_solver.c:397-485); one host look at the residual per cycle.
cfd_status_t mg_solve(cfd_b200_session* s, const cfd_b200_poisson_config_t* cfg, double* x, const double* rhs,
                      poisson_solver_stats_t* out) {
    if (s->desc.nranks > 1) {
        set_error(CFD_ERROR_UNSUPPORTED, "multigrid runs on one GPU only (no slab-decomposed hierarchy yet)");
        return CFD_ERROR_UNSUPPORTED;
    }
    cfd_status_t rc = mg_prepare(s);
    if (rc != CFD_SUCCESS) return rc;
    MgState* st = s->mg;
    const LaunchCtx Lc = s->L();
    const poisson_solver_params_t& pp = cfg->params;
    st->lv[0].x = x;
    st->lv[0].b = const_cast<double*>(rhs);
    unsigned long long* slot = reinterpret_cast<unsigned long long*>(&s->ctl->dot[0]);
    double* h_res = &s->h_ctl[0].dot[0];  // pinned
    auto linf = [&]() -> double {
        cudaMemsetAsync(slot, 0, sizeof(double), s->st);
        residual(Lc, st->lv[0], nullptr, slot);
        cudaMemcpyAsync(h_res, slot, sizeof(double), cudaMemcpyDeviceToHost, s->st);
        cudaStreamSynchronize(s->st);
        return *h_res;
    };
    // capture (or re-use) the graph of one cycle; per-launch profiling needs the individual launches
    const bool want_graph = !(t_prof && t_prof->on) && !st->graph_failed && !getenv("CFD_B200_NO_MG_GRAPH");
    if (want_graph && (!st->exec || st->gx != x || st->gb != rhs)) {
        if (st->exec) cudaGraphExecDestroy(st->exec);
        st->exec = nullptr;
        const unsigned long long l0 = g_launches.load();
        unsigned long long k0[KK_COUNT];
        if (t_prof) memcpy(k0, t_prof->launches, sizeof(k0));
        cudaGraph_t graph = nullptr;
        bool ok = cudaStreamBeginCapture(s->st, cudaStreamCaptureModeThreadLocal) == cudaSuccess;
        if (ok) {
            vcycle(Lc, st, 0);
            cudaMemsetAsync(slot, 0, sizeof(double), s->st);
            residual(Lc, st->lv[0], nullptr, slot);
            cudaMemcpyAsync(h_res, slot, sizeof(double), cudaMemcpyDeviceToHost, s->st);
            ok = cudaStreamEndCapture(s->st, &graph) == cudaSuccess && graph;
        }
        if (ok) ok = cudaGraphInstantiate(&st->exec, graph, 0) == cudaSuccess;
        if (graph) cudaGraphDestroy(graph);
        // nothing ran during the capture: take its launches back and remember them per replay
        st->graph_launches = g_launches.load() - l0;
        g_launches -= st->graph_launches;
        if (t_prof)
            for (int k = 0; k < KK_COUNT; k++) {
                st->graph_kind[k] = t_prof->launches[k] - k0[k];
                t_prof->launches[k] = k0[k];
            }
        if (!ok) {
            cudaGetLastError();
            st->exec = nullptr;
            st->graph_failed = true;  // direct launches from now on
        }
        st->gx = x;
        st->gb = rhs;
    }
    const bool graph = want_graph && st->exec;
    auto cycle_and_residual = [&]() -> double {
        if (!graph) {
            vcycle(Lc, st, 0);
            return linf();
        }
        cudaGraphLaunch(st->exec, s->st);
        g_launches += st->graph_launches;
        if (t_prof)
            for (int k = 0; k < KK_COUNT; k++) t_prof->launches[k] += st->graph_kind[k];
        cudaStreamSynchronize(s->st);
        return *h_res;
    };
    poisson_solver_stats_t ps;
    memset(&ps, 0, sizeof(ps));
    const double r0 = linf();
    double tol = pp.tolerance * r0, res = r0;
    if (tol < pp.absolute_tolerance) tol = pp.absolute_tolerance;
    ps.initial_residual = r0;
    if (r0 < pp.absolute_tolerance) {
        ps.status = POISSON_CONVERGED;
        ps.iterations = 0;
        ps.final_residual = r0;
    } else {
        bool converged = false;
        int iter;
        for (iter = 0; iter < pp.max_iterations; iter++) {
            res = cycle_and_residual();
            if (res < tol || res < pp.absolute_tolerance) {
                converged = true;
                break;
            }
        }
        ps.iterations = converged ? iter + 1 : iter;
        ps.final_residual = res;
        ps.status = converged ? POISSON_CONVERGED : POISSON_MAX_ITER;
    }
    launch_neumann_fill(Lc, x, 0, 0, false);
    cudaError_t ce = cudaGetLastError();
    if (ce != cudaSuccess) {
        set_error(CFD_ERROR, "CUDA failure in multigrid solve: %s", cudaGetErrorString(ce));
        return CFD_ERROR;
    }
    s->last_ps = ps;
    if (out) *out = ps;
    return ps.status == POISSON_CONVERGED ? CFD_SUCCESS : CFD_ERROR_MAX_ITER;
}

}  // namespace cfdb
