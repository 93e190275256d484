// poisson_kernels.cu -- pressure-Poisson kernels: Jacobi, Red-Black SOR, CG/PCG, BiCGSTAB.
//
// Reference semantics followed (file:line under /root/reference/lib/src/solvers/linear):
//   linear_solver.c:304-346 (L-inf residual), :348-392 (Neumann BC), :397-485 (stationary driver)
//   cpu/linear_solver_jacobi.c:76-129, cpu/linear_solver_redblack.c:79-146,
//   cpu/linear_solver_cg.c:290-461, cpu/linear_solver_bicgstab.c:265-495.
//
// Design (B200): every kernel is a 2.5-D sweep -- a block owns an x-y tile and marches a chunk
// of z planes keeping the z-1/z/z+1 values of each stencil operand in registers; in-plane
// neighbours come through L1.  All solver scalars (dots, alpha/beta/omega, tolerances, the
// convergence decision) live in a device control block (Ctl); reductions are finished by the
// last block with a fixed summation order, so there is no host synchronisation and no
// floating-point atomics anywhere inside an iteration.
#include "common.cuh"
#include "kernels.h"
#include "stencil_tma.cuh"

#include <limits.h>
#include <type_traits>

namespace cfdb {
CFDB_ARITH_BEGIN

constexpr int TX = 32, TY = 8;  // thread tile (x fastest): 256 threads, 8 warps

__device__ __forceinline__ bool leader() {
    return (blockIdx.x | blockIdx.y | blockIdx.z | threadIdx.x | threadIdx.y) == 0;
}

// Column owned by this thread and the z chunk of this block.  Returns whether (i,j) is an
// interior column (the only ones that are updated).
__device__ __forceinline__ bool column(const Geom& g, int zchunk, int& i, int& j, int& kb, int& ke) {
    i = blockIdx.x * TX + threadIdx.x;
    j = blockIdx.y * TY + threadIdx.y;
    kb = g.k0 + blockIdx.z * zchunk;
    ke = min(kb + zchunk, g.k1);
    return i >= 1 && i < g.nx - 1 && j >= 1 && j < g.ny - 1;
}

// Store v at interior cell n=(i,j,k) and at every physical ghost cell whose nearest interior
// cell it is.  This is the closed form of the reference's sequential Neumann pass
// (linear_solver.c:367-391 + boundary_conditions_core_impl.h:46-69): ghost = clamped interior.
template <bool D3>
__device__ __forceinline__ void store_with_ghosts(const Geom& g, double* x, int i, int j, int k,
                                                  long long n, double v) {
    x[n] = v;
    const bool xl = (i == 1) && g.glo[0], xh = (i == g.nx - 2) && g.ghi[0];
    const bool yl = (j == 1) && g.glo[1], yh = (j == g.ny - 2) && g.ghi[1];
    const bool zl = D3 && (k == 1) && g.glo[2], zh = D3 && (k == g.nz - 2) && g.ghi[2];
    if (!(xl | xh | yl | yh | zl | zh)) return;
    // every combination of {0, low ghost, high ghost} per axis except (0,0,0); no local arrays
#pragma unroll
    for (int c = 0; c < 3; c++) {
        if ((c == 1 && !zl) || (c == 2 && !zh)) continue;
        const long long oz = (c == 0) ? 0 : (c == 1 ? -g.sz : g.sz);
#pragma unroll
        for (int b = 0; b < 3; b++) {
            if ((b == 1 && !yl) || (b == 2 && !yh)) continue;
            const long long oy = (b == 0) ? 0 : (b == 1 ? -g.sy : g.sy);
#pragma unroll
            for (int a = 0; a < 3; a++) {
                if ((a == 1 && !xl) || (a == 2 && !xh)) continue;
                if (a | b | c) x[n + oz + oy + ((a == 0) ? 0 : (a == 1 ? -1 : 1))] = v;
            }
        }
    }
}

// Same, for the tiled 3-D kernels: `inner` (block-uniform) says that no cell of the block's tile
// touches an x/y ghost, so only the two z-boundary planes take the ghost-writing path.
template <bool D3>
__device__ __forceinline__ void store_with_ghosts_tile(const Geom& g, double* x, int i, int j, int k, long long n,
                                                       double v, bool inner) {
    if (D3 && inner && k != 1 && k != g.nz - 2) x[n] = v;
    else store_with_ghosts<D3>(g, x, i, j, k, n, v);
}
// (3-D tile grids: blockIdx.x / .y index the tile)
__device__ __forceinline__ bool tile_is_inner() {
    return blockIdx.x != 0 && blockIdx.x != gridDim.x - 1 && blockIdx.y != 0 && blockIdx.y != gridDim.y - 1;
}

// ---------------------------------------------------------------------------------------
// Neumann ghost fill as a stand-alone pass ("apply BC" at the start/end of the Krylov solves).
// One thread per BOUNDARY cell of the local slab; ghost cells read their clamped interior source.
// Halo planes (non-physical sides) are left untouched.
// ---------------------------------------------------------------------------------------
__global__ void k_neumann_fill(Geom g, double* x, const Ctl* ctl, int seq, int honour_skip) {
    if (ctl && seq > ctl->done_seq && !honour_skip) return;
    if (ctl && honour_skip && ctl->skip_final_bc) return;
    // one thread per boundary cell (edge / corner cells appear in two or three face lists: they read
    // the same interior source and store the same value)
    const long long t = face_cell(g, (long long)blockIdx.x * blockDim.x + threadIdx.x);
    if (t < 0) return;
    int i = (int)(t % g.nx);
    int j = (int)((t / g.nx) % g.ny);
    int k = (int)(t / ((long long)g.nx * g.ny));
    int ci = i, cj = j, ck = k;
    if (i == 0 && g.glo[0]) ci = 1;
    if (i == g.nx - 1 && g.ghi[0]) ci = g.nx - 2;
    if (j == 0 && g.glo[1]) cj = 1;
    if (j == g.ny - 1 && g.ghi[1]) cj = g.ny - 2;
    if (g.dim3) {
        if (k == 0 && g.glo[2]) ck = 1;
        if (k == g.nz - 1 && g.ghi[2]) ck = g.nz - 2;
    }
    if (ci == i && cj == j && ck == k) return;
    // a ghost cell on a halo plane/row mirrors a cell of that same halo plane: leave it to the exchange
    if (g.dim3 ? ((k == 0 && !g.glo[2]) || (k == g.nz - 1 && !g.ghi[2]))
               : ((j == 0 && !g.glo[1]) || (j == g.ny - 1 && !g.ghi[1])))
        return;
    x[t] = x[(long long)ck * g.nx * g.ny + (long long)cj * g.nx + ci];
}

// ---------------------------------------------------------------------------------------
// Stationary methods: shared convergence decision (linear_solver.c:415-484).
// Called by every thread at the start of the kernel that would begin iteration s, when the
// L-inf residual of x_s sits in dot[s&1].  Returns true when the kernel must not run.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ bool stationary_decide(Ctl* ctl, const KArgs& a) {
    if (a.seq > ctl->done_seq) return true;
    const int s = a.it;
    const double res = ctl->dot[s & 1];
    bool fin = false, checked = true;
    int status = POISSON_CONVERGED, iters = 0;
    double tol;
    if (s == 0) {
        tol = fmax(a.rel_tol * res, a.abs_tol);
        if (res < a.abs_tol) fin = true;
    } else {
        tol = ctl->tol;
        checked = ((s - 1) % a.check) == 0;
        if (checked && (res < tol || res < a.abs_tol)) {
            fin = true;
            iters = s;
        } else if (s >= a.max_iter) {
            fin = true;
            status = POISSON_MAX_ITER;
            iters = a.max_iter + 1;  // reference reports iter+1 after the loop ran out (:478)
        }
    }
    if (leader()) {
        if (s == 0) {
            ctl->r0 = res;
            ctl->tol = tol;
        }
        if (checked) ctl->res = res;
        if (fin) {
            ctl->status = status;
            ctl->iterations = iters;
            ctl->result_buf = s & 1;
            ctl->skip_final_bc = 0;
            ctl->done_seq = a.seq - 1;
        }
    }
    return fin;
}

__global__ void k_stationary_final(Ctl* ctl, KArgs a) { (void)stationary_decide(ctl, a); }

// Jacobi sweep s: reads x_s (ghosts valid), writes x_{s+1} with its Neumann ghosts, and gets
// the L-inf residual of x_s for free from the same loads.  24 B/cell: x in, rhs in, x out.
template <bool D3, bool RES>
__global__ void __launch_bounds__(TX* TY)
k_jacobi(Geom g, const double* __restrict__ xin, double* __restrict__ xout,
         const double* __restrict__ rhs, Ctl* ctl, RedBuf rb, KArgs a) {
    // sweep q decides about x_{q-1}, whose residual sweep q-1 left in dot[(q-1)&1]
    if (a.it > 0) {
        KArgs d = a;
        d.it = a.it - 1;
        if (stationary_decide(ctl, d)) return;
    } else if (a.seq > ctl->done_seq) {
        return;
    }
    int i, j, kb, ke;
    const bool act = column(g, a.zchunk, i, j, kb, ke);
    double rmax[1] = {0.0};
    if (act) {
        long long n = (long long)kb * g.sz + (long long)j * g.sy + i;
        double zm = D3 ? xin[n - g.sz] : 0.0, zc = xin[n], zp = 0.0;
        for (int k = kb; k < ke; k++, n += g.sz) {
            if (D3) zp = xin[n + g.sz];
            const double xe = xin[n + 1], xw = xin[n - 1], yn = xin[n + g.sy], ys = xin[n - g.sy];
            const double f = __ldg(rhs + n);
            if (RES) {
                const double lap = ediv(xe - 2.0 * zc + xw, g.dx2, g.inv_dx2) + ediv(yn - 2.0 * zc + ys, g.dy2, g.inv_dy2) +
                                   (zp + zm - 2.0 * zc) * g.inv_dz2;
                max_into(rmax[0], fabs(lap - f));
            }
            const double v = -(f - ediv(xe + xw, g.dx2, g.inv_dx2) - ediv(yn + ys, g.dy2, g.inv_dy2) -
                               (zp + zm) * g.inv_dz2) * g.inv_factor;
            store_with_ghosts<D3>(g, xout, i, j, k, n, v);
            zm = zc;
            zc = zp;
        }
    }
    if (RES) block_reduce_finish<1, true>(rmax, rb, &ctl->dot[a.it & 1]);
}

// One colour of Red-Black SOR, in place.  colour 0 = first half-sweep = (i+j+k) odd (global indices).
template <bool D3>
__global__ void __launch_bounds__(TX* TY)
k_redblack(Geom g, double* x, const double* __restrict__ rhs, Ctl* ctl, KArgs a, int colour,
           int decide) {
    if (decide ? stationary_decide(ctl, a) : (a.seq > ctl->done_seq)) return;
    int i, j, kb, ke;
    if (!column(g, a.zchunk, i, j, kb, ke)) return;
    long long n = (long long)kb * g.sz + (long long)j * g.sy + i;
    const int par0 = (int)((g.gi0 + i + g.gj0 + j + g.gk0) & 1);
    for (int k = kb; k < ke; k++, n += g.sz) {
        // first half-sweep: i_start = ((j+k)%2==0) ? 1 : 2, i.e. the cells with (i+j+k) ODD
        // (linear_solver_redblack.c:103-105; the "red = even" comment there does not match the code)
        if (((par0 + k) & 1) == colour) continue;
        const double zp = D3 ? x[n + g.sz] : 0.0, zm = D3 ? x[n - g.sz] : 0.0;
        const double pn = -(__ldg(rhs + n) - ediv(x[n + 1] + x[n - 1], g.dx2, g.inv_dx2) -
                            ediv(x[n + g.sy] + x[n - g.sy], g.dy2, g.inv_dy2) - (zp + zm) * g.inv_dz2) *
                          g.inv_factor;
        const double xc = x[n];
        store_with_ghosts<D3>(g, x, i, j, k, n, xc + a.omega * (pn - xc));
    }
}

// L-inf residual of x (after BC) into dot[slot] -- linear_solver.c:304-346.
template <bool D3>
__global__ void __launch_bounds__(TX* TY)
k_linf_residual(Geom g, const double* __restrict__ x, const double* __restrict__ rhs, Ctl* ctl,
                RedBuf rb, KArgs a, int slot) {
    if (a.seq > ctl->done_seq) return;
    int i, j, kb, ke;
    const bool act = column(g, a.zchunk, i, j, kb, ke);
    double rmax[1] = {0.0};
    if (act) {
        long long n = (long long)kb * g.sz + (long long)j * g.sy + i;
        double zm = D3 ? x[n - g.sz] : 0.0, zc = x[n], zp = 0.0;
        for (int k = kb; k < ke; k++, n += g.sz) {
            if (D3) zp = x[n + g.sz];
            const double lap = ediv(x[n + 1] - 2.0 * zc + x[n - 1], g.dx2, g.inv_dx2) +
                               ediv(x[n + g.sy] - 2.0 * zc + x[n - g.sy], g.dy2, g.inv_dy2) +
                               (zp + zm - 2.0 * zc) * g.inv_dz2;
            max_into(rmax[0], fabs(lap - __ldg(rhs + n)));
            zm = zc;
            zc = zp;
        }
    }
    block_reduce_finish<1, true>(rmax, rb, &ctl->dot[slot]);
}


// Pointwise sweeps (axpy-type updates with dots): 1-D blocks of PW_THREADS along x, each block a
// chunk of `rpb` consecutive interior rows (a row = one (j,k) pair), so a thread streams rpb
// independent cells -- the same code serves 3-D grids and 2-D grids (where the z-marching layout
// would give one cell per thread).
constexpr int PW_THREADS = 256;
template <class F>
__device__ __forceinline__ void pointwise_rows(const Geom& g, int rpb, F f) {
    // x tiles start at i = 0 (thread 0 of the first tile idles on the boundary cell): a warp's 256-byte accesses then
    // start on a 256-byte boundary.  Stores that start one element off it run at 0.6 of the aligned rate on this GPU
    // (tools/copy_probe.cu, profiles/r2z_copy_probe.jsonl).
    const int i = blockIdx.x * PW_THREADS + threadIdx.x;
    const int nyi = g.ny - 2;
    const long long rows = (long long)nyi * (g.k1 - g.k0);
    const long long r0 = (long long)blockIdx.y * rpb;
    const long long r1 = r0 + rpb < rows ? r0 + rpb : rows;
    if (i < 1 || i >= g.nx - 1) return;
    int j = 1 + (int)(r0 % nyi), k = g.k0 + (int)(r0 / nyi);
    long long n = (long long)k * g.sz + (long long)j * g.sy + i;
#pragma unroll 4
    for (long long rr = r0; rr < r1; rr++) {
        f(n);
        if (++j == g.ny - 1) {  // next plane: skip the two boundary rows
            j = 1;
            n += 3 * g.sy;
        } else {
            n += g.sy;
        }
    }
}

// Pointwise sweep over the interior cells: 3-D grids keep the z-marching column layout of the
// stencil kernels (all blocks stream the same few planes at a time -- measured 10 % faster than
// row chunks at 512^3, fewer concurrent DRAM streams); 2-D grids use row chunks.
template <bool D3, class F>
__device__ __forceinline__ void pointwise_cells(const Geom& g, int chunk, F f) {
    if (D3) {
        int i, j, kb, ke;
        if (!column(g, chunk, i, j, kb, ke)) return;
        long long n = (long long)kb * g.sz + (long long)j * g.sy + i;
        for (int k = kb; k < ke; k++, n += g.sz) f(n);
    } else {
        pointwise_rows(g, chunk, f);
    }
}

// The same sweep with the loads of U consecutive cells of a thread issued before the first store of the batch.
// An in-place update (x += .., r -= ..) written cell by cell leaves one or two loads in flight per thread: the
// compiler cannot prove the arrays distinct through the functor and orders every load of x / r behind the previous
// cell's stores (SASS of the first version: LDG x -> STG x -> LDG r -> STG r, two memory latencies per cell;
// ncu on 4096^2: 20 % issue slots, everything waiting on the long scoreboard at 57 % of the DRAM peak).
// ld(n) returns the operands of cell n by value, ap(n, v) consumes them; cells are visited in the same order.
template <bool D3, int U, class V, class LD, class AP>
__device__ __forceinline__ void pointwise_cells_batched(const Geom& g, int chunk, LD ld, AP ap) {
    long long n, stride;
    int count;
    if (D3) {
        int i, j, kb, ke;
        if (!column(g, chunk, i, j, kb, ke)) return;
        n = (long long)kb * g.sz + (long long)j * g.sy + i;
        stride = g.sz;
        count = ke - kb;
    } else {  // (2-D grids only: the rows of a chunk are consecutive rows of the one plane)
        const int i = blockIdx.x * PW_THREADS + threadIdx.x;  // (aligned x tiles, see pointwise_rows)
        const int nyi = g.ny - 2;
        const int r0 = blockIdx.y * chunk, r1 = min(r0 + chunk, nyi);
        if (i < 1 || i >= g.nx - 1 || r0 >= r1) return;
        n = (long long)(1 + r0) * g.sy + i;
        stride = g.sy;
        count = r1 - r0;
    }
    int c = 0;
    for (; c + U <= count; c += U, n += U * stride) {
        V v[U];
#pragma unroll
        for (int u = 0; u < U; u++) v[u] = ld(n + u * stride);
#pragma unroll
        for (int u = 0; u < U; u++) ap(n + u * stride, v[u]);
    }
    for (; c < count; c++, n += stride) ap(n, ld(n));
}

// =======================================================================================
// Krylov solvers.  A = -lap on interior cells; r, p, v, s, t keep ZERO on physical boundary
// cells (never written), exactly as the reference's calloc'ed work vectors.
// Dot slots in Ctl::dot  (par = iteration parity, so a kernel never overwrites a value that
// late-starting blocks of the same launch still have to read):
//   CG      : [2*par+0]=r.r  [2*par+1]=r.z   [4+par]=p.Ap
//   BiCGSTAB: [2*par+0]=r.r  [2*par+1]=rhat.r  [4+par]=rhat.v  [8+4*par+{0,1,2}]=s.s, t.s, t.t
// =======================================================================================

__device__ __forceinline__ void finish(Ctl* ctl, const KArgs& a, int status, int iters, double res,
                                       int skip_bc, int after_this_kernel) {
    if (!leader()) return;
    ctl->status = status;
    ctl->iterations = iters;
    ctl->res = res;
    ctl->skip_final_bc = skip_bc;
    ctl->done_seq = after_this_kernel ? a.seq : a.seq - 1;
}

// r = -rhs + lap(x) (linear_solver_cg.c:134-158), d0 = r or dinv*r into p (CG), rhat = r (BiCGSTAB).
// Dots: r.r and r.z (CG) / r.r twice (BiCGSTAB: rhat.r == r.r at start).
template <bool D3>
__global__ void __launch_bounds__(TX* TY)
k_krylov_init(Geom g, const double* __restrict__ x, const double* __restrict__ rhs,
              double* __restrict__ r, double* __restrict__ p0, double* __restrict__ rhat, Ctl* ctl,
              RedBuf rb, KArgs a, int precond) {
    int i, j, kb, ke;
    const bool act = column(g, a.zchunk, i, j, kb, ke);
    double acc[2] = {0.0, 0.0};
    if (act) {
        long long n = (long long)kb * g.sz + (long long)j * g.sy + i;
        double zm = D3 ? x[n - g.sz] : 0.0, zc = x[n], zp = 0.0;
        for (int k = kb; k < ke; k++, n += g.sz) {
            if (D3) zp = x[n + g.sz];
            const double lap = (x[n + 1] - 2.0 * zc + x[n - 1]) * g.inv_dx2 +
                               (x[n + g.sy] - 2.0 * zc + x[n - g.sy]) * g.inv_dy2 +
                               (zp + zm - 2.0 * zc) * g.inv_dz2;
            const double rv = -__ldg(rhs + n) + lap;
            r[n] = rv;
            const double zv = precond ? a.dinv * rv : rv;
            if (p0) p0[n] = zv;
            if (rhat) rhat[n] = rv;
            acc[0] += rv * rv;
            acc[1] += rv * zv;
            zm = zc;
            zc = zp;
        }
    }
    block_reduce_finish<2, false>(acc, rb, &ctl->dot[0]);
}

// Generic fused "direction update + mat-vec + dots" sweep:
//   c(n)   = Op::comb(n)              (a linear combination of up to three vectors)
//   cout   = c                        (written on owned cells and on halo planes)
//   aout   = A c = -lap(c)
//   dots   = Op::dots(...)
// c is re-evaluated at the four in-plane neighbours (their operands come through L1) and
// carried in registers along z, so each operand is read once from HBM.
template <bool D3, class Op>
__device__ __forceinline__ void comb_matvec(const Geom& g, const KArgs& a, const Op& op,
                                            double* __restrict__ cout, double* __restrict__ aout,
                                            double (&acc)[Op::NV]) {
    int i, j, kb, ke;
    if (!column(g, a.zchunk, i, j, kb, ke)) return;
    long long n = (long long)kb * g.sz + (long long)j * g.sy + i;
    double cm = D3 ? op.comb(n - g.sz) : 0.0, cc = op.comb(n), cp = 0.0;
    if (D3 && kb == 1 && !g.glo[2]) cout[n - g.sz] = cm;  // keep the low halo plane of c current
    for (int k = kb; k < ke; k++, n += g.sz) {
        if (D3) cp = op.comb(n + g.sz);
        const double ce = op.comb(n + 1), cw = op.comb(n - 1);
        const double cn = op.comb(n + g.sy), cs = op.comb(n - g.sy);
        const double Ac = -((ce - 2.0 * cc + cw) * g.inv_dx2 + (cn - 2.0 * cc + cs) * g.inv_dy2 +
                            (cp + cm - 2.0 * cc) * g.inv_dz2);
        cout[n] = cc;
        aout[n] = Ac;
        if (!D3) {  // 2-D slabs are cut along y: keep the halo ROWS of c current
            if (j == 1 && !g.glo[1]) cout[n - g.sy] = cs;
            if (j == g.ny - 2 && !g.ghi[1]) cout[n + g.sy] = cn;
        }
        op.dots(n, cc, Ac, acc);
        cm = cc;
        cc = cp;
    }
    if (D3 && ke == g.nz - 1 && !g.ghi[2]) cout[n] = cc;  // high halo plane (n == plane nz-1 now)
}

// ---- CG -------------------------------------------------------------------------------
struct CgDirOp {
    static constexpr int NV = 1, NS = 2;   // TMA operands: r, pold
    __device__ __forceinline__ double combv(const double (&q)[2]) const {
        if (first) return q[1];
        return (precond ? dinv * q[0] : q[0]) + beta * q[1];
    }
    __device__ __forceinline__ double pw(long long) const { return 0.0; }
    __device__ __forceinline__ void dotsv(double, double c, double Ac, double (&acc)[1]) const { acc[0] += c * Ac; }
    const double* r;
    const double* pold;
    double beta, dinv;
    int first, precond;
    __device__ __forceinline__ double comb(long long n) const {
        const double po = __ldg(pold + n);
        if (first) return po;
        const double rv = __ldg(r + n);
        return (precond ? dinv * rv : rv) + beta * po;
    }
    __device__ __forceinline__ void dots(long long, double c, double Ac, double (&acc)[1]) const {
        acc[0] += c * Ac;
    }
};

// Start of CG iteration `it`: convergence / breakdown decisions of iteration it-1
// (linear_solver_cg.c:397-415), then p = z + beta p, Ap = A p, p.Ap.
// Decisions at the start of CG iteration a.it; false => the kernel must not run.
__device__ __forceinline__ bool cg_dir_prologue(Ctl* ctl, const KArgs& a, int precond, double& beta) {
    if (a.seq > ctl->done_seq) return false;
    const int it = a.it, par = it & 1;
    const double rr = ctl->dot[2 * par], rho_new = ctl->dot[2 * par + precond];
    const double res = sqrt(rr);
    beta = 0.0;
    if (it == 0) {
        const double tol = fmax(a.rel_tol * res, a.abs_tol);
        if (leader()) {
            ctl->r0 = res;
            ctl->tol = tol;
        }
        if (res < a.abs_tol) {
            finish(ctl, a, POISSON_CONVERGED, 0, res, 1, 0);
            return false;
        }
    } else {
        const double tol = ctl->tol;
        const double rho_old = ctl->dot[2 * (par ^ 1) + precond];
        const bool conv = (res < tol || res < a.abs_tol);
        if (((it - 1) % a.check) == 0 && conv) {
            finish(ctl, a, POISSON_CONVERGED, it, res, 0, 0);
            return false;
        }
        if (fabs(rho_old) < 1e-30) {
            finish(ctl, a, POISSON_STAGNATED, it, res, 1, 0);
            return false;
        }
        if (it >= a.max_iter) {
            finish(ctl, a, conv ? POISSON_CONVERGED : POISSON_MAX_ITER, a.max_iter, res, 0, 0);
            return false;
        }
        beta = rho_new / rho_old;
    }
    return true;
}

template <bool D3>
__global__ void __launch_bounds__(TX* TY)
k_cg_dir(Geom g, const double* __restrict__ r, const double* __restrict__ pold,
         double* __restrict__ pnew, double* __restrict__ Ap, Ctl* ctl, RedBuf rb, KArgs a,
         int precond, int decide_only) {
    double beta;
    if (!cg_dir_prologue(ctl, a, precond, beta) || decide_only) return;
    const int par = a.it & 1;
    CgDirOp op{r, pold, beta, a.dinv, a.it == 0, precond};
    double acc[1] = {0.0};
    comb_matvec<D3>(g, a, op, pnew, Ap, acc);
    block_reduce_finish<1, false>(acc, rb, &ctl->dot[4 + par]);
}

// alpha = rho / p.Ap ; x += alpha p ; r -= alpha Ap ; r.r and r.z  (linear_solver_cg.c:372-397)
template <bool D3>
__global__ void __launch_bounds__(TX* TY, 4)
k_cg_update(Geom g, double* __restrict__ x, double* __restrict__ r, const double* __restrict__ p,
            const double* __restrict__ Ap, Ctl* ctl, RedBuf rb, KArgs a, int precond) {
    if (a.seq > ctl->done_seq) return;
    const int par = a.it & 1;
    const double pAp = ctl->dot[4 + par];
    if (fabs(pAp) < 1e-30) {
        finish(ctl, a, POISSON_STAGNATED, a.it + 1, a.it == 0 ? ctl->r0 : sqrt(ctl->dot[2 * par]), 1, 0);
        return;
    }
    const double alpha = ctl->dot[2 * par + precond] / pAp;
    double acc[2] = {0.0, 0.0};
    struct Cell {
        double x, r, p, ap;
    };
    pointwise_cells_batched<D3, D3 ? 2 : 4, Cell>(
        g, a.zchunk, [&](long long n) { return Cell{x[n], r[n], __ldg(p + n), __ldg(Ap + n)}; },
        [&](long long n, const Cell& c) {
            x[n] = c.x + alpha * c.p;
            const double rv = c.r - alpha * c.ap;
            r[n] = rv;
            acc[0] += rv * rv;
            acc[1] += rv * (a.dinv * rv);
        });
    block_reduce_finish<2, false>(acc, rb, &ctl->dot[2 * (par ^ 1)]);
}

// ---- BiCGSTAB -------------------------------------------------------------------------
struct BiDirOp {  // p = r + beta (p - omega v) ; v = A p ; rhat.v
    static constexpr int NV = 1, NS = 3;   // TMA operands: r, pold, vold
    __device__ __forceinline__ double combv(const double (&q)[3]) const {
        if (first) return q[0];
        return q[0] + beta * (q[1] - omega * q[2]);
    }
    __device__ __forceinline__ double pw(long long n) const { return __ldg(rhat + n); }
    __device__ __forceinline__ void dotsv(double rh, double, double Ac, double (&acc)[1]) const { acc[0] += rh * Ac; }
    const double *r, *pold, *vold, *rhat;
    double beta, omega;
    int first;  // iteration 0: p = v = 0 (linear_solver_bicgstab.c:307-308) without reading them
    __device__ __forceinline__ double comb(long long n) const {
        const double rv = __ldg(r + n);
        if (first) return rv;
        return rv + beta * (__ldg(pold + n) - omega * __ldg(vold + n));
    }
    __device__ __forceinline__ void dots(long long n, double, double Ac, double (&acc)[1]) const {
        acc[0] += __ldg(rhat + n) * Ac;
    }
};
struct BiSOp {  // s = r - alpha v ; t = A s ; s.s, t.s, t.t
    static constexpr int NV = 3, NS = 2;   // TMA operands: r, v
    __device__ __forceinline__ double combv(const double (&q)[2]) const { return q[0] - alpha * q[1]; }
    __device__ __forceinline__ double pw(long long) const { return 0.0; }
    __device__ __forceinline__ void dotsv(double, double c, double Ac, double (&acc)[3]) const {
        acc[0] += c * c;
        acc[1] += Ac * c;
        acc[2] += Ac * Ac;
    }
    const double *r, *v;
    double alpha;
    __device__ __forceinline__ double comb(long long n) const {
        return __ldg(r + n) - alpha * __ldg(v + n);
    }
    __device__ __forceinline__ void dots(long long, double c, double Ac, double (&acc)[3]) const {
        acc[0] += c * c;
        acc[1] += Ac * c;
        acc[2] += Ac * Ac;
    }
};

// Scalars of BiCGSTAB iteration `it` recomputed from the raw dots (cheap, avoids a control kernel).
__device__ __forceinline__ double bi_alpha(const Ctl* ctl, int it) {
    return ctl->dot[2 * (it & 1) + 1] / ctl->dot[4 + (it & 1)];
}
__device__ __forceinline__ double bi_omega(const Ctl* ctl, int it) {
    return ctl->dot[8 + 4 * (it & 1) + 1] / ctl->dot[8 + 4 * (it & 1) + 2];
}

// Start of iteration `it` (linear_solver_bicgstab.c:338-370 preceded by the tail checks :451-474
// of iteration it-1).
__device__ __forceinline__ bool bicg_dir_prologue(Ctl* ctl, const KArgs& a, double& beta, double& omega_old) {
    if (a.seq > ctl->done_seq) return false;
    const int it = a.it, par = it & 1;
    const double res = sqrt(ctl->dot[2 * par]);
    const double rho_new = ctl->dot[2 * par + 1];
    omega_old = 1.0;
    if (it == 0) {
        const double tol = fmax(a.rel_tol * res, a.abs_tol);
        if (leader()) {
            ctl->r0 = res;
            ctl->tol = tol;
        }
        if (res < a.abs_tol) {
            finish(ctl, a, POISSON_CONVERGED, 0, res, 1, 0);
            return false;
        }
        if (fabs(rho_new) < 1e-30) {
            finish(ctl, a, POISSON_STAGNATED, 1, res, 1, 0);
            return false;
        }
        beta = rho_new;  // (rho_new / 1) * (1 / 1)
    } else {
        const double tol = ctl->tol;
        const bool conv = (res < tol || res < a.abs_tol);
        if (((it - 1) % a.check) == 0 && conv) {
            finish(ctl, a, POISSON_CONVERGED, it, res, 0, 0);
            return false;
        }
        omega_old = bi_omega(ctl, it - 1);
        if (fabs(omega_old) < 1e-30) {
            finish(ctl, a, POISSON_STAGNATED, it, res, 1, 0);
            return false;
        }
        if (it >= a.max_iter) {
            finish(ctl, a, conv ? POISSON_CONVERGED : POISSON_MAX_ITER, a.max_iter, res, 0, 0);
            return false;
        }
        if (fabs(rho_new) < 1e-30) {
            finish(ctl, a, POISSON_STAGNATED, it + 1, res, 1, 0);
            return false;
        }
        const double rho_old = ctl->dot[2 * (par ^ 1) + 1];
        beta = (rho_new / rho_old) * (bi_alpha(ctl, it - 1) / omega_old);
    }
    return true;
}

template <bool D3>
__global__ void __launch_bounds__(TX* TY)
k_bicg_dir(Geom g, const double* __restrict__ r, const double* __restrict__ pold,
           const double* __restrict__ vold, const double* __restrict__ rhat,
           double* __restrict__ pnew, double* __restrict__ vnew, Ctl* ctl, RedBuf rb, KArgs a,
           int decide_only) {
    double beta, omega_old;
    if (!bicg_dir_prologue(ctl, a, beta, omega_old) || decide_only) return;
    BiDirOp op{r, pold, vold, rhat, beta, omega_old, a.it == 0};
    double acc[1] = {0.0};
    comb_matvec<D3>(g, a, op, pnew, vnew, acc);
    block_reduce_finish<1, false>(acc, rb, &ctl->dot[4 + (a.it & 1)]);
}

__device__ __forceinline__ bool bicg_s_prologue(Ctl* ctl, const KArgs& a, double& alpha) {
    if (a.seq > ctl->done_seq) return false;
    const int par = a.it & 1;
    const double rhv = ctl->dot[4 + par];
    if (fabs(rhv) < 1e-30) {
        finish(ctl, a, POISSON_STAGNATED, a.it + 1, a.it == 0 ? ctl->r0 : sqrt(ctl->dot[2 * par]), 1, 0);
        return false;
    }
    alpha = ctl->dot[2 * par + 1] / rhv;
    return true;
}

template <bool D3>
__global__ void __launch_bounds__(TX* TY)
k_bicg_s(Geom g, const double* __restrict__ r, const double* __restrict__ v,
         double* __restrict__ s, double* __restrict__ t, Ctl* ctl, RedBuf rb, KArgs a) {
    double alpha;
    if (!bicg_s_prologue(ctl, a, alpha)) return;
    BiSOp op{r, v, alpha};
    double acc[3] = {0.0, 0.0, 0.0};
    comb_matvec<D3>(g, a, op, s, t, acc);
    block_reduce_finish<3, false>(acc, rb, &ctl->dot[8 + 4 * (a.it & 1)]);
}

// Tail of the iteration (:396-451): early exit on ||s||, omega, x and r updates, r.r and rhat.r.
template <bool D3>
__global__ void __launch_bounds__(TX* TY)
k_bicg_update(Geom g, double* __restrict__ x, double* __restrict__ r, const double* __restrict__ p,
              const double* __restrict__ s, const double* __restrict__ t,
              const double* __restrict__ rhat, Ctl* ctl, RedBuf rb, KArgs a) {
    if (a.seq > ctl->done_seq) return;
    const int par = a.it & 1;
    const double alpha = bi_alpha(ctl, a.it);
    const double sn = sqrt(ctl->dot[8 + 4 * par]);
    const double tt = ctl->dot[8 + 4 * par + 2];
    const double tol = ctl->tol;
    int mode = 0;  // 0 full update, 1 x += alpha p only (then stop)
    if (sn < tol || sn < a.abs_tol) {
        mode = 1;
        finish(ctl, a, POISSON_CONVERGED, a.it + 1, sn, 0, 1);
    } else if (fabs(tt) < 1e-30) {
        mode = 1;
        finish(ctl, a, POISSON_STAGNATED, a.it + 1, sn, 1, 1);
    }
    const double omega = (mode == 0) ? ctl->dot[8 + 4 * par + 1] / tt : 0.0;
    double acc[2] = {0.0, 0.0};
    if (mode == 1) {
        pointwise_cells<D3>(g, a.zchunk, [&](long long n) { x[n] += alpha * __ldg(p + n); });
    } else {
        pointwise_cells<D3>(g, a.zchunk, [&](long long n) {
            const double sv = __ldg(s + n);
            x[n] += alpha * __ldg(p + n) + omega * sv;
            const double rv = sv - omega * __ldg(t + n);
            r[n] = rv;
            acc[0] += rv * rv;
            acc[1] += __ldg(rhat + n) * rv;
        });
        block_reduce_finish<2, false>(acc, rb, &ctl->dot[2 * (par ^ 1)]);
    }
}


#include "poisson_tma.inc"
#include "poisson_tma2.inc"
#include "poisson_tma3.inc"
#include "poisson_tma4.inc"

// =======================================================================================
// Host launchers
// =======================================================================================
#ifndef CFDB_FAST  // (a member of cfdb::LaunchCtx: defined once, by the exact build)
CFDB_ARITH_END
dim3 LaunchCtx::grid() const {
    const int nzi = g.k1 - g.k0;
    return dim3((g.nx + TX - 1) / TX, (g.ny + TY - 1) / TY, (nzi + zchunk - 1) / zchunk);
}
CFDB_ARITH_BEGIN
#endif
static inline dim3 blk() { return dim3(TX, TY); }
// grid of the pointwise sweeps: x tiles of PW_THREADS cells, row chunks sized for ~6 waves of
// 148 x 8 blocks, never more blocks than the partial-sum buffer has rows
static inline dim3 pw_grid(const LaunchCtx& L, int& rpb) {
    const long long xt = (L.g.nx - 1 + PW_THREADS - 1) / PW_THREADS;  // tiles from i = 0 up to i = nx - 2
    const long long rows = (long long)(L.g.ny - 2) * (L.g.k1 - L.g.k0);
    static const int waves = [] {  // tuning knob (default 6 waves of 148 x 8 blocks)
        const char* e = getenv("CFD_B200_PW_WAVES");
        const int v = e ? atoi(e) : 6;
        return v < 1 ? 1 : v;
    }();
    long long r = rows * xt / (148LL * 8 * waves);
    if (r < 4) r = 4;
    if (r > 64) r = 64;
    r = (r + 3) & ~3LL;  // whole batches of the 4-cell pointwise loop
    const long long cap = (long long)L.rb_blocks;
    while (xt * ((rows + r - 1) / r) > cap) r *= 2;
    rpb = (int)r;
    return dim3((unsigned)xt, (unsigned)((rows + r - 1) / r), 1);
}
static inline KArgs with_chunk(const LaunchCtx& L, KArgs a) {
    a.zchunk = L.zchunk;
    return a;
}

void launch_neumann_fill(const LaunchCtx& L, double* x, int seq, int honour_skip, bool use_ctl) {
    const int nb = (int)((face_cells(L.g) + 255) / 256);
    CFDB_LAUNCH(KK_NEUMANN, L.st,
                k_neumann_fill<<<nb, 256, 0, L.st>>>(L.g, x, use_ctl ? L.ctl : nullptr, seq, honour_skip));
}
void launch_jacobi(const LaunchCtx& L, const KArgs& a0, const double* xin, double* xout, const double* rhs) {
    const KArgs a = with_chunk(L, a0);
    // the residual of x_q is only consulted when the reference would have computed it
    const bool res = (a.it == 0) || ((a.it - 1) % a.check == 0);
    {
        TmaSet<1> tm;
        const double* const ops[1] = {xin};
        if (tma_maps(L, ops, tm)) {
            KArgs at = a0;
            at.zchunk = L.zchunk_tma;
            constexpr size_t SM = DirectPipe<1>::SMEM;
            CFDB_ONCE_PER_DEVICE(tma_optin(k_jacobi_tma<true, true>, SM), tma_optin(k_jacobi_tma<true, false>, SM),
                                tma_optin(k_jacobi_tma<false, true>, SM), tma_optin(k_jacobi_tma<false, false>, SM));
#define CFDB_JACT(D3, RES) CFDB_LAUNCH(KK_JACOBI, L.st, (k_jacobi_tma<D3, RES><<<tma_grid(L), Tile::THREADS_PW, SM, L.st>>>(tm, L.g, xout, rhs, L.ctl, L.rb, at)))
            if (L.g.dim3) { if (res) CFDB_JACT(true, true); else CFDB_JACT(true, false); }
            else { if (res) CFDB_JACT(false, true); else CFDB_JACT(false, false); }
#undef CFDB_JACT
            return;
        }
    }
    peer_pull_from_gate(L.st, a0.gate);  // no in-kernel gate on this path: pull the pending halo layers first
#define CFDB_JAC(D3, RES) CFDB_LAUNCH(KK_JACOBI, L.st, (k_jacobi<D3, RES><<<L.grid(), blk(), 0, L.st>>>(L.g, xin, xout, rhs, L.ctl, L.rb, a)))
    if (L.g.dim3) { if (res) CFDB_JAC(true, true); else CFDB_JAC(true, false); }
    else { if (res) CFDB_JAC(false, true); else CFDB_JAC(false, false); }
#undef CFDB_JAC
}
// One-pass Red-Black SOR iteration, TMA-capable grids (nx even).  3-D (poisson_tma3.inc): a single rank, or z-slabs
// with a second halo layer (>= 2 own planes, so that the layer behind a rank's edge plane is its own).  2-D
// (poisson_tma4.inc): a single rank.
bool rbsor_fused_available(const LaunchCtx& L, double* const* halo2) {
    const bool off = getenv("CFD_B200_NO_RB_FUSED") != nullptr;  // read per solve: tests toggle it
    if (off || !L.tc || !L.tc->usable) return false;
    if (!L.g.dim3) return L.g.glo[1] && L.g.ghi[1];
    if (L.g.glo[2] && L.g.ghi[2]) return true;
    return halo2 && halo2[0] && halo2[1] && L.g.nz >= 4;
}
// 2-D: x tiles times chunks of `yc` consecutive 64 x 16 tiles; two blocks per SM, aim at >= 4 waves, 4..64 tiles
// per block (a block's pipeline fill is one tile's load latency; 4096^2: 143 / 122 / 113 / 124 / 128 us per sweep
// with 4 / 7 / 16 / 32 / 64 tiles per block)
static int rbsor_fused2d_ychunk(const LaunchCtx& L, dim3& grid) {
    const int gx = Tile::GX(L.g.nx), ty = (L.g.ny - 2 + Tile::BY - 1) / Tile::BY;
    static const int forced = getenv("CFD_B200_YCHUNK_RB") ? atoi(getenv("CFD_B200_YCHUNK_RB")) : 0;
    int nch = (148 * 2 * 4 + gx - 1) / gx;
    int yc = (ty + nch - 1) / nch;
    yc = yc < 4 ? 4 : (yc > 64 ? 64 : yc);
    if (forced > 0) yc = forced;
    if (yc > ty) yc = ty;
    grid = dim3(gx, (ty + yc - 1) / yc, 1);
    return yc;
}
static bool launch_rbsor_fused2d(const LaunchCtx& L, const KArgs& a0, const double* xin, double* xout, const double* rhs) {
    const CUtensorMap *mx = L.tc->get_wide(xin), *mr = L.tc->get(rhs);
    if (!mx || !mr) return false;
    TmaSet<2> tm;
    tm.m[0] = *mx;
    tm.m[1] = *mr;
    KArgs at = a0;
    dim3 grid;
    at.zchunk = rbsor_fused2d_ychunk(L, grid);
    if ((size_t)grid.x * grid.y > L.rb_blocks) return false;
    const bool res = (at.it == 0) || ((at.it - 1) % at.check == 0);
    constexpr size_t SM = RbFused2D::SMEM;
    CFDB_ONCE_PER_DEVICE(tma_optin(k_rbsor_fused2d_tma<true>, SM), tma_optin(k_rbsor_fused2d_tma<false>, SM));
    if (res) CFDB_LAUNCH(KK_RBSOR_FUSED, L.st, (k_rbsor_fused2d_tma<true><<<grid, RbFused::THREADS, SM, L.st>>>(tm, L.g, xout, L.ctl, L.rb, at)));
    else CFDB_LAUNCH(KK_RBSOR_FUSED, L.st, (k_rbsor_fused2d_tma<false><<<grid, RbFused::THREADS, SM, L.st>>>(tm, L.g, xout, L.ctl, L.rb, at)));
    return true;
}
static int rbsor_fused_zchunk(const LaunchCtx& L, dim3& grid) {
    const int nzi = L.g.k1 - L.g.k0;
    const int gx = Tile::GX(L.g.nx), gy = (L.g.ny - 2 + Tile::BY - 1) / Tile::BY;
    static const int forced = getenv("CFD_B200_ZCHUNK_RB") ? atoi(getenv("CFD_B200_ZCHUNK_RB")) : 0;
    // One block per SM (154 KB of shared memory).  A block of zc planes runs zc + 4 pipeline steps after ~1 step of
    // load latency, and the last wave of blocks is rarely full: choose the chunk count (16..128 planes per block)
    // that maximises  blocks / (waves * 148)  *  zc / (zc + 5).  256^3: 9 chunks of 29 planes (measured 166 us per
    // sweep against 176 with 16 chunks of 16); 512^3 and up: 128 planes (measured flat between 64 and 128).
    int zc = nzi;
    {
        const int tiles = gx * gy;
        const int n_lo = (nzi + 127) / 128, n_hi = nzi / 16 > n_lo ? nzi / 16 : n_lo;
        double best = -1.0;
        // (from many short chunks to few long ones; a longer chunk must win by 2 %: neighbouring blocks of short
        // chunks stay within a few planes of each other and share their halo rows in L2)
        for (int nch = n_hi; nch >= n_lo; nch--) {
            const int c = (nzi + nch - 1) / nch;
            const int real = (nzi + c - 1) / c;  // chunks that length gives
            const long long blocks = (long long)tiles * real;
            const long long waves = (blocks + 147) / 148;
            const double eff = (double)blocks / (double)(waves * 148) * (double)c / (double)(c + 5);
            if (eff > best + 0.02) {
                best = eff;
                zc = c;
            }
        }
    }
    if (forced > 0) zc = forced;
    if (zc > nzi) zc = nzi;
    grid = dim3(gx, gy, (nzi + zc - 1) / zc);
    return zc;
}
bool launch_rbsor_fused(const LaunchCtx& L, const KArgs& a0, const double* xin, double* xout, const double* rhs,
                        double* const* halo2) {
    if (!rbsor_fused_available(L, halo2)) return false;
    if (!L.g.dim3) return launch_rbsor_fused2d(L, a0, xin, xout, rhs);
    const CUtensorMap *mx = L.tc->get_wide(xin), *mr = L.tc->get(rhs);
    if (!mx || !mr) return false;
    TmaSet<4> tm;
    tm.m[0] = *mx;
    tm.m[1] = *mr;
    tm.m[2] = tm.m[3] = *mx;  // (never dereferenced on a side with a physical boundary)
    for (int side = 0; side < 2; side++) {
        if (side ? L.g.ghi[2] : L.g.glo[2]) continue;
        const CUtensorMap* mh = L.tc->get_wide(halo2[side]);  // only plane 0 of this map exists
        if (!mh) return false;
        tm.m[2 + side] = *mh;
    }
    KArgs at = a0;
    dim3 grid;
    at.zchunk = rbsor_fused_zchunk(L, grid);
    if ((size_t)grid.x * grid.y * grid.z > L.rb_blocks) return false;
    const bool res = (at.it == 0) || ((at.it - 1) % at.check == 0);
    constexpr size_t SM = RbFused::SMEM;
    CFDB_ONCE_PER_DEVICE(tma_optin(k_rbsor_fused_tma<true>, SM), tma_optin(k_rbsor_fused_tma<false>, SM));
    if (res) CFDB_LAUNCH(KK_RBSOR_FUSED, L.st, (k_rbsor_fused_tma<true><<<grid, RbFused::THREADS, SM, L.st>>>(tm, L.g, xout, L.ctl, L.rb, at)));
    else CFDB_LAUNCH(KK_RBSOR_FUSED, L.st, (k_rbsor_fused_tma<false><<<grid, RbFused::THREADS, SM, L.st>>>(tm, L.g, xout, L.ctl, L.rb, at)));
    return true;
}

void launch_redblack(const LaunchCtx& L, const KArgs& a0, double* x, const double* rhs, int colour, int decide) {
    {
        TmaSet<1> tm;
        const double* const ops[1] = {x};
        if (tma_maps(L, ops, tm)) {
            KArgs at = a0;
            at.zchunk = L.zchunk_tma;
            constexpr size_t SM = DirectPipe<1>::SMEM;
            CFDB_ONCE_PER_DEVICE(tma_optin(k_redblack_tma<true>, SM), tma_optin(k_redblack_tma<false>, SM));
            if (L.g.dim3) CFDB_LAUNCH(KK_REDBLACK, L.st, (k_redblack_tma<true><<<tma_grid(L), Tile::THREADS_PW, SM, L.st>>>(tm, L.g, x, rhs, L.ctl, at, colour, decide)));
            else CFDB_LAUNCH(KK_REDBLACK, L.st, (k_redblack_tma<false><<<tma_grid(L), Tile::THREADS_PW, SM, L.st>>>(tm, L.g, x, rhs, L.ctl, at, colour, decide)));
            return;
        }
    }
    peer_pull_from_gate(L.st, a0.gate);  // no in-kernel gate on this path: pull the pending halo layers first
    const KArgs a = with_chunk(L, a0);
    if (L.g.dim3) CFDB_LAUNCH(KK_REDBLACK, L.st, k_redblack<true><<<L.grid(), blk(), 0, L.st>>>(L.g, x, rhs, L.ctl, a, colour, decide));
    else CFDB_LAUNCH(KK_REDBLACK, L.st, k_redblack<false><<<L.grid(), blk(), 0, L.st>>>(L.g, x, rhs, L.ctl, a, colour, decide));
}
void launch_linf_residual(const LaunchCtx& L, const KArgs& a0, const double* x, const double* rhs, int slot) {
    {
        TmaSet<1> tm;
        const double* const ops[1] = {x};
        if (tma_maps(L, ops, tm)) {
            KArgs at = a0;
            at.zchunk = L.zchunk_tma;
            constexpr size_t SM = DirectPipe<1>::SMEM;
            CFDB_ONCE_PER_DEVICE(tma_optin(k_linf_residual_tma<true>, SM), tma_optin(k_linf_residual_tma<false>, SM));
            if (L.g.dim3) CFDB_LAUNCH(KK_LINF, L.st, (k_linf_residual_tma<true><<<tma_grid(L), Tile::THREADS_PW, SM, L.st>>>(tm, L.g, rhs, L.ctl, L.rb, at, slot)));
            else CFDB_LAUNCH(KK_LINF, L.st, (k_linf_residual_tma<false><<<tma_grid(L), Tile::THREADS_PW, SM, L.st>>>(tm, L.g, rhs, L.ctl, L.rb, at, slot)));
            return;
        }
    }
    peer_pull_from_gate(L.st, a0.gate);  // no in-kernel gate on this path: pull the pending halo layers first
    const KArgs a = with_chunk(L, a0);
    if (L.g.dim3) CFDB_LAUNCH(KK_LINF, L.st, k_linf_residual<true><<<L.grid(), blk(), 0, L.st>>>(L.g, x, rhs, L.ctl, L.rb, a, slot));
    else CFDB_LAUNCH(KK_LINF, L.st, k_linf_residual<false><<<L.grid(), blk(), 0, L.st>>>(L.g, x, rhs, L.ctl, L.rb, a, slot));
}
void launch_stationary_final(const LaunchCtx& L, const KArgs& a) {
    CFDB_LAUNCH(KK_STAT_FINAL, L.st, k_stationary_final<<<1, 32, 0, L.st>>>(L.ctl, a));
}
void launch_krylov_init(const LaunchCtx& L, const KArgs& a0, const double* x, const double* rhs, double* r,
                        double* p0, double* rhat, int precond) {
    {
        TmaSet<1> tm;
        const double* const ops[1] = {x};
        if (tma_maps(L, ops, tm)) {
            KArgs at = a0;
            at.zchunk = L.zchunk_tma;
            constexpr size_t SM = DirectPipe<1>::SMEM;
            CFDB_ONCE_PER_DEVICE(tma_optin(k_krylov_init_tma<true>, SM), tma_optin(k_krylov_init_tma<false>, SM));
            if (L.g.dim3) CFDB_LAUNCH(KK_KRYLOV_INIT, L.st, (k_krylov_init_tma<true><<<tma_grid(L), Tile::THREADS_PW, SM, L.st>>>(tm, L.g, rhs, r, p0, rhat, L.ctl, L.rb, at, precond)));
            else CFDB_LAUNCH(KK_KRYLOV_INIT, L.st, (k_krylov_init_tma<false><<<tma_grid(L), Tile::THREADS_PW, SM, L.st>>>(tm, L.g, rhs, r, p0, rhat, L.ctl, L.rb, at, precond)));
            return;
        }
    }
    const KArgs a = with_chunk(L, a0);
    if (L.g.dim3) CFDB_LAUNCH(KK_KRYLOV_INIT, L.st, k_krylov_init<true><<<L.grid(), blk(), 0, L.st>>>(L.g, x, rhs, r, p0, rhat, L.ctl, L.rb, a, precond));
    else CFDB_LAUNCH(KK_KRYLOV_INIT, L.st, k_krylov_init<false><<<L.grid(), blk(), 0, L.st>>>(L.g, x, rhs, r, p0, rhat, L.ctl, L.rb, a, precond));
}
void launch_cg_dir(const LaunchCtx& L, const KArgs& a0, const double* r, const double* pold, double* pnew,
                   double* Ap, int precond, int decide_only) {
    if (!decide_only) {
        TmaSet<2> tm;
        const double* const ops[2] = {r, pold};
        if (tma_maps(L, ops, tm)) {
            KArgs a = a0;
            a.zchunk = L.zchunk_tma;
            constexpr size_t SM = TmaPipe<2>::SMEM;
            CFDB_ONCE_PER_DEVICE(tma_optin(k_cg_dir_tma<true>, SM), tma_optin(k_cg_dir_tma<false>, SM));
            if (L.g.dim3) CFDB_LAUNCH(KK_CG_DIR, L.st, (k_cg_dir_tma<true><<<tma_grid(L), Tile::THREADS, SM, L.st>>>(tm, L.g, pnew, Ap, L.ctl, L.rb, a, precond)));
            else CFDB_LAUNCH(KK_CG_DIR, L.st, (k_cg_dir_tma<false><<<tma_grid(L), Tile::THREADS, SM, L.st>>>(tm, L.g, pnew, Ap, L.ctl, L.rb, a, precond)));
            return;
        }
    }
    peer_pull_from_gate(L.st, a0.gate);  // no in-kernel gate on this path: pull the pending halo layers first
    const KArgs a = with_chunk(L, a0);
    const dim3 gr = decide_only ? dim3(1, 1, 1) : L.grid();
    if (L.g.dim3) CFDB_LAUNCH(KK_CG_DIR, L.st, k_cg_dir<true><<<gr, blk(), 0, L.st>>>(L.g, r, pold, pnew, Ap, L.ctl, L.rb, a, precond, decide_only));
    else CFDB_LAUNCH(KK_CG_DIR, L.st, k_cg_dir<false><<<gr, blk(), 0, L.st>>>(L.g, r, pold, pnew, Ap, L.ctl, L.rb, a, precond, decide_only));
}
void launch_cg_update(const LaunchCtx& L, const KArgs& a0, double* x, double* r, const double* p,
                      const double* Ap, int precond) {
    KArgs a = with_chunk(L, a0);
    if (L.g.dim3) {
        CFDB_LAUNCH(KK_CG_UPDATE, L.st, k_cg_update<true><<<L.grid(), blk(), 0, L.st>>>(L.g, x, r, p, Ap, L.ctl, L.rb, a, precond));
    } else {
        const dim3 gr = pw_grid(L, a.zchunk);
        CFDB_LAUNCH(KK_CG_UPDATE, L.st, k_cg_update<false><<<gr, PW_THREADS, 0, L.st>>>(L.g, x, r, p, Ap, L.ctl, L.rb, a, precond));
    }
}
void launch_bicg_dir(const LaunchCtx& L, const KArgs& a0, const double* r, const double* pold,
                     const double* vold, const double* rhat, double* pnew, double* vnew, int first,
                     int decide_only) {
    (void)first;
    if (!decide_only) {
        TmaSet<3> tm;
        const double* const ops[3] = {r, pold, vold};
        if (tma_maps(L, ops, tm)) {
            KArgs a = a0;
            a.zchunk = L.zchunk_tma;
            constexpr size_t SM = TmaPipe<3>::SMEM;
            CFDB_ONCE_PER_DEVICE(tma_optin(k_bicg_dir_tma<true>, SM), tma_optin(k_bicg_dir_tma<false>, SM));
            if (L.g.dim3) CFDB_LAUNCH(KK_BICG_DIR, L.st, (k_bicg_dir_tma<true><<<tma_grid(L), Tile::THREADS, SM, L.st>>>(tm, L.g, rhat, pnew, vnew, L.ctl, L.rb, a)));
            else CFDB_LAUNCH(KK_BICG_DIR, L.st, (k_bicg_dir_tma<false><<<tma_grid(L), Tile::THREADS, SM, L.st>>>(tm, L.g, rhat, pnew, vnew, L.ctl, L.rb, a)));
            return;
        }
    }
    peer_pull_from_gate(L.st, a0.gate);  // no in-kernel gate on this path: pull the pending halo layers first
    const KArgs a = with_chunk(L, a0);
    const dim3 gr = decide_only ? dim3(1, 1, 1) : L.grid();
    if (L.g.dim3) CFDB_LAUNCH(KK_BICG_DIR, L.st, k_bicg_dir<true><<<gr, blk(), 0, L.st>>>(L.g, r, pold, vold, rhat, pnew, vnew, L.ctl, L.rb, a, decide_only));
    else CFDB_LAUNCH(KK_BICG_DIR, L.st, k_bicg_dir<false><<<gr, blk(), 0, L.st>>>(L.g, r, pold, vold, rhat, pnew, vnew, L.ctl, L.rb, a, decide_only));
}
void launch_bicg_s(const LaunchCtx& L, const KArgs& a0, const double* r, const double* v, double* s, double* t) {
    {
        TmaSet<2> tm;
        const double* const ops[2] = {r, v};
        if (tma_maps(L, ops, tm)) {
            KArgs a = a0;
            a.zchunk = L.zchunk_tma;
            constexpr size_t SM = TmaPipe<2>::SMEM;
            CFDB_ONCE_PER_DEVICE(tma_optin(k_bicg_s_tma<true>, SM), tma_optin(k_bicg_s_tma<false>, SM));
            if (L.g.dim3) CFDB_LAUNCH(KK_BICG_S, L.st, (k_bicg_s_tma<true><<<tma_grid(L), Tile::THREADS, SM, L.st>>>(tm, L.g, s, t, L.ctl, L.rb, a)));
            else CFDB_LAUNCH(KK_BICG_S, L.st, (k_bicg_s_tma<false><<<tma_grid(L), Tile::THREADS, SM, L.st>>>(tm, L.g, s, t, L.ctl, L.rb, a)));
            return;
        }
    }
    peer_pull_from_gate(L.st, a0.gate);  // no in-kernel gate on this path: pull the pending halo layers first
    const KArgs a = with_chunk(L, a0);
    if (L.g.dim3) CFDB_LAUNCH(KK_BICG_S, L.st, k_bicg_s<true><<<L.grid(), blk(), 0, L.st>>>(L.g, r, v, s, t, L.ctl, L.rb, a));
    else CFDB_LAUNCH(KK_BICG_S, L.st, k_bicg_s<false><<<L.grid(), blk(), 0, L.st>>>(L.g, r, v, s, t, L.ctl, L.rb, a));
}
void launch_bicg_update(const LaunchCtx& L, const KArgs& a0, double* x, double* r, const double* p,
                        const double* s, const double* t, const double* rhat) {
    KArgs a = with_chunk(L, a0);
    if (L.g.dim3) {
        CFDB_LAUNCH(KK_BICG_UPDATE, L.st, k_bicg_update<true><<<L.grid(), blk(), 0, L.st>>>(L.g, x, r, p, s, t, rhat, L.ctl, L.rb, a));
    } else {
        const dim3 gr = pw_grid(L, a.zchunk);
        CFDB_LAUNCH(KK_BICG_UPDATE, L.st, k_bicg_update<false><<<gr, PW_THREADS, 0, L.st>>>(L.g, x, r, p, s, t, rhat, L.ctl, L.rb, a));
    }
}

CFDB_ARITH_END
}  // namespace cfdb
