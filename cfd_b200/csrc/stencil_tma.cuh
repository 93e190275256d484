// stencil_tma.cuh -- generic TMA plane pipeline for 7-point stencils on NF fields (sm_100a).
//
// One block = one 64 x 16 tile of cells marched along z (tma.cuh: Tile).  The {HX,HY,1} = {68,18,1} boxes
// of the NF stencil fields arrive in a ring of ST stages (one mbarrier per stage, bytes counted
// by the TMA unit).  For plane k a thread takes the in-plane neighbours of its 2 cells from the
// shared plane k, the z neighbours from registers (plane k-1, k) and from the freshly landed
// plane k+1.  Pointwise operands (centre only: rhs, u*, ...) are plain loads issued one plane
// ahead.  There is no block-wide barrier in the loop: a 17th warp is the PRODUCER (waits on the
// stage's empty-barrier, re-arms the full-barrier, issues the TMA boxes), the 16 consumer warps
// release a stage with one mbarrier arrive each, so warps drift apart by up to ST planes and
// absorb each other's latency.
//
// The functor supplies the arithmetic:
//   static constexpr int NF, NPW;                                 // NPW >= 1 (a dummy if none)
//   static constexpr bool STAGED;                                 // op.prime(n, c) is called once per own cell (3-D only)
//   static constexpr bool PRODUCER_WARP;                          // 544-thread launch (regs <= 96) or 512 (<= 128)
//   void pw_load(long long n, double (&pw)[NPW]) const;          // pointwise operands of cell n
//   void cell(const St<NF>& s, const double (&pw)[NPW], long long n, int i, int j, int k);
#pragma once
#include "common.cuh"
#include "tma.cuh"

namespace cfdb {

// Stencil values of NF fields at one cell: centre and z neighbours live in registers, the four
// in-plane neighbours are read from the shared plane at the point of use (keeps the register
// footprint of the multi-field kernels down).
template <int NF>
struct St {
    const double* pl;     // own cell of field 0 in the shared plane k; field f at pl + f * PLANE_PAD
    double zc[NF], zm[NF], zp[NF];
    __device__ __forceinline__ double c(int f) const { return zc[f]; }
    __device__ __forceinline__ double m(int f) const { return zm[f]; }
    __device__ __forceinline__ double p(int f) const { return zp[f]; }
    __device__ __forceinline__ double e(int f) const { return pl[f * Tile::PLANE_PAD + 1]; }
    __device__ __forceinline__ double w(int f) const { return pl[f * Tile::PLANE_PAD - 1]; }
    __device__ __forceinline__ double n(int f) const { return pl[f * Tile::PLANE_PAD + Tile::HX]; }
    __device__ __forceinline__ double s(int f) const { return pl[f * Tile::PLANE_PAD - Tile::HX]; }
};

template <int NF>
struct DirectPipe {
    static constexpr int ST = (NF == 1) ? 6 : (NF == 2 ? 4 : 3);
    static constexpr int RING = ST * NF * Tile::PLANE_PAD;
    static constexpr size_t SMEM = (size_t)RING * sizeof(double) + 2 * ST * sizeof(uint64_t);  // ring + full[] + empty[]
};

#ifdef __CUDACC__

// Chunk a block works on: with a halo gate pending the two edge chunks go LAST (order 1, 2, ..., n-2, 0, n-1), so the
// blocks that have to wait for a neighbour are scheduled after the interior ones.
__device__ __forceinline__ int gate_chunk(const HaloGate* hg, int b, int nb) {
    if (!hg || !hg->on || nb <= 2) return b;
    return b < nb - 2 ? b + 1 : (b == nb - 2 ? 0 : nb - 1);
}
// Called by ALL threads of a block before its first load.  lo / hi: the block reads the low / high halo layer.
// (xa, ya), bw x bh: the block's box in the layer (clipped to the grid here; the origin may be negative);
// d3 = false: the layers are ROWS (y-slabs) and the box is a row segment.  Neighbouring blocks copy overlapping cells
// with identical values (benign).
__device__ __forceinline__ void halo_gate_pull_box(const HaloGate* hg, const Geom& g, int xa, int ya, int bw, int bh,
                                                   bool lo, bool hi, bool d3) {
    if (!hg || !hg->on) return;
    const bool need0 = lo && hg->box[0], need1 = hi && hg->box[1];
    if (!need0 && !need1) return;  // (block-uniform)
    __shared__ int gate_ok;
    if (threadIdx.x == 0) {
        bool ok = true;
        if (need0) ok = wait_flag(hg->flag[0], hg->seq, hg->timeout_flag, 1) && ok;
        if (need1) ok = wait_flag(hg->flag[1], hg->seq, hg->timeout_flag, 2) && ok;
        gate_ok = ok;
    }
    __syncthreads();
    if (gate_ok) {
        const int x0 = max(xa, 0), y0 = d3 ? max(ya, 0) : 0;
        const int w = min(xa + bw, g.nx) - x0;
        const int h = d3 ? min(ya + bh, g.ny) - y0 : 1;
        for (int side = 0; side < 2; side++) {
            if (!(side ? need1 : need0)) continue;
            for (int f = 0; f < hg->nf; f++) {
                const double* src = hg->box[side] + (long long)f * hg->cnt;
                double* dst = hg->halo[side][f];
                for (int e = threadIdx.x; e < w * h; e += blockDim.x) {
                    const int r = e / w, c = e - r * w;
                    const long long off = d3 ? (long long)(y0 + r) * g.nx + x0 + c : (long long)(x0 + c);
                    dst[off] = __ldcg(src + off);  // written by another GPU: read past L1
                }
            }
        }
        __threadfence();
        fence_proxy_async_global();
    }
    __syncthreads();
    fence_proxy_async_global();  // (the thread that issues the TMA loads is one of us)
}
// ... for the {HX, HY} box of a tile with origin (x0, y0)
__device__ __forceinline__ void halo_gate_pull(const HaloGate* hg, const Geom& g, int x0, int y0, bool lo, bool hi,
                                               bool d3) {
    halo_gate_pull_box(hg, g, x0, y0, Tile::HX, Tile::HY, lo, hi, d3);
}

// 2-D grids (nz == 1): the pipeline marches along y instead -- a block owns `ychunk` consecutive
// 64 x 16 tiles of one x column of tiles and streams their {68,18,1} boxes through the same ring
// (each box carries its own y halo; the 2 shared rows per tile come from L2).
template <class Op>
__device__ __forceinline__ void tma_direct_sweep_2d(const TmaSet<Op::NF>& tm, const Geom& g, int ychunk, Op& op,
                                                    const HaloGate* hg = nullptr) {
    constexpr int NF = Op::NF, NPW = Op::NPW, ST = DirectPipe<NF>::ST;
    constexpr int PP = Tile::PLANE_PAD;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    double* ring = reinterpret_cast<double*>(smem_raw);
    uint64_t* full = reinterpret_cast<uint64_t*>(ring + DirectPipe<NF>::RING);
    const int tid = threadIdx.x;
    const int x0 = Tile::X0(blockIdx.x);  // box origin; own cells start XO columns to the right
    const int tiles_y = (g.ny - 2 + Tile::BY - 1) / Tile::BY;
    const int ty0 = gate_chunk(hg, blockIdx.y, gridDim.y) * ychunk;
    const int nt = min(ychunk, tiles_y - ty0);
    uint64_t* empty = full + ST;
    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < ST; s++) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], Tile::WARPS);
        }
        mbar_fence_init();
    }
    __syncthreads();
    halo_gate_pull(hg, g, x0, 0, ty0 == 0, ty0 + nt == tiles_y, false);
    constexpr bool PWARP = Op::PRODUCER_WARP;
    int next_q = 0;
    auto issue = [&](int q) {
        const int slot = q % ST;
        mbar_expect_tx(&full[slot], NF * Tile::PLANE_BYTES);
#pragma unroll
        for (int f = 0; f < NF; f++)
            tma_load_plane(ring + (slot * NF + f) * PP, &tm.m[f], x0, (ty0 + q) * Tile::BY, 0, &full[slot]);
    };
    if (PWARP) {
        if (tid >= Tile::THREADS) {  // producer warp
            if (tid == Tile::THREADS) {
                for (int q = 0; q < nt; q++) {
                    if (q >= ST) mbar_wait(&empty[q % ST], ((q / ST) - 1) & 1);
                    issue(q);
                }
            }
            return;
        }
    } else if (tid == 0) {
        for (; next_q < ST && next_q < nt; next_q++) issue(next_q);
    }
    const int lane = tid & 31;
    const int lx = tid & (Tile::BX - 1), ly = tid >> 6;
    const int i = x0 + Tile::XO + lx;
    const bool actx = (i >= 1) && (i < g.nx - 1);
    int o[2];
#pragma unroll
    for (int c = 0; c < 2; c++) o[c] = Tile::OWN0(lx, ly + 8 * c);
    double pwn[2][NPW];
    auto jrow = [&](int t, int c) { return (ty0 + t) * Tile::BY + 1 + ly + 8 * c; };
#pragma unroll
    for (int c = 0; c < 2; c++)
        if (actx && jrow(0, c) < g.ny - 1) op.pw_load((long long)jrow(0, c) * g.sy + i, pwn[c]);
    for (int t = 0; t < nt; t++) {
        double pw[2][NPW];
#pragma unroll
        for (int c = 0; c < 2; c++) {
#pragma unroll
            for (int v = 0; v < NPW; v++) pw[c][v] = pwn[c][v];
            if (t + 1 < nt && actx && jrow(t + 1, c) < g.ny - 1) op.pw_load((long long)jrow(t + 1, c) * g.sy + i, pwn[c]);
        }
        const int slot = t % ST;
        mbar_wait(&full[slot], (t / ST) & 1);
        const double* cur = ring + slot * NF * PP;
#pragma unroll
        for (int c = 0; c < 2; c++) {
            const int j = jrow(t, c);
            if (!(actx && j < g.ny - 1)) continue;
            St<NF> s;
            s.pl = cur + o[c];
#pragma unroll
            for (int f = 0; f < NF; f++) {
                s.zc[f] = s.pl[f * PP];
                // nz = 1: the reference's z neighbours are the cell itself (sz = 0 / k wraps onto k), so a
                // z-difference is (c - c) * 0 -- zero for finite c but NaN for a non-finite one
                s.zm[f] = s.zp[f] = s.zc[f];
            }
            op.cell(s, pw[c], (long long)j * g.sy + i, i, j, 0);
        }
        if (PWARP) {
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[slot]);
        } else {
            __syncthreads();  // tile t is consumed: its stage may be refilled
            if (tid == 0) {
                fence_proxy_async();
                while (next_q < nt && next_q - ST <= t) issue(next_q++);
            }
        }
    }
}

template <bool D3, class Op>
__device__ __forceinline__ void tma_direct_sweep(const TmaSet<Op::NF>& tm, const Geom& g, int zchunk, Op& op,
                                                 const HaloGate* hg = nullptr) {
    if (!D3) {
        tma_direct_sweep_2d(tm, g, zchunk, op, hg);
        return;
    }
    constexpr int NF = Op::NF, NPW = Op::NPW, ST = DirectPipe<NF>::ST;
    constexpr int HX = Tile::HX, PP = Tile::PLANE_PAD;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    double* ring = reinterpret_cast<double*>(smem_raw);                          // [ST][NF][PP]
    uint64_t* full = reinterpret_cast<uint64_t*>(ring + DirectPipe<NF>::RING);  // [ST]
    const int tid = threadIdx.x;
    const int x0 = Tile::X0(blockIdx.x), y0 = blockIdx.y * Tile::BY;  // box origin
    const int kb = g.k0 + gate_chunk(hg, blockIdx.z, gridDim.z) * zchunk, ke = min(kb + zchunk, g.k1);
    const int np = ke - kb + 2;  // planes kb-1 .. ke
    const int pz0 = kb - 1;
    uint64_t* empty = full + ST;
    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < ST; s++) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], Tile::WARPS);
        }
        mbar_fence_init();
    }
    __syncthreads();
    halo_gate_pull(hg, g, x0, y0, kb == g.k0, ke == g.k1, true);
    constexpr bool PWARP = Op::PRODUCER_WARP;  // false: 512 threads, thread 0 refills after a block barrier
    int next_q = 0;
    auto issue = [&](int q) {
        const int slot = q % ST;
        mbar_expect_tx(&full[slot], NF * Tile::PLANE_BYTES);
#pragma unroll
        for (int f = 0; f < NF; f++)
            tma_load_plane(ring + (slot * NF + f) * PP, &tm.m[f], x0, y0, pz0 + q, &full[slot]);
    };
    if (PWARP) {
        if (tid >= Tile::THREADS) {  // producer warp
            if (tid == Tile::THREADS) {
                for (int q = 0; q < np; q++) {
                    if (q >= ST) mbar_wait(&empty[q % ST], ((q / ST) - 1) & 1);  // all 16 warps left its previous plane
                    issue(q);
                }
            }
            return;
        }
    } else if (tid == 0) {
        for (; next_q < ST && next_q < np; next_q++) issue(next_q);
    }
    const int lane = tid & 31;
    const int lx = tid & (Tile::BX - 1), ly = tid >> 6;
    const int i = x0 + Tile::XO + lx;
    int o[2], j[2];
    bool act[2];
    long long n[2];
#pragma unroll
    for (int c = 0; c < 2; c++) {
        o[c] = Tile::OWN0(lx, ly + 8 * c);
        j[c] = y0 + 1 + ly + 8 * c;
        act[c] = (i >= 1) && (i < g.nx - 1) && (j[c] < g.ny - 1);
        n[c] = (long long)kb * g.sz + (long long)j[c] * g.sy + i;
    }
    double zm[2][NF], zc[2][NF];  // z-1 and centre values of the own cells, carried in registers
    double pwn[2][NPW];
#pragma unroll
    for (int c = 0; c < 2; c++)
        if (act[c]) {
            op.pw_load(n[c], pwn[c]);
            if constexpr (Op::STAGED) op.prime(n[c], c);  // asynchronous staging of further pointwise operands
        }
    auto wait_q = [&](int q) -> const double* {
        const int slot = q % ST;
        mbar_wait(&full[slot], (q / ST) & 1);
        return ring + slot * NF * PP;
    };
    const double* p0 = wait_q(0);
    const double* cur = wait_q(1);  // shared plane k
#pragma unroll
    for (int c = 0; c < 2; c++)
#pragma unroll
        for (int f = 0; f < NF; f++) {
            zm[c][f] = p0[f * PP + o[c]];
            zc[c][f] = cur[f * PP + o[c]];
        }
    if (PWARP) {
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[0]);  // plane kb-1 only fed the registers
    }
    for (int k = kb; k < ke; k++) {
        const int q = k - kb + 2;  // raw plane k+1
        double pw[2][NPW];
#pragma unroll
        for (int c = 0; c < 2; c++) {
#pragma unroll
            for (int v = 0; v < NPW; v++) pw[c][v] = pwn[c][v];
            if (act[c] && k + 1 < ke) op.pw_load(n[c] + g.sz, pwn[c]);
        }
        const double* nxt = wait_q(q);
#pragma unroll
        for (int c = 0; c < 2; c++) {
            St<NF> s;
            s.pl = cur + o[c];
#pragma unroll
            for (int f = 0; f < NF; f++) {
                s.zm[f] = zm[c][f];
                s.zc[f] = zc[c][f];
                s.zp[f] = nxt[f * PP + o[c]];
                zm[c][f] = s.zc[f];
                zc[c][f] = s.zp[f];
            }
            if (act[c]) op.cell(s, pw[c], n[c], i, j[c], k);
            n[c] += g.sz;
        }
        if (PWARP) {
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[(q - 1) % ST]);  // this warp is done with shared plane k
        } else {
            __syncthreads();  // every thread is done with shared plane k (stage of q-1) and older
            if (tid == 0) {
                fence_proxy_async();
                while (next_q < np && next_q - ST <= q - 1) issue(next_q++);
            }
        }
        cur = nxt;
    }
}

template <class K>
static inline void tma_optin(K kernel, size_t smem) {
    cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
}

#endif  // __CUDACC__

}  // namespace cfdb
