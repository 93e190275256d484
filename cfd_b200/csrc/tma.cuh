// tma.cuh -- Blackwell TMA (cp.async.bulk.tensor) + mbarrier plumbing of the plane-pipelined
// stencil kernels (sm_100a).
//
// Every stencil operand is described to the TMA unit as a rank-3 FP64 tensor {nx, ny, nz}
// (x fastest, the reference layout of cfd/core/indexing.h:17-20).  A block owns a BX x BY tile
// of interior cells and marches along z; for each plane it needs the tile plus a one-cell halo,
// i.e. a {BX+2, BY+2, 1} box, which one elected thread requests with a single instruction per
// operand.  Out-of-bounds parts of a box (tiles on the domain edge) are zero-filled by the
// hardware, so the kernels carry no edge branches on the load side.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace cfdb {

// Where a tile's own cells start.  I0 = 0 (default): tile b owns the columns i = b*BX .. b*BX+BX-1, so the 256 bytes a
// warp stores per row start on a 256-byte boundary (the thread on the boundary column i = 0 idles).  A write-only
// stream whose warps start ONE element off that boundary runs at 0.6 of the aligned rate on this GPU
// (tools/copy_probe.cu, profiles/r2z_copy_probe.jsonl: 3.9 against 6.3 TB/s for the tile-row pattern of these
// kernels).  The price: the box has to start on an even column (TMA coordinate rule, below), i.e. two columns left
// of the own cells, and is one column wider than the stencil needs.  I0 = 1 is the round-1 geometry (own cells
// start at b*BX + 1, the first interior column; box = tile + one halo column each side), kept for A/B builds.
#ifndef CFDB_TILE_I0
#define CFDB_TILE_I0 0
#endif
struct Tile {
    static constexpr int BX = 64, BY = 16;            // interior cells per block and plane
    static constexpr int I0 = CFDB_TILE_I0;           // own column 0 of tile b is global column b*BX + I0
    static constexpr int XO = 2 - I0;                 // ... and sits in column XO of the box
    static constexpr int HX = BX + 2 * XO, HY = BY + 2;  // box with halo
    static constexpr int PLANE = HX * HY;             // doubles per operand plane
    static constexpr int PLANE_PAD = (PLANE + 15) / 16 * 16;  // padded so every plane starts 128-B aligned
    static constexpr int PLANE_BYTES = PLANE * 8;     // bytes the TMA delivers per operand plane
    static constexpr int THREADS = 512;               // 2 cells per thread: rows ly and ly + 8
    static constexpr int WARPS = THREADS / 32;
    static constexpr int THREADS_PW = THREADS + 32;   // + one producer warp (stencil_tma.cuh)
    __host__ __device__ static constexpr int OWN0(int lx, int ly) { return (ly + 1) * HX + lx + XO; }
    __host__ __device__ static constexpr int X0(int b) { return b * BX + I0 - XO; }   // box origin (even)
    __host__ __device__ static constexpr int GX(int nx) { return (nx - 1 - I0 + BX - 1) / BX; }  // tiles along x
    __host__ __device__ static constexpr int GY(int ny) { return (ny - 2 + BY - 1) / BY; }
};
static_assert((Tile::HX * 8) % 16 == 0, "TMA box rows must be a multiple of 16 bytes");
static_assert((Tile::PLANE_PAD * 8) % 128 == 0, "TMA destinations must stay 128-byte aligned");
static_assert((Tile::I0 - Tile::XO) % 2 == 0, "box origins must be even columns");

// Box with a TWO-cell halo, for sweeps that apply two dependent stencil stages per pass
// (poisson_tma3.inc): stage 1 is evaluated on the tile plus a one-cell ring, stage 2 on the tile.
// The innermost start coordinate of a TMA box must be 16-byte aligned (FP64: EVEN -- an odd x
// start raises "illegal instruction" on sm_100a, measured; negative starts are fine), so the box starts at the
// even column b*BX - 2: XOFF = 2 + I0 cells left of the first own cell.
struct WideTile {
    static constexpr int XOFF = 2 + Tile::I0, YOFF = 2;   // own cell (lx, ly) sits at (ly + YOFF) * WX + lx + XOFF
    static constexpr int WX = Tile::BX + 2 * XOFF, WY = Tile::BY + 4;
    static constexpr int PLANE = WX * WY;
    static constexpr int PLANE_PAD = (PLANE + 15) / 16 * 16;
    static constexpr int PLANE_BYTES = PLANE * 8;
    __host__ __device__ static constexpr int X0(int b) { return b * Tile::BX + Tile::I0 - XOFF; }
};
static_assert((WideTile::WX * 8) % 16 == 0 && (WideTile::PLANE_PAD * 8) % 128 == 0, "wide box alignment");

// Up to 4 tensor maps by value in kernel parameter space (must be __grid_constant__).
template <int N>
struct TmaSet {
    CUtensorMap m[N];
};

#ifdef __CUDACC__

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
// make the barrier initialisation visible to the async (TMA) proxy
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// generic-proxy reads of a stage are finished (ordered by the preceding __syncthreads); order
// them before the async-proxy write that refills the stage
__device__ __forceinline__ void fence_proxy_async() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
// generic-proxy WRITES to global memory (halo layers copied out of a mailbox) before TMA loads of the same addresses
__device__ __forceinline__ void fence_proxy_async_global() {
    asm volatile("fence.proxy.async.global;" ::: "memory");
}
// one {HX, HY, 1} box of `map` at (x0, y0, z) -> smem dst; completion counted in bytes on bar
__device__ __forceinline__ void tma_load_plane(void* dst, const CUtensorMap* map, int x0, int y0, int z,
                                               uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes "
        "[%0], [%1, {%2, %3, %4}], [%5];" ::"r"(smem_u32(dst)),
        "l"(map), "r"(x0), "r"(y0), "r"(z), "r"(smem_u32(bar))
        : "memory");
}
// 8-byte asynchronous global -> shared copy (LDGSTS): no destination register, so a thread can keep
// many loads in flight without register pressure; completion per thread via commit / wait groups.
__device__ __forceinline__ void cp_async8(void* dst, const void* src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}

#endif  // __CUDACC__

}  // namespace cfdb
