// tma_host.cu -- host side of the TMA path: tensor-map creation (cuTensorMapEncodeTiled, reached
// through cudaGetDriverEntryPoint so the library does not link libcuda) and a per-session cache
// keyed by device pointer (the arrays of a session are few and long-lived).
#include "kernels.h"
#include "tma.cuh"

#include <cudaTypedefs.h>
#include <stdlib.h>

namespace cfdb {

static PFN_cuTensorMapEncodeTiled encode_fn() {
    static PFN_cuTensorMapEncodeTiled fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = (PFN_cuTensorMapEncodeTiled)p;
        else
            cudaGetLastError();
    }
    return fn;
}

static const CUtensorMap* lookup(TmaCache& tc, std::unordered_map<const double*, CUtensorMap>& maps, const double* p,
                                 int bx, int by);

void TmaCache::configure(const Geom& g) {
    nx = g.nx; ny = g.ny; nz = g.nz;
    maps.clear();
    wide_maps.clear();
    // TMA needs 16-byte global strides: nx even (the odd-nx grids, e.g. 129^2, take the direct kernels)
    usable = (nx % 2 == 0) && nx >= 4 && encode_fn() != nullptr && !getenv("CFD_B200_NO_TMA");
}

const CUtensorMap* TmaCache::get(const double* p) { return lookup(*this, maps, p, Tile::HX, Tile::HY); }
const CUtensorMap* TmaCache::get_wide(const double* p) {
    return lookup(*this, wide_maps, p, WideTile::WX, WideTile::WY);
}

static const CUtensorMap* lookup(TmaCache& tc, std::unordered_map<const double*, CUtensorMap>& maps, const double* p,
                                 int bx, int by) {
    const int nx = tc.nx, ny = tc.ny, nz = tc.nz;
    bool& usable = tc.usable;
    if (!usable || !p) return nullptr;
    auto it = maps.find(p);
    if (it != maps.end()) return &it->second;
    if (((uintptr_t)p & 15) != 0) return nullptr;
    CUtensorMap m;
    const cuuint64_t dims[3] = {(cuuint64_t)nx, (cuuint64_t)ny, (cuuint64_t)nz};
    const cuuint64_t strides[2] = {(cuuint64_t)nx * 8, (cuuint64_t)nx * ny * 8};
    const cuuint32_t box[3] = {(cuuint32_t)bx, (cuuint32_t)by, 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    // L2 promotion of the box rows (544 bytes starting 16 bytes before a 512-byte boundary): 256-byte lines by
    // default; CFD_B200_TMA_L2PROMO=128|64|0 for experiments
    static const CUtensorMapL2promotion promo = [] {
        const char* e = getenv("CFD_B200_TMA_L2PROMO");
        const int v = e ? atoi(e) : 256;
        return v == 128 ? CU_TENSOR_MAP_L2_PROMOTION_L2_128B
                        : (v == 64 ? CU_TENSOR_MAP_L2_PROMOTION_L2_64B
                                   : (v == 0 ? CU_TENSOR_MAP_L2_PROMOTION_NONE : CU_TENSOR_MAP_L2_PROMOTION_L2_256B));
    }();
    CUresult r = encode_fn()(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, (void*)p, dims, strides, box, estr,
                             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, promo,
                             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        usable = false;  // never retry: every later launch takes the direct kernels
        return nullptr;
    }
    return &maps.emplace(p, m).first->second;
}

}  // namespace cfdb
