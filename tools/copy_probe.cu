// copy_probe.cu -- what a 1:1 read/write stream reaches on this GPU as a function of the access width.
//
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o gpurun_out/copy_probe tools/copy_probe.cu
//   gpurun_out/copy_probe [cells]            (default 2^30 doubles per array)
//
// Background: the kernels that write half of their bytes (copy, krylov_init, bicg_s_matvec) sit at 0.82-0.85 of the
// measured copy peak (MEASURED_PEAKS.json: torch copy_, 16-byte vector accesses) while read-heavy ones reach 0.95-1.0.
// Variants: 8-byte accesses (what the library's pointwise stores are), 16-byte, 32-byte (two 16-byte), 8-byte with
// the interior-cell misalignment (start at element 1), and a TMA bulk copy through shared memory.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define OK(x)                                                                      \
    do {                                                                           \
        cudaError_t e_ = (x);                                                      \
        if (e_ != cudaSuccess) {                                                   \
            printf("%s failed: %s\n", #x, cudaGetErrorString(e_));                 \
            return 1;                                                              \
        }                                                                          \
    } while (0)

__global__ void copy8(double* __restrict__ d, const double* __restrict__ s, size_t n) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += stride) d[t] = s[t];
}
template <int U>
__global__ void copy8u(double* __restrict__ d, const double* __restrict__ s, size_t n) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; t + (U - 1) * stride < n; t += U * stride) {
        double v[U];
#pragma unroll
        for (int u = 0; u < U; u++) v[u] = s[t + u * stride];
#pragma unroll
        for (int u = 0; u < U; u++) d[t + u * stride] = v[u];
    }
    for (; t < n; t += stride) d[t] = s[t];
}
template <int U>
__global__ void copy16u(double2* __restrict__ d, const double2* __restrict__ s, size_t n2) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; t + (U - 1) * stride < n2; t += U * stride) {
        double2 v[U];
#pragma unroll
        for (int u = 0; u < U; u++) v[u] = s[t + u * stride];
#pragma unroll
        for (int u = 0; u < U; u++) d[t + u * stride] = v[u];
    }
    for (; t < n2; t += stride) d[t] = s[t];
}
// 32 bytes per thread and access group: two adjacent 16-byte vectors
template <int U>
__global__ void copy32u(double2* __restrict__ d, const double2* __restrict__ s, size_t n2) {
    const size_t stride = (size_t)gridDim.x * blockDim.x * 2;
    size_t t = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 2;
    for (; t + (U - 1) * stride + 1 < n2; t += U * stride) {
        double2 v[U][2];
#pragma unroll
        for (int u = 0; u < U; u++) {
            v[u][0] = s[t + u * stride];
            v[u][1] = s[t + u * stride + 1];
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            d[t + u * stride] = v[u][0];
            d[t + u * stride + 1] = v[u][1];
        }
    }
}
// block-contiguous: a block streams its own contiguous range (the layout of the library's column / row kernels)
template <int U>
__global__ void copy8_blockrange(double* __restrict__ d, const double* __restrict__ s, size_t n) {
    const size_t per = (n + gridDim.x - 1) / gridDim.x;
    const size_t b0 = (size_t)blockIdx.x * per, b1 = b0 + per < n ? b0 + per : n;
    size_t t = b0 + threadIdx.x;
    for (; t + (U - 1) * blockDim.x < b1; t += U * blockDim.x) {
        double v[U];
#pragma unroll
        for (int u = 0; u < U; u++) v[u] = s[t + u * blockDim.x];
#pragma unroll
        for (int u = 0; u < U; u++) d[t + u * blockDim.x] = v[u];
    }
    for (; t < b1; t += blockDim.x) d[t] = s[t];
}

// the stencil kernels' store pattern: a block of 512 threads owns 64 x 8 cells of a 1024-cell-wide plane per step
// (rows of 512 B at `off` elements from the line start) and marches down the rows
template <int U>
__global__ void copy8_tilerows(double* __restrict__ d, const double* __restrict__ s, size_t n, int off) {
    const size_t nrows = n / 1024;
    const int lx = threadIdx.x & 63, ly = threadIdx.x >> 6;      // 64 x 8 threads
    const int tx = blockIdx.x & 15;                              // 16 tile columns (the last one is 2 cells short)
    const size_t chunk = (nrows + (gridDim.x / 16) - 1) / (gridDim.x / 16);
    const size_t r0 = (size_t)(blockIdx.x / 16) * chunk, r1 = r0 + chunk < nrows ? r0 + chunk : nrows;
    const int i = tx * 64 + lx + off;
    if (i >= 1023) return;
    for (size_t r = r0 + ly; r + 8 * (U - 1) < r1; r += 8 * U) {
        double v[U];
#pragma unroll
        for (int u = 0; u < U; u++) v[u] = s[(r + 8 * u) * 1024 + i];
#pragma unroll
        for (int u = 0; u < U; u++) d[(r + 8 * u) * 1024 + i] = v[u];
    }
}

// TMA bulk copy: global -> shared -> global, 16 KB chunks, 4 stages, one thread drives it
__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__global__ void __launch_bounds__(32) copy_bulk(double* __restrict__ d, const double* __restrict__ s, size_t n) {
    constexpr int CH = 2048, ST = 4;  // doubles per chunk
    extern __shared__ __align__(128) unsigned char raw[];
    double* buf = reinterpret_cast<double*>(raw);
    uint64_t* full = reinterpret_cast<uint64_t*>(buf + ST * CH);
    if (threadIdx.x == 0) {
        for (int q = 0; q < ST; q++)
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&full[q])) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        const size_t nch = n / CH;
        size_t c = blockIdx.x;
        int issued = 0, done = 0;
        size_t pend[ST];
        auto load = [&](size_t ch) {
            const int q = issued % ST;
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(&full[q])), "r"(CH * 8) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                             s32(buf + q * CH)),
                         "l"(s + ch * CH), "r"(CH * 8), "r"(s32(&full[q]))
                         : "memory");
            pend[q] = ch;
            issued++;
        };
        for (; issued < ST - 1 && c < nch; c += gridDim.x) load(c);
        while (done < issued) {
            const int q = done % ST;
            const unsigned ph = (done / ST) & 1;
            asm volatile(
                "{\n.reg .pred p;\nW_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D_%=;\nbra W_%=;\nD_%=:\n}\n" ::"r"(
                    s32(&full[q])),
                "r"(ph)
                : "memory");
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(d + pend[q] * CH), "r"(s32(buf + q * CH)),
                         "r"(CH * 8)
                         : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            done++;
            if (c < nch) {
                // the stage refilled next was stored ST-1 groups ago: wait until its reads are done
                asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(ST - 2) : "memory");
                load(c);
                c += gridDim.x;
            }
        }
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
}

int main(int argc, char** argv) {
    const size_t n = argc > 1 ? (size_t)atoll(argv[1]) : ((size_t)1 << 30);
    double *a, *b;
    OK(cudaMalloc(&a, (n + 16) * 8));
    OK(cudaMalloc(&b, (n + 16) * 8));
    OK(cudaMemset(a, 1, (n + 16) * 8));
    OK(cudaMemset(b, 0, (n + 16) * 8));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int blocks = 148 * 8;
    auto timeit = [&](const char* name, auto launch) {
        float best = 1e30f;
        for (int rep = 0; rep < 6; rep++) {
            cudaEventRecord(e0);
            launch();
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            if (rep > 0 && ms < best) best = ms;
        }
        const cudaError_t e = cudaGetLastError();
        printf("{\"variant\": \"%s\", \"ms\": %.4f, \"GBps\": %.1f, \"err\": \"%s\"}\n", name, best, 16.0 * n / best / 1e6,
               e == cudaSuccess ? "" : cudaGetErrorString(e));
    };
    timeit("cudaMemcpyAsync d2d", [&] { cudaMemcpyAsync(b, a, n * 8, cudaMemcpyDeviceToDevice); });
    timeit("8B grid-stride (library k_copy)", [&] { copy8<<<blocks, 256>>>(b, a, n); });
    timeit("8B grid-stride, 4 loads before 4 stores", [&] { copy8u<4><<<blocks, 256>>>(b, a, n); });
    timeit("8B grid-stride, 8 loads before 8 stores", [&] { copy8u<8><<<blocks, 256>>>(b, a, n); });
    timeit("8B grid-stride x4, 32 blocks per SM", [&] { copy8u<4><<<148 * 32, 256>>>(b, a, n); });
    timeit("8B misaligned by one element x4", [&] { copy8u<4><<<blocks, 256>>>(b + 1, a + 1, n); });
    timeit("8B x4: aligned loads, stores misaligned by one element", [&] { copy8u<4><<<blocks, 256>>>(b + 1, a, n); });
    timeit("8B x4: loads misaligned by one element, aligned stores", [&] { copy8u<4><<<blocks, 256>>>(b, a + 1, n); });
    timeit("8B x4 misaligned both, 32 blocks per SM", [&] { copy8u<4><<<148 * 32, 256>>>(b + 1, a + 1, n); });
    timeit("8B x4 misaligned by two elements (16 B)", [&] { copy8u<4><<<blocks, 256>>>(b + 2, a + 2, n); });
    timeit("8B x4 misaligned by four elements (32 B)", [&] { copy8u<4><<<blocks, 256>>>(b + 4, a + 4, n); });
    timeit("8B tile rows: 64-cell rows at element 1 of 1024-cell lines, x4", [&] { copy8_tilerows<4><<<148 * 16, 512>>>(b, a, n, 1); });
    timeit("8B tile rows: 64-cell rows at element 0 of 1024-cell lines, x4", [&] { copy8_tilerows<4><<<148 * 16, 512>>>(b, a, n, 0); });
    timeit("8B block-contiguous ranges x4", [&] { copy8_blockrange<4><<<148 * 64, 256>>>(b, a, n); });
    timeit("16B x4", [&] { copy16u<4><<<blocks, 256>>>((double2*)b, (const double2*)a, n / 2); });
    timeit("16B x2", [&] { copy16u<2><<<blocks, 256>>>((double2*)b, (const double2*)a, n / 2); });
    timeit("2 x 16B adjacent x2", [&] { copy32u<2><<<blocks, 256>>>((double2*)b, (const double2*)a, n / 2); });
    {
        const int sm = 4 * 2048 * 8 + 64;
        cudaFuncSetAttribute(copy_bulk, cudaFuncAttributeMaxDynamicSharedMemorySize, sm);
        timeit("TMA bulk 16 KB chunks, 4 stages, 8 blocks/SM", [&] { copy_bulk<<<148 * 8, 32, sm>>>(b, a, n); });
        timeit("TMA bulk 16 KB chunks, 4 stages, 3 blocks/SM", [&] { copy_bulk<<<148 * 3, 32, sm>>>(b, a, n); });
    }
    return 0;
}
