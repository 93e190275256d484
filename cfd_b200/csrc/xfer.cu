// xfer.cu -- host <-> device movement of whole fields for the drop-in entry points.
//
// The reference hands us flow_field arrays from its own allocator (cfd_aligned_malloc: pageable
// memory, cpu/solver_explicit_euler.c:79-110).  A cudaMemcpy from pageable memory is staged by the
// driver through one pinned buffer on one thread (~15 GB/s measured on the B200 box, a quarter of
// what PCIe 5 delivers), and at 512^3 and up the transfers ARE the step time of a stateless
// solve_*_method_gpu call.  Here pinned / registered host memory still goes straight to the copy
// engine; pageable memory goes through a pipeline of pinned slots fed by several host threads
// (memcpy into a slot | async copy of the previous slot), each thread on its own stream.
#include "session.h"

#include <atomic>
#include <mutex>
#include <stdlib.h>
#include <string.h>
#include <thread>
#include <vector>

namespace cfdb {

namespace {

constexpr size_t SLOT_BYTES = 16u << 20;     // one staging slot
constexpr size_t STAGE_MIN = 32u << 20;      // smaller pageable copies are left to the driver
constexpr int MAX_THREADS = 8;

struct Lane {  // one host thread: a stream and two pinned slots
    cudaStream_t st = nullptr;
    void* slot[2] = {nullptr, nullptr};
    cudaEvent_t ev[2] = {nullptr, nullptr};
};

int lane_count() {
    static const int n = [] {
        if (const char* e = getenv("CFD_B200_XFER_THREADS")) return atoi(e);  // 0: never stage
        const unsigned hc = std::thread::hardware_concurrency();
        int t = (int)(hc / 2);
        if (t < 2) t = 2;
        if (t > MAX_THREADS) t = MAX_THREADS;
        return t;
    }();
    return n < 0 ? 0 : (n > MAX_THREADS ? MAX_THREADS : n);
}

bool is_pageable(const void* host) {
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, host) != cudaSuccess) {
        cudaGetLastError();
        return true;  // older runtimes report unregistered memory as an error
    }
    return at.type == cudaMemoryTypeUnregistered;
}

}  // namespace

struct Stager {
    std::vector<Lane> lanes;
    ~Stager() {
        for (Lane& l : lanes) {
            for (int q = 0; q < 2; q++) {
                if (l.ev[q]) cudaEventDestroy(l.ev[q]);
                if (l.slot[q]) cudaFreeHost(l.slot[q]);
            }
            if (l.st) cudaStreamDestroy(l.st);
        }
    }
    bool ensure(int n) {
        while ((int)lanes.size() < n) {
            Lane l;
            bool ok = cudaStreamCreateWithFlags(&l.st, cudaStreamNonBlocking) == cudaSuccess;
            for (int q = 0; q < 2 && ok; q++)
                ok = cudaHostAlloc(&l.slot[q], SLOT_BYTES, cudaHostAllocPortable) == cudaSuccess &&
                     cudaEventCreateWithFlags(&l.ev[q], cudaEventDisableTiming) == cudaSuccess;
            lanes.push_back(l);  // (a partially built lane is still released by the destructor)
            if (!ok) {
                cudaGetLastError();
                return false;
            }
        }
        return true;
    }
};

// One staging pipeline per device for the whole process (256 MB of pinned memory, allocated on the first
// large pageable transfer and kept): sessions come and go -- every Poisson solver object owns one -- and
// pinning memory costs more than the transfer it would speed up.
static std::mutex g_stage_mu;
static Stager* g_stager[64] = {};

struct Chunk {
    char* dev;
    char* host;
    size_t bytes;
};

// Move every segment; to_device selects the direction.  The caller has drained the session stream.
static cfd_status_t staged(cfd_b200_session* s, const std::vector<Chunk>& chunks, bool to_device, int nl) {
    std::lock_guard<std::mutex> lk(g_stage_mu);
    if (s->dev < 0 || s->dev >= 64) return CFD_ERROR_INVALID;
    if (!g_stager[s->dev]) g_stager[s->dev] = new Stager();
    Stager* const stg = g_stager[s->dev];
    if (!stg->ensure(nl)) {
        set_error(CFD_ERROR_NOMEM, "pinned staging buffers for host transfers could not be allocated");
        return CFD_ERROR_NOMEM;
    }
    std::atomic<int> err{(int)cudaSuccess};
    const int dev = s->dev;
    auto work = [&](int t) {
        Lane& l = stg->lanes[t];
        cudaError_t e = cudaSetDevice(dev);
        int k = 0, prev = -1;  // prev: chunk whose device -> slot copy is in flight (download)
        for (size_t c = t; c < chunks.size() && e == cudaSuccess; c += (size_t)nl, k++) {
            const Chunk& ch = chunks[c];
            const int q = k & 1;
            if (to_device) {
                if (k >= 2) e = cudaEventSynchronize(l.ev[q]);  // the copy that last read this slot is done
                if (e != cudaSuccess) break;
                memcpy(l.slot[q], ch.host, ch.bytes);
                e = cudaMemcpyAsync(ch.dev, l.slot[q], ch.bytes, cudaMemcpyHostToDevice, l.st);
                if (e == cudaSuccess) e = cudaEventRecord(l.ev[q], l.st);
            } else {
                e = cudaMemcpyAsync(l.slot[q], ch.dev, ch.bytes, cudaMemcpyDeviceToHost, l.st);
                if (e == cudaSuccess) e = cudaEventRecord(l.ev[q], l.st);
                if (prev >= 0 && e == cudaSuccess) {  // drain the previous slot while this copy runs
                    e = cudaEventSynchronize(l.ev[q ^ 1]);
                    if (e == cudaSuccess) memcpy(chunks[prev].host, l.slot[q ^ 1], chunks[prev].bytes);
                }
                prev = (int)c;
            }
        }
        if (e == cudaSuccess) e = cudaStreamSynchronize(l.st);
        if (!to_device && prev >= 0 && e == cudaSuccess)
            memcpy(chunks[prev].host, l.slot[(k - 1) & 1], chunks[prev].bytes);
        if (e != cudaSuccess) err.store((int)e);
    };
    std::vector<std::thread> th;
    for (int t = 1; t < nl; t++) th.emplace_back(work, t);
    work(0);
    for (std::thread& x : th) x.join();
    if (err.load() != (int)cudaSuccess) {
        set_error(CFD_ERROR, "staged host transfer failed: %s", cudaGetErrorString((cudaError_t)err.load()));
        return CFD_ERROR;
    }
    return CFD_SUCCESS;
}

// n segments of `bytes` each.  Synchronous: the data has arrived when this returns.
cfd_status_t xfer_fields(cfd_b200_session* s, double* const* dev, double* const* host, int n, size_t bytes,
                         bool to_device) {
    CFDB_CUDA_OK(cudaSetDevice(s->dev));
    const int nl = lane_count();
    std::vector<Chunk> chunks;
    for (int q = 0; q < n; q++) {
        if (!dev[q] || !host[q]) continue;
        if (nl >= 1 && bytes >= STAGE_MIN && is_pageable(host[q])) {
            for (size_t off = 0; off < bytes; off += SLOT_BYTES)
                chunks.push_back({(char*)dev[q] + off, (char*)host[q] + off, bytes - off < SLOT_BYTES ? bytes - off : SLOT_BYTES});
        } else if (to_device) {
            CFDB_CUDA_OK(cudaMemcpyAsync(dev[q], host[q], bytes, cudaMemcpyHostToDevice, s->st));
        } else {
            CFDB_CUDA_OK(cudaMemcpyAsync(host[q], dev[q], bytes, cudaMemcpyDeviceToHost, s->st));
        }
    }
    CFDB_CUDA_OK(cudaStreamSynchronize(s->st));  // direct copies done; nothing on the session stream touches the buffers
    if (chunks.empty()) return CFD_SUCCESS;
    return staged(s, chunks, to_device, nl);
}

}  // namespace cfdb
