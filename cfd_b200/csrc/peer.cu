// peer.cu -- halo exchange and small all-reduces over NVLink PEER MEMORY, no communication library.
//
// SURVEY 8e names the transport ("NCCL send/recv or P2P ... peer-mapped direct stores -- all peers are uniform
// over NVSwitch"); this is the P2P one.  Every rank owns a box (PeerBoxHeader + halo mailboxes) that the others
// map: cudaDeviceEnablePeerAccess for the slabs of one process (gang.cu), cudaIpcOpenMemHandle for one process
// per GPU (cfd_b200_session_peer_export / _attach).  Nothing here calls the host during a solve:
//
//   halo exchange #n  = k_halo_push  (my edge layers -> the neighbours' mailboxes, buffer n&1, coalesced 16-byte
//                       stores over NVLink; the block that finishes last publishes sequence number n with a
//                       system-scope release)
//                     + k_halo_pull  (wait for the neighbours' n, mailbox -> my halo layers)
//   all-reduce   #m   = k_allreduce_peer, one block: my <= 8 words into slot [me] of EVERY box, then flag m;
//                       wait for the R flags in my own box; combine the R slots in RANK ORDER -- every rank adds
//                       the same numbers in the same order, so all ranks hold the same bits and take the same
//                       stopping decisions.
//
// Two buffers suffice without any acknowledgement: a neighbour can only issue exchange n+2 after it has pulled
// n+1, which it can only do after I pushed n+1, which my stream orders after my pull of n (same argument for the
// reductions).  Waits are bounded (~4 s of spinning): a lost peer sets Ctl::comm_timeout instead of hanging the GPU.
#include "session.h"

#include <stdlib.h>
#include <string.h>

namespace cfdb {

namespace {

struct PushArgs {
    const double* lo_src[PEER_MAX_FIELDS];   // my first owned layer of every field
    const double* hi_src[PEER_MAX_FIELDS];   // my last owned layer
    double* lo_dst;                          // lo neighbour's HI-side mailbox (buffer n&1), or null
    double* hi_dst;                          // hi neighbour's LO-side mailbox, or null
    unsigned long long* lo_flag;             // the flags that go with them
    unsigned long long* hi_flag;
    unsigned* ticket;
    unsigned long long seq;
    long long cnt;                           // doubles per layer
    int nf;
};

__global__ void __launch_bounds__(256) k_halo_push(PushArgs a) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long t0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const bool vec = (a.cnt & 1) == 0;  // layers of an even number of doubles start 16-byte aligned (cudaMalloc, even nx)
    for (int side = 0; side < 2; side++) {
        double* dst = side ? a.hi_dst : a.lo_dst;
        if (!dst) continue;
        for (int q = 0; q < a.nf; q++) {
            const double* src = side ? a.hi_src[q] : a.lo_src[q];
            double* d = dst + (long long)q * a.cnt;
            if (vec && ((((size_t)src) | ((size_t)d)) & 15) == 0) {
                const double2* s2 = reinterpret_cast<const double2*>(src);
                double2* d2 = reinterpret_cast<double2*>(d);
                for (long long t = t0; t < a.cnt / 2; t += stride) d2[t] = s2[t];
            } else {
                for (long long t = t0; t < a.cnt; t += stride) d[t] = src[t];
            }
        }
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned done = atomicInc(a.ticket, gridDim.x - 1);  // wraps to 0 after the last block
        if (done == gridDim.x - 1) {
            __threadfence_system();
            if (a.lo_flag) st_release_sys(a.lo_flag, a.seq);
            if (a.hi_flag) st_release_sys(a.hi_flag, a.seq);
        }
    }
}

struct PullArgs {
    double* lo_halo[PEER_MAX_FIELDS];        // my layer 0 of every field
    double* hi_halo[PEER_MAX_FIELDS];        // my layer n-1
    const double* lo_box;                    // my LO-side mailbox (buffer n&1), or null when there is no lo neighbour
    const double* hi_box;
    const unsigned long long* lo_flag;
    const unsigned long long* hi_flag;
    int* timeout_flag;
    unsigned long long seq;
    long long cnt;
    int nf;
};

__global__ void __launch_bounds__(256) k_halo_pull(PullArgs a) {
    __shared__ int ok;
    if (threadIdx.x == 0) {
        bool good = true;
        if (a.lo_box) good = wait_flag(a.lo_flag, a.seq, a.timeout_flag, 1) && good;
        if (a.hi_box) good = wait_flag(a.hi_flag, a.seq, a.timeout_flag, 2) && good;
        ok = good;
    }
    __syncthreads();
    if (!ok) return;
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long t0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const bool vec = (a.cnt & 1) == 0;
    for (int side = 0; side < 2; side++) {
        const double* box = side ? a.hi_box : a.lo_box;
        if (!box) continue;
        for (int q = 0; q < a.nf; q++) {
            double* d = side ? a.hi_halo[q] : a.lo_halo[q];
            const double* src = box + (long long)q * a.cnt;
            if (vec && ((((size_t)src) | ((size_t)d)) & 15) == 0) {
                const double2* s2 = reinterpret_cast<const double2*>(src);
                double2* d2 = reinterpret_cast<double2*>(d);
                for (long long t = t0; t < a.cnt / 2; t += stride) {
                    double2 v;  // the mailbox was written by another GPU: read it past L1
                    asm volatile("ld.global.cg.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(s2 + t));
                    d2[t] = v;
                }
            } else {
                for (long long t = t0; t < a.cnt; t += stride) d[t] = __ldcg(src + t);
            }
        }
    }
}

struct ReduceArgs {
    char* box[PEER_MAX_RANKS];   // every rank's box as mapped on this device
    void* vals;                  // in place: double[n] or int[n]
    int* timeout_flag;
    unsigned long long seq;
    int R, r, n, op;             // op 0: sum f64, 1: max f64, 2: max i32
};

__global__ void __launch_bounds__(32) k_allreduce_peer(ReduceArgs a) {
    const int t = threadIdx.x;
    const int buf = (int)(a.seq & 1);
    __shared__ double mine[PEER_RED_WORDS];
    __shared__ int ok;
    if (t == 0) ok = 1;
    if (t < a.n) mine[t] = (a.op == 2) ? (double)((const int*)a.vals)[t] : ((const double*)a.vals)[t];
    __syncwarp();
    if (t < a.R) {  // lane t serves peer t: payload, fence, flag
        PeerBoxHeader* h = reinterpret_cast<PeerBoxHeader*>(a.box[t]);
        for (int j = 0; j < a.n; j++) h->rval[buf][a.r][j] = mine[j];
        __threadfence_system();
        st_release_sys(&h->rflag[a.r][buf], a.seq);
    }
    PeerBoxHeader* me = reinterpret_cast<PeerBoxHeader*>(a.box[a.r]);
    if (t < a.R && !wait_flag(&me->rflag[t][buf], a.seq, a.timeout_flag, 16 + t)) ok = 0;
    __syncwarp();
    if (!ok) return;
    if (t < a.n) {
        const volatile double* v = &me->rval[buf][0][t];
        double acc = v[0];
        for (int q = 1; q < a.R; q++) {  // rank order, the same on every rank
            const double x = v[(size_t)q * PEER_RED_WORDS];
            if (a.op == 0) acc = acc + x;
            else if (x > acc) acc = x;  // the kernels' own running maximum: `if (r > m) m = r`
        }
        if (a.op == 2) ((int*)a.vals)[t] = (int)acc;
        else ((double*)a.vals)[t] = acc;
    }
}

// The waiting kernels ask for the same shared-memory carveout as the TMA stencil kernels (maximum shared memory):
// with several slabs on ONE GPU (virtual ranks) a stencil kernel of one slab must be able to start on an SM while
// another slab's waiting kernel is resident elsewhere; measured on B200: with the default carveout the stencil
// kernel is held back until the waiting kernel exits (SM carveout reconfiguration), i.e. until its time-out.
void peer_kernel_setup() {
    CFDB_ONCE_PER_DEVICE(
        cudaFuncSetAttribute(k_halo_push, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        cudaFuncSetAttribute(k_halo_pull, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        cudaFuncSetAttribute(k_allreduce_peer, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        cudaGetLastError());
}

size_t mailbox_offset(const Peer& p, int side, int buf) {
    return sizeof(PeerBoxHeader) + ((size_t)(side * 2 + buf) * PEER_MAX_FIELDS) * p.cnt * sizeof(double);
}

}  // namespace

cfd_status_t peer_alloc_box(cfd_b200_session* s) {
    Peer& p = s->peer;
    if (p.box) return CFD_SUCCESS;
    if (s->desc.nranks > PEER_MAX_RANKS) {
        set_error(CFD_ERROR_UNSUPPORTED, "peer transport supports up to %d ranks", PEER_MAX_RANKS);
        return CFD_ERROR_UNSUPPORTED;
    }
    CFDB_CUDA_OK(cudaSetDevice(s->dev));
    peer_kernel_setup();
    p.cnt = s->g.dim3 ? (size_t)s->g.nx * s->g.ny : (size_t)s->g.nx;
    const size_t mail_bytes = sizeof(PeerBoxHeader) + 4 * PEER_MAX_FIELDS * p.cnt * sizeof(double);
    const size_t halo2_off = (mail_bytes + 255) / 256 * 256;
    p.box_bytes = halo2_off + 2 * p.cnt * sizeof(double);
    if (cudaMalloc((void**)&p.box, p.box_bytes) != cudaSuccess) {
        cudaGetLastError();
        p.box = nullptr;
        set_error(CFD_ERROR_NOMEM, "peer transport: %zu bytes for the halo mailboxes could not be allocated", p.box_bytes);
        return CFD_ERROR_NOMEM;
    }
    CFDB_CUDA_OK(cudaMemset(p.box, 0, sizeof(PeerBoxHeader)));
    p.halo2[0] = reinterpret_cast<double*>(p.box + halo2_off);
    p.halo2[1] = p.halo2[0] + p.cnt;
    CFDB_CUDA_OK(cudaMemset(p.halo2[0], 0, 2 * p.cnt * sizeof(double)));
    s->bytes_allocated += p.box_bytes;
    p.peer_box[s->desc.rank] = p.box;
    return CFD_SUCCESS;
}

void peer_release(cfd_b200_session* s) {
    Peer& p = s->peer;
    for (int q = 0; q < PEER_MAX_RANKS; q++)
        if (p.ipc_mapped[q] && p.peer_box[q]) cudaIpcCloseMemHandle(p.peer_box[q]);
    if (p.box) cudaFree(p.box);
    p = Peer();
    s->rb.peer = nullptr;
}

cfd_status_t peer_connected(cfd_b200_session* s) {
    Peer& p = s->peer;
    CFDB_CUDA_OK(cudaSetDevice(s->dev));
    PeerBoxHeader h;
    CFDB_CUDA_OK(cudaMemcpy(&h, p.box, sizeof(h), cudaMemcpyDeviceToHost));
    for (int q = 0; q < PEER_MAX_RANKS; q++) h.boxes[q] = q < s->desc.nranks ? p.peer_box[q] : nullptr;
    h.timeout_flag = &s->ctl->comm_timeout;
    h.R = s->desc.nranks;
    h.r = s->desc.rank;
    CFDB_CUDA_OK(cudaMemcpy(p.box, &h, sizeof(h), cudaMemcpyHostToDevice));
    p.on = true;
    s->rb.peer = getenv("CFD_B200_NO_FUSED_REDUCE") ? nullptr : reinterpret_cast<PeerBoxHeader*>(p.box);
    // edge blocks that spin inside a sweep would starve the OTHER slabs' kernels when slabs share one GPU
    const char* vr = getenv("CFD_B200_VIRTUAL_RANKS");
    s->gate_enabled = !getenv("CFD_B200_NO_HALO_GATE") && !(vr && vr[0] && vr[0] != '0');
    return CFD_SUCCESS;
}

// gate != nullptr: launch the push only and describe the pull in *gate (nf <= 2): the next kernel that reads these
// fields does it in its edge blocks (stencil_tma.cuh: halo_gate_pull), or peer_pull_from_gate() does it stand-alone.
// deep2: ONE field, two layers per side (the second one lands in Peer::halo2), sent as two "fields" of one exchange.
static cfd_status_t peer_exchange_impl(cfd_b200_session* s, double* const* fields, int nf, bool ring, HaloGate* gate,
                                       bool deep2 = false);

cfd_status_t peer_halo_exchange(cfd_b200_session* s, double* const* fields, int nf, bool ring) {
    return peer_exchange_impl(s, fields, nf, ring, nullptr);
}
cfd_status_t peer_halo_exchange2(cfd_b200_session* s, double* field, HaloGate* gate) {
    if (gate) *gate = HaloGate();
    double* f[1] = {field};
    return peer_exchange_impl(s, f, 1, false, gate, true);
}
cfd_status_t peer_halo_push_gated(cfd_b200_session* s, double* const* fields, int nf, HaloGate* gate) {
    *gate = HaloGate();
    if (nf > 2) return peer_exchange_impl(s, fields, nf, false, nullptr);
    return peer_exchange_impl(s, fields, nf, false, gate);
}
void peer_pull_from_gate(cudaStream_t st, const HaloGate& hg) {
    if (!hg.on) return;
    PullArgs pl;
    memset(&pl, 0, sizeof(pl));
    for (int q = 0; q < hg.nf; q++) {
        pl.lo_halo[q] = hg.halo[0][q];
        pl.hi_halo[q] = hg.halo[1][q];
    }
    pl.lo_box = hg.box[0];
    pl.hi_box = hg.box[1];
    pl.lo_flag = hg.flag[0];
    pl.hi_flag = hg.flag[1];
    pl.timeout_flag = hg.timeout_flag;
    pl.seq = hg.seq;
    pl.cnt = hg.cnt;
    pl.nf = hg.nf;
    const long long work = (long long)hg.nf * hg.cnt / 2;
    int blocks = (int)((work + 256 * 8 - 1) / (256 * 8));
    blocks = blocks < 1 ? 1 : (blocks > 148 ? 148 : blocks);
    CFDB_LAUNCH(KK_HALO, st, k_halo_pull<<<blocks, 256, 0, st>>>(pl));
}

static cfd_status_t peer_exchange_impl(cfd_b200_session* s, double* const* fields, int nf, bool ring, HaloGate* gate,
                                       bool deep2) {
    Peer& p = s->peer;
    const int R = s->desc.nranks, r = s->desc.rank;
    if (nf > PEER_MAX_FIELDS) {
        for (int q = 0; q < nf; q += PEER_MAX_FIELDS) {
            const cfd_status_t rc = peer_exchange_impl(s, fields + q, nf - q < PEER_MAX_FIELDS ? nf - q : PEER_MAX_FIELDS, ring, nullptr);
            if (rc != CFD_SUCCESS) return rc;
        }
        return CFD_SUCCESS;
    }
    const Geom& g = s->g;
    const long long n_ax = g.dim3 ? g.nz : g.ny;
    const long long st = (long long)p.cnt;
    int lo = r - 1, hi = r + 1;
    if (ring) {
        lo = (r + R - 1) % R;
        hi = (r + 1) % R;
    }
    const bool has_lo = lo >= 0 && lo < R, has_hi = hi >= 0 && hi < R;
    const unsigned long long seq = ++p.hseq;
    const int buf = (int)(seq & 1);
    PeerBoxHeader* mine = reinterpret_cast<PeerBoxHeader*>(p.box);
    PushArgs pa;
    memset(&pa, 0, sizeof(pa));
    PullArgs pl;
    memset(&pl, 0, sizeof(pl));
    for (int q = 0; q < nf; q++) {
        pa.lo_src[q] = fields[q] + 1 * st;
        pa.hi_src[q] = fields[q] + (n_ax - 2) * st;
        pl.lo_halo[q] = fields[q];
        pl.hi_halo[q] = fields[q] + (n_ax - 1) * st;
    }
    if (deep2) {
        if (nf != 1 || n_ax < 4 || !p.halo2[0]) return CFD_ERROR_INVALID;
        pa.lo_src[1] = fields[0] + 2 * st;
        pa.hi_src[1] = fields[0] + (n_ax - 3) * st;
        pl.lo_halo[1] = p.halo2[0];
        pl.hi_halo[1] = p.halo2[1];
        nf = 2;
    }
    if (has_lo) {  // my low edge is the lo neighbour's HIGH halo
        PeerBoxHeader* h = reinterpret_cast<PeerBoxHeader*>(p.peer_box[lo]);
        pa.lo_dst = reinterpret_cast<double*>(p.peer_box[lo] + mailbox_offset(p, 1, buf));
        pa.lo_flag = &h->hflag[1][buf];
        pl.lo_box = reinterpret_cast<const double*>(p.box + mailbox_offset(p, 0, buf));
        pl.lo_flag = &mine->hflag[0][buf];
    }
    if (has_hi) {
        PeerBoxHeader* h = reinterpret_cast<PeerBoxHeader*>(p.peer_box[hi]);
        pa.hi_dst = reinterpret_cast<double*>(p.peer_box[hi] + mailbox_offset(p, 0, buf));
        pa.hi_flag = &h->hflag[0][buf];
        pl.hi_box = reinterpret_cast<const double*>(p.box + mailbox_offset(p, 1, buf));
        pl.hi_flag = &mine->hflag[1][buf];
    }
    if (!has_lo && !has_hi) return CFD_SUCCESS;
    pa.ticket = &mine->ticket;
    pa.seq = pl.seq = seq;
    pa.cnt = pl.cnt = st;
    pa.nf = pl.nf = nf;
    pl.timeout_flag = &s->ctl->comm_timeout;
    const long long work = (long long)nf * st / 2;
    int blocks = (int)((work + 256 * 8 - 1) / (256 * 8));
    if (blocks < 1) blocks = 1;
    if (blocks > 148) blocks = 148;
    CFDB_LAUNCH(KK_HALO, s->st, k_halo_push<<<blocks, 256, 0, s->st>>>(pa));
    if (gate) {
        gate->on = 1;
        gate->nf = nf;
        gate->cnt = st;
        gate->seq = seq;
        gate->box[0] = pl.lo_box;
        gate->box[1] = pl.hi_box;
        gate->flag[0] = pl.lo_flag;
        gate->flag[1] = pl.hi_flag;
        for (int q = 0; q < nf; q++) {
            gate->halo[0][q] = pl.lo_halo[q];
            gate->halo[1][q] = pl.hi_halo[q];
        }
        gate->timeout_flag = pl.timeout_flag;
        return CFD_SUCCESS;
    }
    CFDB_LAUNCH(KK_HALO, s->st, k_halo_pull<<<blocks, 256, 0, s->st>>>(pl));
    return CFD_SUCCESS;
}

cfd_status_t peer_allreduce(cfd_b200_session* s, void* dev, int n, int op) {
    Peer& p = s->peer;
    if (n > PEER_RED_WORDS) return CFD_ERROR_INVALID;
    ReduceArgs a;
    memset(&a, 0, sizeof(a));
    for (int q = 0; q < s->desc.nranks; q++) a.box[q] = p.peer_box[q];
    a.vals = dev;
    a.timeout_flag = &s->ctl->comm_timeout;
    a.seq = ++p.rseq;
    a.R = s->desc.nranks;
    a.r = s->desc.rank;
    a.n = n;
    a.op = op;
    CFDB_LAUNCH(KK_ALLREDUCE, s->st, k_allreduce_peer<<<1, 32, 0, s->st>>>(a));
    static const bool hostsync = getenv("CFD_B200_PEER_HOSTSYNC") != nullptr;  // debugging aid
    if (hostsync) cudaStreamSynchronize(s->st);
    return CFD_SUCCESS;
}

}  // namespace cfdb

using namespace cfdb;

extern "C" {

// One process per GPU: export my box as a 64-byte cudaIpcMemHandle, gather all of them (the caller's job, e.g.
// torch.distributed.all_gather), attach.  After a successful attach the session uses the peer transport for every
// halo exchange and all-reduce; without it, the NCCL communicator of cfd_b200_session_init_comm.
cfd_status_t cfd_b200_session_peer_export(cfd_b200_session_t* s, char out[64]) {
    if (!s || !out) return CFD_ERROR_INVALID;
    cfd_status_t rc = peer_alloc_box(s);
    if (rc != CFD_SUCCESS) return rc;
    cudaIpcMemHandle_t h;
    static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
    if (cudaIpcGetMemHandle(&h, s->peer.box) != cudaSuccess) {
        set_error(CFD_ERROR_UNSUPPORTED, "cudaIpcGetMemHandle failed: %s", cudaGetErrorString(cudaGetLastError()));
        return CFD_ERROR_UNSUPPORTED;
    }
    memcpy(out, &h, 64);
    return CFD_SUCCESS;
}

cfd_status_t cfd_b200_session_peer_attach(cfd_b200_session_t* s, const char* handles) {
    if (!s || !handles || !s->peer.box) return CFD_ERROR_INVALID;
    CFDB_CUDA_OK(cudaSetDevice(s->dev));
    Peer& p = s->peer;
    for (int q = 0; q < s->desc.nranks; q++) {
        if (q == s->desc.rank) continue;
        cudaIpcMemHandle_t h;
        memcpy(&h, handles + (size_t)q * 64, 64);
        void* ptr = nullptr;
        if (cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
            set_error(CFD_ERROR_UNSUPPORTED, "cudaIpcOpenMemHandle(rank %d) failed: %s", q,
                      cudaGetErrorString(cudaGetLastError()));
            return CFD_ERROR_UNSUPPORTED;
        }
        p.peer_box[q] = (char*)ptr;
        p.ipc_mapped[q] = true;
    }
    return peer_connected(s);
}

int cfd_b200_session_transport(const cfd_b200_session_t* s) {  // 0 none, 1 NCCL, 2 peer memory
    if (!s || s->desc.nranks <= 1) return 0;
    return s->peer.on ? 2 : (s->nccl.comm ? 1 : 0);
}

}  // extern "C"
