// session_io.cu -- host<->device field movement, coordinate tables, halo exchange and
// all-reduce over NCCL (one process per GPU; the communicator is bootstrapped by the caller,
// e.g. through torch.distributed, and handed to us as a 128-byte ncclUniqueId).
#include "session.h"

#include <dlfcn.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

namespace cfdb {

static double* const* field_ptrs(cfd_b200_session* s, double* out[6]) {
    out[0] = s->u; out[1] = s->v; out[2] = s->w; out[3] = s->p; out[4] = s->rho; out[5] = s->T;
    return out;
}

cfd_status_t session_upload(cfd_b200_session* s, const flow_field* f, int mask) {
    if (!s || !f) return CFD_ERROR_INVALID;
    if (f->nx != s->desc.nx || f->ny != s->desc.ny || f->nz != s->desc.nz) {
        set_error(CFD_ERROR_INVALID, "upload: field is %zux%zux%zu but the session is %zux%zux%zu", f->nx,
                  f->ny, f->nz, s->desc.nx, s->desc.ny, s->desc.nz);
        return CFD_ERROR_INVALID;
    }
    CFDB_CUDA_OK(cudaSetDevice(s->dev));
    if ((mask & FM_RHO) && !s->rho) {
        cfd_status_t rc = dalloc(s, &s->rho);
        if (rc != CFD_SUCCESS) return rc;
    }
    if ((mask & FM_T) && !s->T) {
        cfd_status_t rc = dalloc(s, &s->T);
        if (rc != CFD_SUCCESS) return rc;
    }
    double* dv[6];
    field_ptrs(s, dv);
    double* hv[6] = {f->u, f->v, f->w, f->p, f->rho, f->T};
    for (int q = 0; q < 6; q++) {
        if (!(mask & (1 << q))) { dv[q] = nullptr; continue; }
        if (!hv[q]) return CFD_ERROR_INVALID;
    }
    if (f->rho) s->rho0 = f->rho[0];
    return xfer_fields(s, dv, hv, 6, s->N * sizeof(double), true);
}

cfd_status_t session_download(cfd_b200_session* s, flow_field* f, int mask) {
    if (!s || !f) return CFD_ERROR_INVALID;
    if (f->nx != s->desc.nx || f->ny != s->desc.ny || f->nz != s->desc.nz) return CFD_ERROR_INVALID;
    CFDB_CUDA_OK(cudaSetDevice(s->dev));
    double* dv[6];
    field_ptrs(s, dv);
    double* hv[6] = {f->u, f->v, f->w, f->p, f->rho, f->T};
    for (int q = 0; q < 6; q++)
        if (!(mask & (1 << q))) dv[q] = nullptr;
    return xfer_fields(s, dv, hv, 6, s->N * sizeof(double), false);
}

cfd_status_t session_set_coords(cfd_b200_session* s, const double* x, const double* y) {
    if (x) s->xs.assign(x, x + s->desc.nx);
    if (y) s->ys.assign(y, y + s->desc.ny);
    s->tab_valid = false;
    return CFD_SUCCESS;
}

// ---- NCCL plumbing -------------------------------------------------------------------
cfd_status_t load_nccl(Nccl& n, const char* path) {
    if (n.lib) return CFD_SUCCESS;
    // one NCCL per process: when the host program (e.g. torch) has already loaded libnccl.so.2, take that
    // copy whatever `path` says; only otherwise load the named / default library
    if (!getenv("CFD_B200_NCCL_FORCE_PATH")) n.lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);
    if (!n.lib) n.lib = dlopen(path && path[0] ? path : "libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!n.lib) {
        set_error(CFD_ERROR_UNSUPPORTED, "dlopen(%s) failed: %s", path ? path : "libnccl.so.2", dlerror());
        return CFD_ERROR_UNSUPPORTED;
    }
#define SYM(field, name)                                                   \
    *(void**)(&n.field) = dlsym(n.lib, name);                              \
    if (!n.field) {                                                        \
        set_error(CFD_ERROR_UNSUPPORTED, "NCCL symbol %s not found", name); \
        return CFD_ERROR_UNSUPPORTED;                                      \
    }
    SYM(GetUniqueId, "ncclGetUniqueId")
    SYM(CommInitRank, "ncclCommInitRank")
    SYM(CommDestroy, "ncclCommDestroy")
    SYM(AllReduce, "ncclAllReduce")
    SYM(Send, "ncclSend")
    SYM(Recv, "ncclRecv")
    SYM(GroupStart, "ncclGroupStart")
    SYM(GroupEnd, "ncclGroupEnd")
    SYM(CommInitAll, "ncclCommInitAll")
#undef SYM
    return CFD_SUCCESS;
}

// One-cell halo exchange along the decomposed axis for nf fields: send my first/last owned
// plane (row in 2-D), receive into my halo plane.  periodic_ring also links rank 0 <-> last
// (RK steppers: periodic wrap across the decomposed axis).  Stream-ordered, no host sync.
cfd_status_t halo_exchange(cfd_b200_session* s, double* const* fields, int nf, bool periodic_ring) {
    const int R = s->desc.nranks, r = s->desc.rank;
    if (R <= 1) return CFD_SUCCESS;
    if (s->peer.on) return peer_halo_exchange(s, fields, nf, periodic_ring);
    if (!s->nccl.comm) {
        set_error(CFD_ERROR, "multi-rank session without a communicator (call cfd_b200_session_init_comm)");
        if (s->comm_err == CFD_SUCCESS) s->comm_err = CFD_ERROR;
        return CFD_ERROR;
    }
    const Geom& g = s->g;
    const size_t cnt = g.dim3 ? (size_t)g.nx * g.ny : (size_t)g.nx;
    const long long n_ax = g.dim3 ? g.nz : g.ny;
    const long long st = g.dim3 ? (long long)g.nx * g.ny : g.nx;
    int lo = r - 1, hi = r + 1;
    if (periodic_ring) {
        lo = (r + R - 1) % R;
        hi = (r + 1) % R;
    }
    const Nccl& n = s->nccl;
    n.GroupStart();
    for (int q = 0; q < nf; q++) {
        double* f = fields[q];
        // sends in the order (lo, hi), receives in the order (hi, lo): with two ranks in a
        // periodic ring both neighbours are the same peer and NCCL matches messages in order.
        const bool has_lo = lo >= 0 && lo < R, has_hi = hi >= 0 && hi < R;
        if (has_lo) n.Send(f + 1 * st, cnt, 8, lo, n.comm, s->st);
        if (has_hi) n.Send(f + (n_ax - 2) * st, cnt, 8, hi, n.comm, s->st);
        if (has_hi) n.Recv(f + (n_ax - 1) * st, cnt, 8, hi, n.comm, s->st);
        if (has_lo) n.Recv(f + 0 * st, cnt, 8, lo, n.comm, s->st);
    }
    if (n.GroupEnd() != 0) {
        set_error(CFD_ERROR, "NCCL halo exchange failed");
        if (s->comm_err == CFD_SUCCESS) s->comm_err = CFD_ERROR;
        return CFD_ERROR;
    }
    return CFD_SUCCESS;
}

static cfd_status_t allreduce_any(cfd_b200_session* s, void* dev, int cnt, int dtype, int op) {
    if (s->desc.nranks <= 1) return CFD_SUCCESS;
    if (s->peer.on) return peer_allreduce(s, dev, cnt, dtype == 2 ? 2 : (op == 2 ? 1 : 0));
    if (!s->nccl.comm) {
        set_error(CFD_ERROR, "multi-rank session without a communicator (call cfd_b200_session_init_comm)");
        if (s->comm_err == CFD_SUCCESS) s->comm_err = CFD_ERROR;
        return CFD_ERROR;
    }
    if (s->nccl.AllReduce(dev, dev, (size_t)cnt, dtype, op, s->nccl.comm, s->st) != 0) {
        set_error(CFD_ERROR, "NCCL all-reduce failed");
        if (s->comm_err == CFD_SUCCESS) s->comm_err = CFD_ERROR;
        return CFD_ERROR;
    }
    return CFD_SUCCESS;
}
// ncclDouble = 8, ncclInt32 = 2; ncclSum = 0, ncclMax = 2 (nccl.h 2.27 / 2.28)
cfd_status_t allreduce(cfd_b200_session* s, double* dev, int cnt, bool is_max) {
    return allreduce_any(s, dev, cnt, 8, is_max ? 2 : 0);
}
cfd_status_t allreduce_int_max(cfd_b200_session* s, int* dev, int cnt) { return allreduce_any(s, dev, cnt, 2, 2); }
cfd_status_t allreduce_produced(cfd_b200_session* s, double* dev, int cnt, bool is_max) {
    if (s->rb.peer) return CFD_SUCCESS;  // block_reduce_finish of the producing kernel already combined the ranks
    return allreduce(s, dev, cnt, is_max);
}

cfd_status_t comm_ready(cfd_b200_session* s) {
    if (s->desc.nranks > 1 && !s->nccl.comm && !s->peer.on) {
        set_error(CFD_ERROR, "session is rank %d of %d but has no communicator (call cfd_b200_session_init_comm)",
                  s->desc.rank, s->desc.nranks);
        return CFD_ERROR;
    }
    return CFD_SUCCESS;
}
// (call after a host synchronisation that refreshed h_ctl[0])
cfd_status_t comm_check(cfd_b200_session* s) {
    cfd_status_t rc = s->comm_err;
    s->comm_err = CFD_SUCCESS;
    if (rc == CFD_SUCCESS && s->peer.on && s->h_ctl[0].comm_timeout) {
        set_error(CFD_ERROR, "peer transport: rank %d waited too long for a neighbour (lost rank or mismatched call sequence)",
                  s->desc.rank);
        cudaMemsetAsync(&s->ctl->comm_timeout, 0, 2 * sizeof(int), s->st);
        if (getenv("CFD_B200_PEER_DEBUG")) {  // which exchange / reduction never arrived
            PeerBoxHeader h;
            cudaMemcpy(&h, s->peer.box, sizeof(h), cudaMemcpyDeviceToHost);
            fprintf(stderr, "[peer] first wait that gave up: code %d seq %d; ", s->h_ctl[0].comm_timeout, s->h_ctl[0].pad[0]);
            fprintf(stderr, "[peer] rank %d/%d issued halo #%llu reduce #%llu; box: halo flags lo{%llu,%llu} hi{%llu,%llu}; reduce flags",
                    s->desc.rank, s->desc.nranks, s->peer.hseq, s->peer.rseq, h.hflag[0][0], h.hflag[0][1], h.hflag[1][0],
                    h.hflag[1][1]);
            for (int q = 0; q < s->desc.nranks; q++) fprintf(stderr, " r%d{%llu,%llu}", q, h.rflag[q][0], h.rflag[q][1]);
            fprintf(stderr, "\n");
        }
        rc = CFD_ERROR;
    }
    return rc;
}
size_t global_dec_extent(const cfd_b200_session* s) {
    if (s->gdec) return s->gdec;
    const size_t loc = s->g.dim3 ? s->desc.nz : s->desc.ny;
    return s->desc.nranks > 1 ? (loc - 2) * (size_t)s->desc.nranks + 2 : loc;
}

}  // namespace cfdb

using namespace cfdb;

extern "C" {

cfd_status_t cfd_b200_nccl_unique_id(const char* nccl_lib, char out[128]) {
    Nccl n;
    cfd_status_t rc = load_nccl(n, nccl_lib);
    if (rc != CFD_SUCCESS) return rc;
    NcclId id;
    if (n.GetUniqueId(&id) != 0) return CFD_ERROR;
    memcpy(out, id.internal, 128);
    return CFD_SUCCESS;
}

cfd_status_t cfd_b200_session_init_comm(cfd_b200_session_t* s, const char* nccl_lib, const char id[128]) {
    if (!s || !id) return CFD_ERROR_INVALID;
    cfd_status_t rc = load_nccl(s->nccl, nccl_lib);
    if (rc != CFD_SUCCESS) return rc;
    CFDB_CUDA_OK(cudaSetDevice(s->dev));
    NcclId nid;
    memcpy(nid.internal, id, 128);
    if (s->nccl.CommInitRank(&s->nccl.comm, s->desc.nranks, nid, s->desc.rank) != 0) {
        set_error(CFD_ERROR, "ncclCommInitRank(rank %d of %d) failed", s->desc.rank, s->desc.nranks);
        return CFD_ERROR;
    }
    return CFD_SUCCESS;
}

cfd_status_t cfd_b200_session_set_slab(cfd_b200_session_t* s, size_t global_layers, size_t global_lo) {
    if (!s || global_layers < 3) return CFD_ERROR_INVALID;
    const size_t loc = s->g.dim3 ? s->desc.nz : s->desc.ny;
    if (global_lo + loc > global_layers) {
        set_error(CFD_ERROR_INVALID, "slab [%zu, %zu) does not fit in %zu global layers", global_lo, global_lo + loc,
                  global_layers);
        return CFD_ERROR_INVALID;
    }
    s->gdec = global_layers;
    s->rb_slab_ok = -1;
    if (s->g.dim3) s->g.gk0 = (long long)global_lo;
    else s->g.gj0 = (long long)global_lo;
    // physical faces follow from the position, not from the rank number
    s->g.glo[s->dec_axis] = (global_lo == 0);
    s->g.ghi[s->dec_axis] = (global_lo + loc == global_layers);
    return CFD_SUCCESS;
}

cfd_status_t cfd_b200_session_set_coords(cfd_b200_session_t* s, const double* x, const double* y) {
    if (!s) return CFD_ERROR_INVALID;
    return session_set_coords(s, x, y);
}
cfd_status_t cfd_b200_session_upload(cfd_b200_session_t* s, const flow_field* host) {
    return session_upload(s, host, FM_ALL);
}
cfd_status_t cfd_b200_session_download(cfd_b200_session_t* s, flow_field* host) {
    return session_download(s, host, FM_ALL);
}
cfd_status_t cfd_b200_session_upload_mask(cfd_b200_session_t* s, const flow_field* host, int mask) {
    return session_upload(s, host, mask);
}
cfd_status_t cfd_b200_session_download_mask(cfd_b200_session_t* s, flow_field* host, int mask) {
    return session_download(s, host, mask);
}

cfd_status_t cfd_b200_session_init_synthetic(cfd_b200_session_t* s, uint64_t seed, size_t global_k0,
                                             size_t global_nz) {
    if (!s) return CFD_ERROR_INVALID;
    CFDB_CUDA_OK(cudaSetDevice(s->dev));
    t_prof = &s->prof;
    launch_synth_init(s->L(), seed, (long long)global_k0, (long long)global_nz, s->u, s->v, s->w, s->p,
                      s->T, s->rho);
    s->rho0 = 1.0;
    CFDB_CUDA_OK(cudaStreamSynchronize(s->st));
    return CFD_SUCCESS;
}

cfd_status_t cfd_b200_session_upload_poisson(cfd_b200_session_t* s, const double* x, const double* rhs) {
    if (!s) return CFD_ERROR_INVALID;
    CFDB_CUDA_OK(cudaSetDevice(s->dev));
    cfd_status_t rc = dalloc(s, &s->rhs);
    if (rc != CFD_SUCCESS) return rc;
    double* dv[2] = {x ? s->p : nullptr, rhs ? s->rhs : nullptr};
    double* hv[2] = {const_cast<double*>(x), const_cast<double*>(rhs)};
    return xfer_fields(s, dv, hv, 2, s->N * sizeof(double), true);
}
cfd_status_t cfd_b200_session_download_poisson(cfd_b200_session_t* s, double* x) {
    if (!s || !x) return CFD_ERROR_INVALID;
    double* dv[1] = {s->p};
    double* hv[1] = {x};
    return xfer_fields(s, dv, hv, 1, s->N * sizeof(double), false);
}

}  // extern "C"
