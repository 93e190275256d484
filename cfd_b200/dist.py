"""One-process-per-GPU plumbing: slab decomposition of the global grid along z (along y when
nz == 1), torch.distributed bootstrap of the NCCL communicator the C library uses for halo
exchange and Krylov all-reduces, and scatter/gather helpers for host fields.

The reference has no distributed code; this is new design.  Layout:
rank r owns `n_own` interior layers of the decomposed axis and keeps one extra layer on each
side: the physical boundary layer on the first/last rank, a halo layer otherwise.  Uniform
slabs are required (global interior layers divisible by the number of ranks).
"""
import ctypes as C
import glob
import os

import numpy as np

from . import abi


def dec_axis(nz):
    """Decomposed axis in (nz, ny, nx) array order: 0 (z) for 3-D grids, 1 (y) for 2-D."""
    return 0 if nz > 1 else 1


def slab_extent(n_global, nranks):
    """Local layer count (including the two boundary/halo layers) of a uniform slab."""
    interior = n_global - 2
    if interior % nranks != 0:
        raise ValueError("global interior layers (%d) must divide by the rank count (%d)"
                         % (interior, nranks))
    return interior // nranks + 2


def slab_range(n_global, nranks, rank):
    """[lo, hi) global layer indices held by `rank` (boundary/halo layers included)."""
    own = (n_global - 2) // nranks
    lo = rank * own
    return lo, lo + own + 2


def local_grid(gnx, gny, gnz, bounds, nranks, rank):
    """HostGrid of the slab of `rank`; bounds = (xmin, xmax, ymin, ymax, zmin, zmax) global."""
    xmin, xmax, ymin, ymax, zmin, zmax = bounds
    if gnz > 1:
        dz = (zmax - zmin) / (gnz - 1)
        lo, hi = slab_range(gnz, nranks, rank)
        return abi.HostGrid(gnx, gny, hi - lo, xmin, xmax, ymin, ymax, zmin + lo * dz,
                            zmin + (hi - 1) * dz)
    dy = (ymax - ymin) / (gny - 1)
    lo, hi = slab_range(gny, nranks, rank)
    return abi.HostGrid(gnx, hi - lo, 1, xmin, xmax, ymin + lo * dy, ymin + (hi - 1) * dy)


def scatter(global_field, nranks, rank):
    """Copy the slab of `rank` (with its halo layers) out of a global HostField."""
    ax = dec_axis(global_field.nz)
    n = global_field.nz if ax == 0 else global_field.ny
    lo, hi = slab_range(n, nranks, rank)
    shape = [global_field.nz, global_field.ny, global_field.nx]
    shape[ax] = hi - lo
    loc = abi.HostField(shape[2], shape[1], shape[0])
    sl = [slice(None)] * 3
    sl[ax] = slice(lo, hi)
    for name in abi.HostField.NAMES:
        getattr(loc, name)[...] = getattr(global_field, name)[tuple(sl)]
    return loc


def gather_into(global_field, local_field, nranks, rank):
    """Write the layers OWNED by `rank` (plus the physical boundary layers it holds) back."""
    ax = dec_axis(global_field.nz)
    n = global_field.nz if ax == 0 else global_field.ny
    lo, hi = slab_range(n, nranks, rank)
    a = 0 if rank == 0 else 1
    b = (hi - lo) if rank == nranks - 1 else (hi - lo - 1)
    gs, ls = [slice(None)] * 3, [slice(None)] * 3
    gs[ax] = slice(lo + a, lo + b)
    ls[ax] = slice(a, b)
    for name in abi.HostField.NAMES:
        getattr(global_field, name)[tuple(gs)] = getattr(local_field, name)[tuple(ls)]


def find_nccl():
    """Path of the libnccl.so.2 torch itself uses (bundled wheel first, then the system one)."""
    try:
        import nvidia.nccl
        cand = glob.glob(os.path.join(os.path.dirname(nvidia.nccl.__file__), "lib", "libnccl.so*"))
        if cand:
            return sorted(cand)[0]
    except Exception:
        pass
    for p in ("/usr/lib/x86_64-linux-gnu/libnccl.so.2", "/usr/local/cuda/lib64/libnccl.so.2"):
        if os.path.exists(p):
            return p
    return "libnccl.so.2"


def broadcast_unique_id(make_id, rank, src=0):
    """torch.distributed broadcast of the 128-byte ncclUniqueId produced on `src`."""
    import torch
    import torch.distributed as td
    buf = torch.zeros(128, dtype=torch.uint8)
    if rank == src:
        raw = make_id()
        buf = torch.tensor(list(raw), dtype=torch.uint8)
    dev = None
    if td.get_backend() == "nccl":
        dev = torch.device("cuda", torch.cuda.current_device())
        buf = buf.to(dev)
    td.broadcast(buf, src=src)
    return bytes(buf.cpu().tolist())


def attach_peers(session, rank, nranks):
    """Peer-memory transport across processes: every rank exports its box as a cudaIpcMemHandle, the handles
    are all-gathered, every rank maps the others' boxes.  Returns True when EVERY rank succeeded."""
    import torch.distributed as td
    try:
        mine = session.peer_export()
    except Exception:
        mine = None
    got = [None] * nranks
    td.all_gather_object(got, mine)
    ok = all(h is not None for h in got)
    if ok:
        ok = session.peer_attach(b"".join(got)) == 0
    flags = [None] * nranks
    td.all_gather_object(flags, bool(ok))
    return all(flags)


def attach_comm(session, rank, nranks, transport=None):
    """Connect the slab of `session` to its neighbours (all ranks must call): the peer-memory transport
    (csrc/peer.cu) when every rank can map the others' memory, else an NCCL communicator owned by the library.
    transport / env CFD_B200_TRANSPORT: "peer" (default, falls back to NCCL), "nccl"."""
    from .lib import check, load
    if nranks <= 1:
        return
    transport = transport or os.environ.get("CFD_B200_TRANSPORT", "peer")
    if transport == "peer" and attach_peers(session, rank, nranks):
        return
    L = load()
    path = find_nccl()

    def make_id():
        out = C.create_string_buffer(128)
        check(L.cfd_b200_nccl_unique_id(path.encode(), out), "cfd_b200_nccl_unique_id")
        return out.raw

    uid = broadcast_unique_id(make_id, rank)
    session.init_comm(path, uid)


def host_halo_exchange(arr, rank, nranks, periodic_ring=False):
    """CPU restatement of the device halo exchange on a (nz, ny, nx) numpy array using
    torch.distributed point-to-point (any backend).  Used by the gloo tests of the slab logic."""
    import torch
    import torch.distributed as td
    ax = dec_axis(arr.shape[0])
    n = arr.shape[ax]

    def layer(i):
        sl = [slice(None)] * 3
        sl[ax] = i
        return tuple(sl)

    lo, hi = rank - 1, rank + 1
    if periodic_ring:
        lo, hi = (rank - 1) % nranks, (rank + 1) % nranks
    # sends ordered (lo, hi), receives ordered (hi, lo): see halo_exchange in session_io.cu
    sends, recvs, recv = [], [], []
    if 0 <= lo < nranks:
        sends.append(td.P2POp(td.isend, torch.from_numpy(np.ascontiguousarray(arr[layer(1)])), lo))
    if 0 <= hi < nranks:
        sends.append(td.P2POp(td.isend, torch.from_numpy(np.ascontiguousarray(arr[layer(n - 2)])), hi))
        r = torch.empty_like(torch.from_numpy(np.ascontiguousarray(arr[layer(n - 1)])))
        recvs.append(td.P2POp(td.irecv, r, hi))
        recv.append((n - 1, r))
    if 0 <= lo < nranks:
        r = torch.empty_like(torch.from_numpy(np.ascontiguousarray(arr[layer(0)])))
        recvs.append(td.P2POp(td.irecv, r, lo))
        recv.append((0, r))
    ops = sends + recvs
    if ops:
        for w in td.batch_isend_irecv(ops):
            w.wait()
    for i, r in recv:
        arr[layer(i)] = r.numpy()


def host_halo_exchange2(arr, rank, nranks):
    """CPU restatement of peer_halo_exchange2 (csrc/peer.cu): the usual one-layer exchange of `arr` plus the layer
    BEHIND each neighbour's edge layer, returned as (lo2, hi2) -- None on a side without a neighbour.  The one-pass
    Red-Black SOR sweep on slabs reads x two layers deep (csrc/poisson_tma3.inc)."""
    import torch
    import torch.distributed as td
    ax = dec_axis(arr.shape[0])
    n = arr.shape[ax]

    def layer(i):
        sl = [slice(None)] * 3
        sl[ax] = i
        return tuple(sl)

    host_halo_exchange(arr, rank, nranks)
    lo, hi = rank - 1, rank + 1
    ops, got = [], {}
    if 0 <= lo < nranks:
        ops.append(td.P2POp(td.isend, torch.from_numpy(np.ascontiguousarray(arr[layer(2)])), lo))
    if 0 <= hi < nranks:
        ops.append(td.P2POp(td.isend, torch.from_numpy(np.ascontiguousarray(arr[layer(n - 3)])), hi))
        got["hi"] = torch.empty_like(torch.from_numpy(np.ascontiguousarray(arr[layer(0)])))
        ops.append(td.P2POp(td.irecv, got["hi"], hi))
    if 0 <= lo < nranks:
        got["lo"] = torch.empty_like(torch.from_numpy(np.ascontiguousarray(arr[layer(0)])))
        ops.append(td.P2POp(td.irecv, got["lo"], lo))
    if ops:
        for w in td.batch_isend_irecv(ops):
            w.wait()
    return (got["lo"].numpy() if "lo" in got else None, got["hi"].numpy() if "hi" in got else None)

