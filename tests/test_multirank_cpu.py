"""World-size-2 tests (gloo, CPU) of the host-side slab decomposition that bench.py and the
multi-GPU sessions rely on: slab ranges, scatter/gather, halo exchange ordering (including the
two-rank periodic ring), the unique-id broadcast, and a decomposed Jacobi iteration that must
equal the undecomposed one bit for bit (no reduction order changes)."""
import os
import socket

import numpy as np
import pytest

from cfd_b200 import abi, dist


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _jacobi_sweep(x, rhs, dx, dy, dz, phys_lo, phys_hi):
    """One Jacobi sweep + Neumann ghosts on a slab (z-decomposed 3-D or y-decomposed 2-D)."""
    nz = x.shape[0]
    ix, iy = 1.0 / (dx * dx), 1.0 / (dy * dy)
    iz = 1.0 / (dz * dz) if nz > 1 else 0.0
    inv = 1.0 / (2.0 * (ix + iy + iz))
    new = x.copy()
    c = (slice(1, -1) if nz > 1 else slice(None), slice(1, -1), slice(1, -1))
    zsum = (x[2:, 1:-1, 1:-1] + x[:-2, 1:-1, 1:-1]) if nz > 1 else 0.0
    new[c] = -(rhs[c] - (x[c[0], 1:-1, 2:] + x[c[0], 1:-1, :-2]) * ix
               - (x[c[0], 2:, 1:-1] + x[c[0], :-2, 1:-1]) * iy - zsum * iz) * inv
    ax = dist.dec_axis(nz)
    # ghosts: x faces always physical; the decomposed axis only on the end ranks
    new[:, :, 0], new[:, :, -1] = new[:, :, 1], new[:, :, -2]
    if ax == 0:
        new[:, 0, :], new[:, -1, :] = new[:, 1, :], new[:, -2, :]
        if phys_lo:
            new[0] = new[1]
        if phys_hi:
            new[-1] = new[-2]
    else:
        if phys_lo:
            new[:, 0, :] = new[:, 1, :]
        if phys_hi:
            new[:, -1, :] = new[:, -2, :]
    return new


def _rb_relax(x, rhs, mask, dx, dy, dz, omega):
    """x + omega (pn - x) on the interior cells selected by `mask` (cpu/linear_solver_redblack.c:107-121)."""
    ix, iy, iz = 1.0 / (dx * dx), 1.0 / (dy * dy), 1.0 / (dz * dz)
    inv = 1.0 / (2.0 * (ix + iy + iz))
    c = (slice(1, -1), slice(1, -1), slice(1, -1))
    pn = -(rhs[c] - (x[1:-1, 1:-1, 2:] + x[1:-1, 1:-1, :-2]) * ix - (x[1:-1, 2:, 1:-1] + x[1:-1, :-2, 1:-1]) * iy
           - (x[2:, 1:-1, 1:-1] + x[:-2, 1:-1, 1:-1]) * iz) * inv
    new = x.copy()
    new[c] = np.where(mask[c], x[c] + omega * (pn - x[c]), x[c])
    return new


def _colour(shape, k0):
    """(i + j + k) & 1 with GLOBAL k = k0 + local k."""
    nz, ny, nx = shape
    return ((np.arange(nz)[:, None, None] + k0 + np.arange(ny)[None, :, None] + np.arange(nx)[None, None, :]) & 1)


def _neumann_xy(x):
    x[:, :, 0], x[:, :, -1] = x[:, :, 1], x[:, :, -2]
    x[:, 0, :], x[:, -1, :] = x[:, 1, :], x[:, -2, :]


def _rb_global(x, rhs, dx, dy, dz, omega):
    """The reference's iteration: odd half-sweep, even half-sweep (in place), then the Neumann ghosts."""
    col = _colour(x.shape, 0)
    x = _rb_relax(x, rhs, col == 1, dx, dy, dz, omega)
    x = _rb_relax(x, rhs, col == 0, dx, dy, dz, omega)
    _neumann_xy(x)
    x[0], x[-1] = x[1], x[-2]
    return x


def _rb_slab_one_pass(x, rhs, lo2, hi2, k0, dx, dy, dz, omega):
    """The one-pass sweep on a z-slab as poisson_tma3.inc does it: stage A (odd cells) on the own planes AND on the
    neighbour's edge plane (the halo layer), which needs the layer behind it (lo2 / hi2); stage B (even cells) on
    the own planes from stage A's results; ghosts; out of place.  Halo layers of the result come from the exchange."""
    zero = np.zeros_like(x[0])
    ext = np.concatenate([(lo2 if lo2 is not None else zero)[None], x, (hi2 if hi2 is not None else zero)[None]])
    rext = np.concatenate([zero[None], rhs, zero[None]])
    col = _colour(ext.shape, k0 - 1)
    a_planes = np.zeros(ext.shape, bool)
    a_planes[2:-2] = True                 # own planes
    a_planes[1] = lo2 is not None         # the lo neighbour's edge plane (my halo layer)
    a_planes[-2] = hi2 is not None
    R = _rb_relax(ext, rext, (col == 1) & a_planes, dx, dy, dz, omega)
    b_planes = np.zeros(ext.shape, bool)
    b_planes[2:-2] = True
    new = _rb_relax(R, rext, (col == 0) & b_planes, dx, dy, dz, omega)[1:-1]
    _neumann_xy(new)
    if lo2 is None:
        new[0] = new[1]
    if hi2 is None:
        new[-1] = new[-2]
    return new


def _worker(rank, world, port, shape, out):
    import torch.distributed as td
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    td.init_process_group("gloo", rank=rank, world_size=world)
    try:
        nx, ny, nz = shape
        rng = np.random.default_rng(42)
        gf = abi.HostField(nx, ny, nz)
        for name in gf.NAMES:
            getattr(gf, name)[...] = rng.standard_normal((nz, ny, nx))
        # 1. unique-id broadcast
        uid = dist.broadcast_unique_id(lambda: bytes(range(128)), rank)
        assert uid == bytes(range(128))
        # 2. scatter / gather round trip
        loc = dist.scatter(gf, world, rank)
        back = abi.HostField(nx, ny, nz)
        dist.gather_into(back, loc, world, rank)
        ax = dist.dec_axis(nz)
        n = nz if ax == 0 else ny
        lo, hi = dist.slab_range(n, world, rank)
        sl = [slice(None)] * 3
        sl[ax] = slice(lo + (0 if rank == 0 else 1), hi - (0 if rank == world - 1 else 1))
        np.testing.assert_array_equal(back.u[tuple(sl)], gf.u[tuple(sl)])
        # 3. halo exchange fills halos with the neighbour's owned layers
        a = loc.p.copy()
        hs = [slice(None)] * 3
        if rank > 0:
            hs[ax] = 0
            a[tuple(hs)] = -1.0
        if rank < world - 1:
            hs[ax] = a.shape[ax] - 1
            a[tuple(hs)] = -1.0
        dist.host_halo_exchange(a, rank, world)
        np.testing.assert_array_equal(a, loc.p)
        # 4. periodic ring: layer 0 <- last rank's layer n-2, layer n-1 <- first rank's layer 1
        b = loc.T.copy()
        dist.host_halo_exchange(b, rank, world, periodic_ring=True)
        g0 = [slice(None)] * 3
        if rank == 0:
            g0[ax] = n - 2
            hs[ax] = 0
            np.testing.assert_array_equal(b[tuple(hs)], gf.T[tuple(g0)])
        if rank == world - 1:
            g0[ax] = 1
            hs[ax] = b.shape[ax] - 1
            np.testing.assert_array_equal(b[tuple(hs)], gf.T[tuple(g0)])
        # 5. decomposed Jacobi == global Jacobi, bitwise
        dx, dy, dz = 0.1, 0.07, 0.13
        xg, rg = gf.p.copy(), gf.T.copy()
        xl, rl = loc.p.copy(), loc.T.copy()
        for _ in range(4):
            xg = _jacobi_sweep(xg, rg, dx, dy, dz, True, True)
            xl = _jacobi_sweep(xl, rl, dx, dy, dz, rank == 0, rank == world - 1)
            dist.host_halo_exchange(xl, rank, world)
        gsl = [slice(None)] * 3
        gsl[ax] = slice(lo, hi)
        np.testing.assert_array_equal(xl, xg[tuple(gsl)])
        # 6. all-reduce(max) of a local residual == global residual
        import torch
        own = [slice(None)] * 3
        own[ax] = slice(1, -1)
        t = torch.tensor([np.max(np.abs(xl[tuple(own)]))])
        td.all_reduce(t, op=td.ReduceOp.MAX)
        inner = [slice(None)] * 3
        inner[ax] = slice(1, -1)
        assert float(t) == np.max(np.abs(xg[tuple(inner)]))
        # 7. one-pass Red-Black SOR on slabs (two-layer exchange once per iteration) == the reference's two half-sweeps
        #    on the undecomposed grid, bitwise
        if ax == 0:
            omega = 1.6
            xg, xl = gf.p.copy(), loc.p.copy()
            _neumann_xy(xg)
            xg[0], xg[-1] = xg[1], xg[-2]
            xl = xg[lo:hi].copy()
            for _ in range(3):
                lo2, hi2 = dist.host_halo_exchange2(xl, rank, world)
                if rank > 0:
                    np.testing.assert_array_equal(lo2, xg[lo - 1])
                if rank < world - 1:
                    np.testing.assert_array_equal(hi2, xg[hi])
                xg_new = _rb_global(xg.copy(), rg, dx, dy, dz, omega)
                xl = _rb_slab_one_pass(xl, rl, lo2, hi2, lo, dx, dy, dz, omega)
                own = slice(1, -1)
                np.testing.assert_array_equal(xl[own], xg_new[lo:hi][own])
                xg = xg_new
            dist.host_halo_exchange(xl, rank, world)
            np.testing.assert_array_equal(xl, xg[lo:hi])
        out.put((rank, "ok"))
    except Exception as e:  # pragma: no cover - reported to the parent
        import traceback
        out.put((rank, traceback.format_exc() + repr(e)))
    finally:
        td.destroy_process_group()


@pytest.mark.parametrize("shape", [(9, 8, 10), (11, 14, 1)])
def test_two_rank_slab_logic(shape):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, shape, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in res:
        assert msg == "ok", "rank %d: %s" % (rank, msg)


def test_slab_helpers():
    assert dist.slab_extent(1024 * 8 - 14, 8) == 1024  # 1022 owned layers + 2 per rank
    assert dist.slab_range(10, 2, 0) == (0, 6) and dist.slab_range(10, 2, 1) == (4, 10)
    with pytest.raises(ValueError):
        dist.slab_extent(10, 3)
    g = dist.local_grid(8, 8, 10, (0, 1, 0, 1, 0, 9), 2, 1)
    assert g.nz == 6 and abs(g.z[0] - 4.0) < 1e-12 and abs(g.spacing[2] - 1.0) < 1e-12
    assert dist.dec_axis(1) == 1 and dist.dec_axis(5) == 0
