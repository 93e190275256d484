"""Multi-GPU parity worker: run under torchrun with one rank per GPU.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 \
        --master-port P tests/multigpu_parity.py

Every rank builds the same global problem, takes its z-slab (y-slab for 2-D), steps it on its own
GPU with halo exchange / all-reduce over NCCL inside libcfd_b200, and rank 0 compares the gathered
result with the UNDECOMPOSED oracle: bitwise equal for Jacobi, Red-Black SOR,
RK2/RK4 and explicit Euler (no reduction order changes), <= 1e-8 for CG / PCG, BiCGSTAB at a fixed iteration count.
Prints one line per case and "MULTIGPU PARITY OK" at the end; exit code 1 on any mismatch.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


class Ctx:
    """Per-process state of a parity run: rank / world, the oracle, torch.distributed."""

    def __init__(self, td, rank, world, local):
        import oracle
        self.td, self.rank, self.world, self.local = td, rank, world, local
        self.orc = oracle.Oracle()
        self.fails = []
        self.results = []
        self.transport = "?"


def global_case(kind, nx, ny, nzg):
    import cases
    if kind == "natconv":
        return cases.natural_convection(nx, nzg, dt=1e-3) if nzg > 1 else cases.natural_convection(nx, 1, dt=1e-3)
    if kind == "natconv_y":  # 2-D natural convection on an nx x ny grid (C4 shape: y-slabs)
        from cfd_b200 import abi
        g, f, p = cases.natural_convection(nx, 1, dt=2e-5)
        g = abi.HostGrid(nx, ny, 1, 0.0, 1.0, 0.0, 1.0)
        f = abi.HostField(nx, ny, 1)
        f.T[...] = 310.0 - 20.0 * g.x[None, None, :]
        return g, f, p
    g, f = cases.random_field(nx, ny, nzg, seed=21, scale=0.05)
    f.rho[...] = 1.0 + 0.1 * np.random.default_rng(2).random(f.rho.shape)
    return g, f, cases.params(dt=1e-3, amp=(0.1, 0.05))


def run_case(cx, name, kind, shape, stepper, method=None, nsteps=2, tol=0.0, poison=None, **pkw):
    """One decomposed run against the UNDECOMPOSED oracle.  poison = (k, j, i): seed a NaN in that global cell
    in the pressure
    (every rank must leave the step loop at the same step with the same status)."""
    import cases
    from cfd_b200 import abi, api, dist
    td, rank, world, local, orc = cx.td, cx.rank, cx.world, cx.local, cx.orc
    nx, ny, nzg = shape
    g, f, p = global_case(kind, nx, ny, nzg)
    if poison:  # pressure is never clamped (solver_rk4.c:47-63), so the NaN survives to the scan
        f.p[poison] = np.nan
    fo = f.copy()
    pc = abi.poisson_config(method or "cg", on_nonconvergence=abi.CFD_B200_NONCONV_CONTINUE, **pkw)
    if stepper == "projection":
        rc_o, _ = orc.projection(fo, g, p, pc, nsteps)
    elif stepper == "euler":
        rc_o = orc.explicit_euler(fo, g, p, nsteps)
    else:
        rc_o = orc.rk(fo, g, p, int(stepper[2]), nsteps)
    loc = dist.scatter(f, world, rank)
    bounds = (g.x[0], g.x[-1], g.y[0], g.y[-1], g.z[0] if nzg > 1 else 0.0, g.z[-1] if nzg > 1 else 0.0)
    lg = dist.local_grid(nx, ny, nzg, bounds, world, rank)
    dx, dy, dz = g.spacing
    s = api.Session(lg.nx, lg.ny, lg.nz, dx, dy, dz, float(lg.x[0]), float(lg.y[0]),
                    float(lg.z[0]) if nzg > 1 else 0.0, rank=rank, nranks=world, device=local)
    dist.attach_comm(s, rank, world)
    cx.transport = s.transport()
    # bit-identical coordinates on every rank: the slice of the GLOBAL grid arrays
    if nzg > 1:
        s.set_coords(g.x, g.y)
    else:
        lo, hi = dist.slab_range(ny, world, rank)
        s.set_coords(g.x, g.y[lo:hi])
    s.upload(loc)
    if stepper == "projection":
        rc, _ = s.projection(p, pc, nsteps)
    elif stepper == "euler":
        rc, _ = s.rk(1, p, nsteps)
    else:
        rc, _ = s.rk(int(stepper[2]), p, nsteps)
    s.download(loc)
    s.close()
    parts = [None] * world
    td.all_gather_object(parts, {"rc": rc, **{k: getattr(loc, k) for k in ("u", "v", "w", "p", "T")}})
    res = None
    if rank == 0:
        out = abi.HostField(nx, ny, nzg)
        for r in range(world):
            lf = abi.HostField(loc.nx, loc.ny, loc.nz)
            for k, a in parts[r].items():
                if k != "rc":
                    getattr(lf, k)[...] = a
            dist.gather_into(out, lf, world, r)
        worst = 0.0
        ok = all(parts[r]["rc"] == rc_o for r in range(world))
        for k in ("u", "v", "w", "p", "T"):
            a, b = getattr(out, k), getattr(fo, k)
            if poison:  # a diverged step: statuses must agree; fields are whatever the step left behind
                continue
            if tol == 0.0:
                same = np.array_equal(a, b)
                worst = max(worst, float(np.max(np.abs(a - b))))
                ok = ok and same
            else:
                e = max(cases.rel_err(a, b)) if np.max(np.abs(b)) > 1e-12 else float(np.max(np.abs(a - b)))
                worst = max(worst, e)
                ok = ok and e <= tol
        print("%-44s world=%d rc=%s/%d worst=%.3e tol=%s %s" % (name, world, [parts[r]["rc"] for r in range(world)] if poison else rc,
                                                                  rc_o, worst, "bitwise" if tol == 0.0 else "%.0e" % tol,
                                                                  "ok" if ok else "FAIL"), flush=True)
        if not ok:
            cx.fails.append(name)
        res = {"case": name, "worst": worst, "tol": tol, "ok": bool(ok)}
        cx.results.append(res)
    td.barrier()
    return res


def bench_parity(cx):
    """The decomposed-vs-oracle cases bench.py --gpus N (N > 1) runs before its timed region: the headline
    method on a box that spans 2 x 2 TMA tiles and >= 2 z-chunks per rank, and the bitwise Jacobi twin.
    Returns {"case", "worst", "ok"} on rank 0 (None elsewhere)."""
    w = cx.world
    a = run_case(cx, "projection+bicgstab x12 130x34x(40w+2)", "random", (130, 34, 40 * w + 2), "projection", "bicgstab",
                 tol=1e-8, max_iterations=12)
    b = run_case(cx, "projection+jacobi x20 130x34x(40w+2)", "random", (130, 34, 40 * w + 2), "projection", "jacobi",
                 max_iterations=20)
    if cx.rank != 0:
        return None
    return {"case": "%s | %s" % (a["case"], b["case"]), "worst": max(a["worst"], b["worst"]),
            "worst_bitwise_case": b["worst"], "tolerance": "1e-8 (BiCGSTAB x12) / bitwise (Jacobi)",
            "ok": bool(a["ok"] and b["ok"]), "world": w}


def main():
    import torch
    import torch.distributed as td
    from cfd_b200 import abi, api
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    td.init_process_group("nccl", device_id=torch.device("cuda", local))
    cx = Ctx(td, rank, world, local)
    fails = cx.fails

    def case(*a, **kw):
        return run_case(cx, *a, **kw)

    nzg = 6 * world + 2      # 6 owned planes per rank
    nyg = 8 * world + 2
    case("projection+jacobi 3d", "random", (20, 14, nzg), "projection", "jacobi", max_iterations=40)
    case("projection+rbsor 3d", "random", (20, 14, nzg), "projection", "rbsor", max_iterations=30)
    case("projection+cg 3d", "random", (20, 14, nzg), "projection", "cg", tol=1e-8)
    case("projection+pcg 3d", "random", (20, 14, nzg), "projection", "pcg", tol=1e-8)
    case("projection+bicgstab x12 3d", "random", (20, 14, nzg), "projection", "bicgstab", tol=1e-8, max_iterations=12)
    case("rk4 3d (periodic ring)", "random", (20, 14, nzg), "rk4")
    case("rk2 3d (periodic ring)", "random", (20, 14, nzg), "rk2")
    case("projection+rbsor 2d y-slabs", "random", (22, nyg, 1), "projection", "rbsor", max_iterations=30)
    case("projection+pcg 2d y-slabs", "random", (22, nyg, 1), "projection", "pcg", tol=1e-8)
    case("rk4 2d y-slabs", "random", (22, nyg, 1), "rk4")
    case("explicit euler 3d", "random", (20, 14, nzg), "euler", nsteps=3)
    case("explicit euler 3d nx odd", "random", (21, 14, nzg), "euler", nsteps=3)
    case("explicit euler 2d y-slabs", "random", (22, nyg, 1), "euler", nsteps=3)
    case("explicit euler + energy 3d", "natconv", (14, 14, nzg), "euler", nsteps=3)
    # even nx: the TMA kernels (odd-nx grids above take the direct kernels... 20/22 are even: both covered)
    case("projection+jacobi 3d nx odd", "random", (21, 14, nzg), "projection", "jacobi", max_iterations=25)
    case("projection+cg 3d nx odd", "random", (21, 14, nzg), "projection", "cg", tol=1e-8)
    # multi-tile / multi-chunk slabs: 2 x 2 TMA tiles of 64 x 16 in the plane, 40 planes (>= 2 z-chunks, the halo
    # plane written by the combining engine) per rank; 2-D: 3 row tiles per rank, 2 tiles across
    big3, big2 = (130, 34, 40 * world + 2), (130, 34 * world + 2, 1)
    case("multi-tile jacobi 3d", "random", big3, "projection", "jacobi", max_iterations=20)
    case("multi-tile rbsor 3d", "random", big3, "projection", "rbsor", max_iterations=20)
    case("multi-tile cg 3d", "random", big3, "projection", "cg", tol=1e-8, max_iterations=40)
    case("multi-tile bicgstab x12 3d", "random", big3, "projection", "bicgstab", tol=1e-8, max_iterations=12)
    case("multi-tile rk4 3d", "random", big3, "rk4")
    case("multi-tile euler 3d", "random", big3, "euler")
    case("multi-tile jacobi 2d y-slabs", "random", big2, "projection", "jacobi", max_iterations=20)
    case("multi-tile rbsor 2d y-slabs", "random", big2, "projection", "rbsor", max_iterations=20)
    case("multi-tile pcg 2d y-slabs", "random", big2, "projection", "pcg", tol=1e-8, max_iterations=40)
    case("multi-tile rk4 2d y-slabs", "random", big2, "rk4")
    # C4 shape: natural convection (energy + buoyancy + thermal BCs) + PCG on y-slabs
    case("C4 natconv pcg 258x(64w+2) y-slabs", "natconv_y", (258, 64 * world + 2, 1), "projection", "pcg", tol=1e-8,
         nsteps=3, max_iterations=60)
    # a NaN seeded in ONE slab: every rank must report CFD_ERROR_DIVERGED at the same step (no hang)
    case("NaN in one slab: rk4", "random", (20, 14, nzg), "rk4", poison=(2, 5, 7))
    case("NaN in one slab: euler", "random", (20, 14, nzg), "euler", poison=(2, 5, 7))
    if world > 1:  # the multigrid hierarchy is single-GPU: a decomposed session must refuse it, not mis-solve
        s = api.Session(20, 14, 8, 0.1, 0.1, 0.1, rank=rank, nranks=world, device=local)
        rc, _ = s.poisson(abi.poisson_config("multigrid"))
        s.close()
        if rank == 0:
            ok = rc == abi.CFD_ERROR_UNSUPPORTED
            print("%-34s world=%d rc=%d %s" % ("multigrid on a slab is refused", world, rc, "ok" if ok else "FAIL"), flush=True)
            if not ok:
                fails.append("multigrid refusal")
    td.barrier()
    td.destroy_process_group()
    if rank == 0:
        if fails:
            print("MULTIGPU PARITY FAILED:", fails)
            sys.exit(1)
        print("MULTIGPU PARITY OK (world=%d transport=%s)" % (world, cx.transport))


if __name__ == "__main__":
    main()
