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


def main():
    import torch
    import torch.distributed as td
    import cases
    import oracle
    from cfd_b200 import abi, api, dist
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    td.init_process_group("nccl", device_id=torch.device("cuda", local))
    orc = oracle.Oracle()
    fails = []

    def global_case(kind, nx, ny, nzg):
        if kind == "natconv":
            g, f, p = cases.natural_convection(nx, nzg, dt=1e-3) if nzg > 1 else cases.natural_convection(nx, 1, dt=1e-3)
            return g, f, p
        g, f = cases.random_field(nx, ny, nzg, seed=21, scale=0.05)
        f.rho[...] = 1.0 + 0.1 * np.random.default_rng(2).random(f.rho.shape)
        return g, f, cases.params(dt=1e-3, amp=(0.1, 0.05))

    def run_case(name, kind, shape, stepper, method=None, nsteps=2, tol=0.0, **pkw):
        nx, ny, nzg = shape
        g, f, p = global_case(kind, nx, ny, nzg)
        if stepper != "projection":
            f.rho[...] = f.rho  # RK reads rho everywhere
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
        td.all_gather_object(parts, {k: getattr(loc, k) for k in ("u", "v", "w", "p", "T")})
        if rank == 0:
            out = abi.HostField(nx, ny, nzg)
            for r in range(world):
                lf = abi.HostField(loc.nx, loc.ny, loc.nz)
                for k, a in parts[r].items():
                    getattr(lf, k)[...] = a
                dist.gather_into(out, lf, world, r)
            worst = 0.0
            ok = (rc == rc_o)
            for k in ("u", "v", "w", "p", "T"):
                a, b = getattr(out, k), getattr(fo, k)
                if tol == 0.0:
                    same = np.array_equal(a, b)
                    worst = max(worst, float(np.max(np.abs(a - b))))
                    ok = ok and same
                else:
                    e = max(cases.rel_err(a, b)) if np.max(np.abs(b)) > 1e-12 else float(np.max(np.abs(a - b)))
                    worst = max(worst, e)
                    ok = ok and e <= tol
            print("%-34s world=%d rc=%d/%d worst=%.3e tol=%s %s" % (name, world, rc, rc_o, worst,
                                                                      "bitwise" if tol == 0.0 else "%.0e" % tol,
                                                                      "ok" if ok else "FAIL"), flush=True)
            if not ok:
                fails.append(name)
        td.barrier()

    nzg = 6 * world + 2      # 6 owned planes per rank
    nyg = 8 * world + 2
    run_case("projection+jacobi 3d", "random", (20, 14, nzg), "projection", "jacobi", max_iterations=40)
    run_case("projection+rbsor 3d", "random", (20, 14, nzg), "projection", "rbsor", max_iterations=30)
    run_case("projection+cg 3d", "random", (20, 14, nzg), "projection", "cg", tol=1e-8)
    run_case("projection+pcg 3d", "random", (20, 14, nzg), "projection", "pcg", tol=1e-8)
    run_case("projection+bicgstab x12 3d", "random", (20, 14, nzg), "projection", "bicgstab", tol=1e-8, max_iterations=12)
    run_case("rk4 3d (periodic ring)", "random", (20, 14, nzg), "rk4")
    run_case("rk2 3d (periodic ring)", "random", (20, 14, nzg), "rk2")
    run_case("projection+rbsor 2d y-slabs", "random", (22, nyg, 1), "projection", "rbsor", max_iterations=30)
    run_case("projection+pcg 2d y-slabs", "random", (22, nyg, 1), "projection", "pcg", tol=1e-8)
    run_case("rk4 2d y-slabs", "random", (22, nyg, 1), "rk4")
    run_case("explicit euler 3d", "random", (20, 14, nzg), "euler", nsteps=3)
    run_case("explicit euler 3d nx odd", "random", (21, 14, nzg), "euler", nsteps=3)
    run_case("explicit euler 2d y-slabs", "random", (22, nyg, 1), "euler", nsteps=3)
    run_case("explicit euler + energy 3d", "natconv", (14, 14, nzg), "euler", nsteps=3)
    # even nx: the TMA kernels (odd-nx grids above take the direct kernels... 20/22 are even: both covered)
    run_case("projection+jacobi 3d nx odd", "random", (21, 14, nzg), "projection", "jacobi", max_iterations=25)
    run_case("projection+cg 3d nx odd", "random", (21, 14, nzg), "projection", "cg", tol=1e-8)
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
        print("MULTIGPU PARITY OK")


if __name__ == "__main__":
    main()
