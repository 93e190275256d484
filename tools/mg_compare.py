#!/usr/bin/env python
"""Time-to-solution of the multigrid solver against the stationary methods it shares its discrete system
with (Jacobi / Red-Black SOR: mirrored boundary cells), on one GPU, device-resident.

    python tools/mg_compare.py --shapes 129,129,1 256,256,256 --methods multigrid,rbsor

x0 = 0, zero-mean random right-hand side, the reference's default stopping rule (1e-6 relative on the
L-infinity residual, linear_solver.c:37-47).  Reports iterations (V-cycles for multigrid), the summed
kernel time from the per-launch CUDA events (cfd_b200_session_profile) and the wall time of the solve
call with the operands already on the device.
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shapes", nargs="+", default=["129,129,1", "256,256,256"])
    ap.add_argument("--methods", default="multigrid,rbsor")
    ap.add_argument("--tol", type=float, default=1e-6)
    ap.add_argument("--cap", type=int, default=20000)
    a = ap.parse_args()
    import numpy as np
    from cfd_b200 import abi, api
    for shape in a.shapes:
        nx, ny, nz = (int(v) for v in shape.split(","))
        hx, hy, hz = 1.0 / (nx - 1), 1.0 / (ny - 1), (1.0 / (nz - 1) if nz > 1 else 0.0)
        rhs = abi.aligned_zeros(nx * ny * nz).reshape(nz, ny, nx)
        sl = (slice(1, -1) if nz > 1 else slice(None), slice(1, -1), slice(1, -1))
        rng = np.random.default_rng(5)
        for k in range(rhs.shape[0]):
            if nz == 1 or 0 < k < nz - 1:
                rhs[k, 1:-1, 1:-1] = rng.standard_normal((ny - 2, nx - 2))
        rhs[sl] -= rhs[sl].mean()
        x0 = abi.aligned_zeros(nx * ny * nz).reshape(nz, ny, nx)
        for m in a.methods.split(","):
            s = api.Session(nx, ny, nz, hx, hy, hz)
            pc = abi.poisson_config(m, tolerance=a.tol, max_iterations=(200 if m == "multigrid" else a.cap))
            s.poisson(pc, x0, rhs)  # warm-up: allocations, hierarchy, module load
            s.sync()

            def solve():
                api.check(s.L.cfd_b200_session_upload_poisson(s.h, abi._ptr(x0), abi._ptr(rhs)), "upload_poisson")
                s.sync()
                t0 = time.perf_counter()
                rc, _ = s.poisson(pc)
                s.sync()
                return rc, time.perf_counter() - t0
            best = min(solve()[1] for _ in range(3))  # as a user runs it (multigrid: one CUDA graph per cycle)
            s.reset_kernel_stats()
            s.profile(True)                           # per-launch events: individual launches
            for _ in range(2):
                rc, _ = solve()
            s.profile(False)
            st = s.last_poisson_stats()
            ks = [k for k in s.kernel_stats() if k["total_ms"] > 0]
            out = {"shape": [nx, ny, nz], "method": m, "rc": rc,
                   "iterations": st.iterations, "initial_residual": st.initial_residual,
                   "final_residual": st.final_residual, "solve_wall_ms": round(best * 1e3, 3),
                   "kernel_ms_per_solve": round(sum(k["total_ms"] for k in ks) / 2, 3),
                   "launches_per_solve": sum(k["launches"] for k in ks) // 2,
                   "kernels": {k["name"]: [k["launches"] // 2, round(k["total_ms"] / 2, 3)] for k in ks}}
            print(json.dumps(out), flush=True)
            s.close()


if __name__ == "__main__":
    main()
