#!/usr/bin/env python
"""ref_gpu_compare.py -- the reference's OWN CUDA backend timed beside cfd_b200 on the same B200.

    python tools/ref_gpu_compare.py [--n 512] [--steps 4] [--iters 50] [--out gpurun_out/ref_gpu_compare.jsonl]

SURVEY 8(d): "secondary comparator on the GPU: the reference's own projection_gpu / rk4_gpu built
for sm_100, same inputs, via solve()".  oracle/_ref/libcfd_ref_cuda.so is the UNMODIFIED reference
backend (oracle/Makefile: refcuda, Release flags); libcfd_b200.so is ours.  Both are driven through
the SAME C symbols with the SAME host flow_field:

  * solve_projection_method_gpu(field, grid, params{max_iter = S}, config): the reference runs CG
    (its only Poisson method inside the step) capped at `--iters` iterations with tolerance 0
    (gpu_config_t.poisson_max_iter / poisson_tolerance, solver_projection_gpu.cu:703); ours runs
    CG with the same fixed count.  Called with S = 1 and S = 1 + steps: the difference is the
    device-resident cost of `steps` steps (upload / download / allocation cancel), the S = 1 call
    is the end-to-end single-step time.
  * solve_rk4_method_gpu, same scheme.
  * the four Poisson solver objects create_*_gpu_solver (init + solve on host x / rhs, fixed
    iteration count).

Every run is its own process (the two libraries export the same symbol names).  This is a TIMING
comparison only: the reference's GPU kernels are not a parity oracle (they drop the source terms
and use 1/dt for rho/dt); parity is pinned against its CPU path elsewhere.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

LIBS = {"reference_cuda": os.path.join(ROOT, "oracle", "_ref", "libcfd_ref_cuda.so"),
        "cfd_b200": os.path.join(ROOT, "cfd_b200", "libcfd_b200.so")}


def host_state(n):
    import numpy as np
    from cfd_b200 import abi
    g = abi.HostGrid(n, n, n, 0.0, 1.0, 0.0, 1.0, 0.0, 1.0)
    f = abi.HostField(n, n, n)
    x = np.linspace(0.0, 1.0, n)
    sx, cx = np.sin(2 * np.pi * x), np.cos(2 * np.pi * x)
    plane_u, plane_v = 0.1 * np.outer(cx, sx), -0.1 * np.outer(sx, cx)
    ripple = 1e-3 * np.sin(37.0 * np.pi * x)[None, :] * np.cos(29.0 * np.pi * x)[:, None]
    for k in range(n):
        f.u[k] = cx[k] * plane_u + ripple
        f.v[k] = cx[k] * plane_v - ripple
        f.w[k] = sx[k] * ripple
    f.p[...] = 0.0
    f.rho[...] = 1.0
    f.T[...] = 300.0
    return g, f


def worker(impl, what, n, steps, iters):
    import numpy as np
    from cfd_b200 import abi
    L = C.CDLL(LIBS[impl], mode=C.RTLD_GLOBAL)
    FF, GR, NP, GC = (C.POINTER(abi.FlowField), C.POINTER(abi.Grid), C.POINTER(abi.NSSolverParams),
                      C.POINTER(abi.GpuConfig))
    L.gpu_config_default.restype = abi.GpuConfig
    cfg = L.gpu_config_default()
    cfg.min_grid_size, cfg.min_steps = 1, 1
    cfg.poisson_max_iter, cfg.poisson_tolerance = iters, 0.0
    if impl == "cfd_b200":
        pc = abi.poisson_config("cg", max_iterations=iters, tolerance=0.0, absolute_tolerance=0.0,
                                on_nonconvergence=abi.CFD_B200_NONCONV_CONTINUE)
        L.cfd_b200_set_poisson_config.argtypes = [C.POINTER(abi.PoissonConfig)]
        assert L.cfd_b200_set_poisson_config(C.byref(pc)) == 0
    out = {"impl": impl, "workload": what, "n": n, "cells": n ** 3}
    if what in ("projection", "rk4", "rk2", "explicit_euler"):
        fn = getattr(L, "solve_%s_method_gpu" % what)
        fn.restype, fn.argtypes = C.c_int, [FF, GR, NP, GC]
        g, f0 = host_state(n)
        p = abi.ns_params_default()
        p.dt, p.mu = 1e-4, 0.01
        p.source_amplitude_u = p.source_amplitude_v = 0.0

        def call(S):
            f = f0.copy()
            p.max_iter = S
            t0 = time.perf_counter()
            rc = fn(C.byref(f.c), C.byref(g.c), C.byref(p), C.byref(cfg))
            dt = time.perf_counter() - t0
            if rc != 0:
                raise RuntimeError("%s %s rc=%d" % (impl, what, rc))
            assert np.all(np.isfinite(f.u[n // 2]))
            return dt
        call(1)  # warm-up: context, allocations, cached session
        t1 = min(call(1), call(1))
        tS = min(call(1 + steps), call(1 + steps))
        dev = (tS - t1) / steps
        out.update({"poisson": "cg x%d (fixed)" % iters if what == "projection" else None, "steps": steps,
                    "e2e_single_step_ms": t1 * 1e3, "e2e_single_step_mlups": n ** 3 / t1 / 1e6,
                    "device_ms_per_step": dev * 1e3, "device_mlups": n ** 3 / dev / 1e6})
    else:  # stand-alone Poisson object: poisson_<method>
        meth = {"jacobi": (abi.POISSON_METHOD_JACOBI, "create_jacobi_gpu_solver"),
                "rbsor": (abi.POISSON_METHOD_REDBLACK_SOR, "create_redblack_gpu_solver"),
                "cg": (abi.POISSON_METHOD_CG, "create_cg_gpu_solver"),
                "bicgstab": (abi.POISSON_METHOD_BICGSTAB, "create_bicgstab_gpu_solver")}[what.split("_")[1]]
        fac = getattr(L, meth[1])
        fac.restype = C.POINTER(abi.PoissonSolver)
        PSP = C.POINTER(abi.PoissonSolver)
        h = 1.0 / (n - 1)
        rng = np.random.default_rng(5)
        rhs = abi.aligned_zeros(n ** 3).reshape(n, n, n)
        for k in range(1, n - 1):
            rhs[k, 1:-1, 1:-1] = rng.standard_normal((n - 2, n - 2))
        rhs[1:-1, 1:-1, 1:-1] -= rhs[1:-1, 1:-1, 1:-1].mean()
        x0 = abi.aligned_zeros(n ** 3).reshape(n, n, n)
        xt = abi.aligned_zeros(n ** 3).reshape(n, n, n)

        def solve(nit):
            """poisson_solver_create + _init + _solve + _destroy (linear_solver.c:150-298), solve timed."""
            s = fac()
            assert s, "factory returned NULL"
            sv = s.contents
            pp = abi.poisson_params_default()
            pp.max_iterations, pp.tolerance, pp.absolute_tolerance, pp.check_interval = nit, 0.0, 0.0, 1
            sv.nx, sv.ny, sv.nz, sv.dx, sv.dy, sv.dz, sv.params = n, n, n, h, h, h, pp
            init = C.CFUNCTYPE(C.c_int, PSP, C.c_size_t, C.c_size_t, C.c_size_t, C.c_double, C.c_double,
                               C.c_double, C.POINTER(abi.PoissonParams))(sv.init)
            rc = init(s, n, n, n, h, h, h, C.byref(sv.params))
            assert rc == 0, "init rc=%d" % rc
            x = x0.copy()
            st = abi.PoissonStats()
            fn = C.CFUNCTYPE(C.c_int, PSP, abi.c_double_p, abi.c_double_p, abi.c_double_p,
                             C.POINTER(abi.PoissonStats))(sv.solve)
            t0 = time.perf_counter()
            fn(s, abi._ptr(x), abi._ptr(xt), abi._ptr(rhs), C.byref(st))
            dt = time.perf_counter() - t0
            C.CFUNCTYPE(None, PSP)(sv.destroy)(s)
            C.CDLL(None).free(C.cast(s, C.c_void_p))
            return dt, st.iterations
        solve(2)
        # solve time is linear in the iteration count: least-squares slope over four counts (two samples each)
        counts = [iters, 2 * iters, 3 * iters, 5 * iters]
        pts = []
        for c in counts:
            t, it = min(solve(c), solve(c))
            pts.append((it, t))
        mx, my = sum(p[0] for p in pts) / len(pts), sum(p[1] for p in pts) / len(pts)
        per = sum((p[0] - mx) * (p[1] - my) for p in pts) / sum((p[0] - mx) ** 2 for p in pts)
        out.update({"iterations": [p[0] for p in pts], "solve_ms": [round(p[1] * 1e3, 2) for p in pts],
                    "ms_per_iteration": per * 1e3, "mlups_per_iteration": n ** 3 / per / 1e6})
    print("RESULT " + json.dumps(out), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=512)
    ap.add_argument("--steps", type=int, default=4)
    ap.add_argument("--iters", type=int, default=50)
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "ref_gpu_compare.jsonl"))
    ap.add_argument("--workloads", default="projection,rk4,poisson_cg,poisson_bicgstab,poisson_jacobi,poisson_rbsor")
    ap.add_argument("--worker", nargs=2)
    a = ap.parse_args()
    if a.worker:
        worker(a.worker[0], a.worker[1], a.n, a.steps, a.iters)
        return
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    rows = []
    for what in a.workloads.split(","):
        for impl in ("reference_cuda", "cfd_b200"):
            if not os.path.exists(LIBS[impl]):
                rows.append({"impl": impl, "workload": what, "error": "library not built: " + LIBS[impl]})
                continue
            r = subprocess.run([sys.executable, os.path.abspath(__file__), "--n", str(a.n), "--steps", str(a.steps),
                                "--iters", str(a.iters), "--worker", impl, what], capture_output=True, text=True,
                               timeout=900)
            got = [ln for ln in r.stdout.splitlines() if ln.startswith("RESULT ")]
            if r.returncode != 0 or not got:
                rows.append({"impl": impl, "workload": what, "error": (r.stderr or r.stdout)[-400:]})
            else:
                rows.append(json.loads(got[-1][7:]))
            print(json.dumps(rows[-1]), flush=True)
    with open(a.out, "w") as fh:
        for r in rows:
            fh.write(json.dumps(r) + "\n")


if __name__ == "__main__":
    main()
