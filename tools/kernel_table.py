#!/usr/bin/env python
"""Per-kernel roofline table for any grid shape / stepper on ONE GPU (exploration companion of
bench.py; same timing method: per-launch CUDA events on the session stream, inputs larger than L2
unless the grid itself is small -- flagged as cache-resident then).

    python tools/kernel_table.py --shape 4096,4096,1 --method pcg --iters 100
    python tools/kernel_table.py --shape 128,128,128 --stepper rk4 --steps 20
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="512,512,512")
    ap.add_argument("--stepper", default="projection", choices=["projection", "euler", "rk2", "rk4", "poisson"])
    ap.add_argument("--method", default="bicgstab")
    ap.add_argument("--iters", type=int, default=50)
    ap.add_argument("--tol", type=float, default=0.0,
                    help="relative Poisson tolerance (0 = run exactly --iters iterations; > 0: time to solution, --iters is the cap)")
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--rhs", default="random", choices=["random", "smooth"], help="--stepper poisson: right-hand side")
    ap.add_argument("--energy", action="store_true", help="alpha > 0, beta != 0 (natural-convection kernels)")
    ap.add_argument("--fast", action="store_true", help="the fast arithmetic build (cfd_b200_set_arithmetic(1))")
    ap.add_argument("--no-profile", action="store_true", help="no per-launch events: whole-step time only")
    a = ap.parse_args()
    import torch
    from cfd_b200 import abi, api, lib
    nx, ny, nz = (int(v) for v in a.shape.split(","))
    api.set_arithmetic(1 if a.fast else 0)
    hx, hy, hz = 1.0 / (nx - 1), 1.0 / (ny - 1), (1.0 / (nz - 1) if nz > 1 else 0.0)
    s = api.Session(nx, ny, nz, hx, hy, hz)
    s.init_synthetic(12345, 0, nz)
    p = abi.ns_params_default()
    p.dt, p.mu = 1e-4, 0.01
    p.source_amplitude_u = p.source_amplitude_v = 0.0
    if a.energy:
        p.alpha, p.beta, p.T_ref = 3.035e-2, 3.333e-3, 300.0
        p.gravity[1] = -9.81
        p.dt = 2.5e-7
    pc = abi.poisson_config(a.method, max_iterations=a.iters, tolerance=a.tol, absolute_tolerance=0.0 if a.tol == 0.0 else 1e-300,
                            on_nonconvergence=abi.CFD_B200_NONCONV_CONTINUE)

    if a.stepper == "poisson":  # stand-alone solves: x = 0 and a zero-mean rhs, uploaded once per solve
        import numpy as np
        rhs = abi.aligned_zeros(nx * ny * nz).reshape(nz, ny, nx)
        sl = (slice(1, -1) if nz > 1 else slice(None), slice(1, -1), slice(1, -1))
        if a.rhs == "random":
            rhs[sl] = np.random.default_rng(5).standard_normal(rhs[sl].shape)
        else:
            xs = np.linspace(0.0, 1.0, nx)
            rhs[sl] = (np.cos(np.pi * xs)[None, None, :] * np.ones((nz, ny, 1)))[sl]
        rhs[sl] -= rhs[sl].mean()
        x0 = abi.aligned_zeros(nx * ny * nz).reshape(nz, ny, nx)

    def run(k):
        if a.stepper == "projection":
            rc, _ = s.projection(p, pc, k)
        elif a.stepper == "poisson":
            rc = 0
            for _ in range(k):
                s.poisson(pc, x0, rhs)
        elif a.stepper == "euler":
            rc, _ = s.rk(1, p, k)
        else:
            rc, _ = s.rk(int(a.stepper[2]), p, k)
        if rc != 0:
            raise RuntimeError("step failed rc=%d: %s" % (rc, lib.last_error()))

    run(a.warmup)
    s.reset_kernel_stats()
    s.profile(not a.no_profile)
    st = torch.cuda.ExternalStream(s.stream())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record(st)
    run(a.steps)
    e1.record(st)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    s.profile(False)
    interior = (nx - 2) * (ny - 2) * ((nz - 2) if nz > 1 else 1)
    peak = 6548.2
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    rows = []
    for k in sorted(s.kernel_stats(), key=lambda k: -k["total_ms"]):
        if k["total_ms"] <= 0:
            continue
        avg = k["total_ms"] / k["launches"]
        gbs = k["bytes_per_cell"] * interior / (avg * 1e-3) / 1e9
        rows.append({"kernel": k["name"], "launches": k["launches"], "avg_us": round(avg * 1e3, 2),
                     "bytes_per_cell": k["bytes_per_cell"], "GBps": round(gbs, 1), "frac_of_measured_peak": round(gbs / peak, 4)})
    out = {"shape": [nx, ny, nz], "arithmetic": "fast" if a.fast else "exact", "stepper": a.stepper, "method": a.method, "iters": a.iters, "steps": a.steps,
           "tol": a.tol, "poisson_iterations": s.last_poisson_stats().iterations, "ms_per_step": ms / a.steps, "MLUPS": nx * ny * nz * a.steps / (ms * 1e-3) / 1e6,
           "env": {k: v for k, v in os.environ.items() if k.startswith("CFD_B200_")},
           "field_MB": 8 * nx * ny * nz / 1e6, "cache_resident": 8 * nx * ny * nz * 6 < 126e6, "kernels": rows}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
