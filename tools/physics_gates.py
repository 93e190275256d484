#!/usr/bin/env python
"""The reference's long physics gates, run once per round on the GPU path and (for the numbers kept under
tests/golden/physics_gates_oracle.json) on the CPU oracle:

  ghia129   lid-driven cavity Re = 100, 129 x 129, 50 000 steps, dt = 2e-4 -- the CAVITY_FULL_VALIDATION run of
            tests/validation/test_cavity_backends.c:57-59; gate RMS_u, RMS_v < 0.10 against Ghia et al. (1982)
  dvd       de Vahl Davis natural convection Ra = 1e3, 41 x 41, dt = 2e-3, marched to a kinetic-energy residual of
            1e-6 -- tests/validation/test_natural_convection.c:139-307; gates u_max*, v_max*, Nu within 10 %

    python tools/physics_gates.py --impl gpu|oracle [--case ghia129|dvd|all] [--steps N]

One JSON line per case.  The GPU runs go through the ns_solver_t "projection_gpu" object (the reference's call
sequence: BCs on the host, solver_step, repeat); ghia129 uses CFD_B200_SYNC_DOWNLOAD_LAZY -- its boundary values
never change, so nothing crosses PCIe until the final download."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402

GHIA_Y = [0.0000, 0.0547, 0.0625, 0.0703, 0.1016, 0.1719, 0.2813, 0.4531, 0.5000, 0.6172, 0.7344, 0.8516, 0.9531,
          0.9609, 0.9688, 0.9766, 1.0000]
GHIA_U = [0.00000, -0.03717, -0.04192, -0.04775, -0.06434, -0.10150, -0.15662, -0.21090, -0.20581, -0.13641, 0.00332,
          0.23151, 0.68717, 0.73722, 0.78871, 0.84123, 1.00000]
GHIA_X = [0.0000, 0.0625, 0.0703, 0.0781, 0.0938, 0.1563, 0.2266, 0.2344, 0.5000, 0.8047, 0.8594, 0.9063, 0.9453,
          0.9531, 0.9609, 0.9688, 1.0000]
GHIA_V = [0.00000, 0.09233, 0.10091, 0.10890, 0.12317, 0.16077, 0.17507, 0.17527, 0.05454, -0.24533, -0.22445,
          -0.16914, -0.10313, -0.08864, -0.07391, -0.05906, 0.00000]


def ghia_rms(f, n):
    c = np.linspace(0.0, 1.0, n)
    u_mid, v_mid = f.u[0, :, n // 2], f.v[0, n // 2, :]
    ru = float(np.sqrt(np.mean((np.interp(GHIA_Y, c, u_mid) - np.array(GHIA_U)) ** 2)))
    rv = float(np.sqrt(np.mean((np.interp(GHIA_X, c, v_mid) - np.array(GHIA_V)) ** 2)))
    return ru, rv


def ghia129(impl, steps, n=129, dt=2e-4):
    import cases
    from cfd_b200 import abi
    g, f, p = cases.cavity(n, 1, dt=dt)
    pc = abi.poisson_config("cg", on_nonconvergence=abi.CFD_B200_NONCONV_CONTINUE)
    cases.apply_cavity_bc(f)
    t0 = time.time()
    if impl == "gpu":
        from cfd_b200 import api
        api.set_poisson(pc)
        slv = api.GpuSolver("projection_gpu")
        slv.set_host_sync(api.SYNC_DOWNLOAD_LAZY)
        for step in range(steps):
            rc, _ = slv.step(f, g, p)
            assert rc == 0, (step, rc)
        slv.sync_host(f)
        slv.close()
    else:
        import oracle
        orc = oracle.Oracle()
        for step in range(steps):
            cases.apply_cavity_bc(f)
            rc, _ = orc.projection(f, g, p, pc, 1)
            assert rc == 0, (step, rc)
    ru, rv = ghia_rms(f, n)
    return {"case": "ghia129", "impl": impl, "grid": n, "steps": steps, "dt": dt, "rms_u": ru, "rms_v": rv,
            "gate": 0.10, "pass": bool(ru < 0.10 and rv < 0.10), "u_center": float(f.u[0, n // 2, n // 2]),
            "seconds": round(time.time() - t0, 1)}


def dvd(impl, max_steps=30000, n=41, dt=2e-3):
    import cases
    from cfd_b200 import abi
    g, f, p = cases.natural_convection(n, 1, dt=dt)
    pc = abi.poisson_config("cg", on_nonconvergence=abi.CFD_B200_NONCONV_CONTINUE)
    alpha = p.alpha

    def wall_bcs():
        for a in (f.u, f.v):
            a[0, :, 0] = a[0, :, -1] = 0.0
            a[0, 0, :] = a[0, -1, :] = 0.0

    def ke():
        return 0.5 * float(np.sum(f.u[0] ** 2 + f.v[0] ** 2))

    t0 = time.time()
    if impl == "gpu":
        from cfd_b200 import api
        api.set_poisson(pc)
        slv = api.GpuSolver("projection_gpu")
        step_fn = lambda: slv.step(f, g, p)[0]
    else:
        import oracle
        orc = oracle.Oracle()
        step_fn = lambda: orc.projection(f, g, p, pc, 1)[0]
    prev, done, conv = ke(), 0, False
    for step in range(max_steps):
        wall_bcs()
        assert step_fn() == 0, step
        wall_bcs()
        k = ke()
        res = abs(k - prev) / (prev + 1e-10)
        prev, done = k, step + 1
        if step > 200 and res < 1e-6:
            conv = True
            break
    if impl == "gpu":
        slv.close()
    scale = 1.0 / alpha
    umax = float(np.max(np.abs(f.u[0, :, n // 2]))) * scale
    vmax = float(np.max(np.abs(f.v[0, n // 2, :]))) * scale
    dx = 1.0 / (n - 1)
    Ts = (f.T[0] - 290.0) / 20.0
    nu_loc = -(-3.0 * Ts[:, 0] + 4.0 * Ts[:, 1] - Ts[:, 2]) / (2.0 * dx)
    w = np.ones(n)
    w[0] = w[-1] = 0.5
    nu = float(np.sum(w * nu_loc) * dx)
    ref = {"u": 3.649, "v": 3.697, "nu": 1.117}
    err = {"u": abs(umax - ref["u"]) / ref["u"], "v": abs(vmax - ref["v"]) / ref["v"], "nu": abs(nu - ref["nu"]) / ref["nu"]}
    return {"case": "dvd_ra1e3", "impl": impl, "grid": n, "dt": dt, "steps": done, "converged": conv, "u_max": umax,
            "v_max": vmax, "Nu": nu, "rel_err": err, "gate": 0.10, "pass": bool(conv and max(err.values()) < 0.10),
            "seconds": round(time.time() - t0, 1)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--impl", default="gpu", choices=["gpu", "oracle"])
    ap.add_argument("--case", default="all", choices=["all", "ghia129", "dvd"])
    ap.add_argument("--steps", type=int, default=50000)
    a = ap.parse_args()
    if a.case in ("all", "dvd"):
        print(json.dumps(dvd(a.impl)), flush=True)
    if a.case in ("all", "ghia129"):
        print(json.dumps(ghia129(a.impl, a.steps)), flush=True)


if __name__ == "__main__":
    main()
