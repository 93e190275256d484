#!/usr/bin/env python
"""numpy experiment behind DESIGN section 9 ("multigrid as the preconditioner of the reference's CG") -- CPU only.

PCG on the reference's Krylov operator (A = -lap on interior cells, boundary cells of the vectors held at zero,
linear_solver_cg.c:103-123) preconditioned by ONE V(2,2) cycle of the EXISTING hierarchy (mirrored ghosts, the system of
the stationary solvers; tools/mg_prototype.py), with the plain colour order ("mg") and with the post-smoothing in
reversed colour order ("mg_sym").  Iterations to 1e-6 relative, measured here:
    129^2  CG 322   mg  80   mg_sym  35        130^2  CG 324   mg  97   mg_sym  93
    513^2  CG 1245  mg 147   mg_sym  64        258^2  CG 638   mg 134   mg_sym 125
    65^3   CG 182   mg  63   mg_sym  34        66^3   CG 184   mg  66   mg_sym  65
A preconditioned iteration costs a V-cycle (~10 fine-grid sweeps), so 3-20x fewer iterations is no time-to-solution win:
the boundary treatments do not match.  A hierarchy for the Krylov operator itself (zero ghosts, wall-distance rows) is
what the row needs: tools/mgpcg_dirichlet_prototype.py (7-8 iterations on every grid).
"""
import os
import sys
import time
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import numpy as np
import mg_prototype as mp

def sweep_order(L, x, b, order):
    for col in order:
        mp.mirror(x)
        xn = (L.nbsum(x) - b[L.inner]) / L.diag
        x[L.inner] = np.where(L.colour == col, xn, x[L.inner])
    mp.mirror(x)

def vcycle_sym(levels, l, x, b, sym):
    L = levels[l]
    if L.kinds is None:
        for _ in range(15):
            sweep_order(L, x, b, (0, 1))
        for _ in range(15):
            sweep_order(L, x, b, (1, 0) if sym else (0, 1))
        return
    sweep_order(L, x, b, (0, 1)); sweep_order(L, x, b, (0, 1))
    rc = L.residual(x, b)
    for ax, k in enumerate(L.kinds):
        if k:
            rc = mp.restrict1(rc, ax, k, L.vol[ax], levels[l + 1].vol[ax], False)
    E = np.zeros(levels[l + 1].shape)
    vcycle_sym(levels, l + 1, E, rc, sym)
    mp.mirror(E)
    e = E
    for ax, k in enumerate(L.kinds):
        if k:
            e = mp.prolong1(e, ax, k, L.shape[ax] - 2)
    x[L.inner] += e[L.inner]
    o = (1, 0) if sym else (0, 1)
    sweep_order(L, x, b, o); sweep_order(L, x, b, o)

def A_krylov(L, p):
    """A = -lap on interior cells, boundary cells of p are zero (linear_solver_cg.c:103-123)."""
    q = np.zeros_like(p)
    pz = p.copy()
    for ax in range(p.ndim):
        if not L.act[ax]: continue
        s = [slice(None)] * p.ndim
        s[ax] = 0; pz[tuple(s)] = 0
        s[ax] = -1; pz[tuple(s)] = 0
    lap = 0.0
    for ax in range(p.ndim):
        if not L.act[ax]: continue
        pl, mi = list(L.inner), list(L.inner)
        pl[ax], mi[ax] = slice(2, None), slice(0, -2)
        lap = lap + (pz[tuple(pl)] - 2 * pz[L.inner] + pz[tuple(mi)]) / L.h[ax] ** 2
    q[L.inner] = -lap
    return q

def pcg(shape, mode, tol=1e-6, maxit=3000):
    rng = np.random.default_rng(0)
    h = [1.0 / (n - 1) if n > 1 else 1.0 for n in shape]
    levels = mp.build(shape, h, True)
    L = levels[0]
    b = np.zeros(shape); b[L.inner] = rng.standard_normal(b[L.inner].shape)
    x = np.zeros(shape)
    r = b - A_krylov(L, x)
    def prec(r):
        if mode == "none": return r.copy()
        y = np.zeros(shape)
        vcycle_sym(levels, 0, y, -r, mode == "mg_sym")   # lap y = -r  =>  y ~ A^-1 r
        z = np.zeros(shape); z[L.inner] = y[L.inner]
        return z
    z = prec(r); p = z.copy()
    rz = np.vdot(r, z); r0 = np.linalg.norm(r)
    for it in range(maxit):
        Ap = A_krylov(L, p)
        alpha = rz / np.vdot(p, Ap)
        x += alpha * p; r -= alpha * Ap
        res = np.linalg.norm(r)
        if res < tol * r0:
            return it + 1, res / r0
        z = prec(r)
        rz_new = np.vdot(r, z)
        p = z + (rz_new / rz) * p
        rz = rz_new
    return maxit, res / r0

for shape in [(1, 129, 129), (1, 130, 130), (1, 258, 258), (1, 513, 513), (34, 34, 34), (66, 66, 66), (65, 65, 65)]:
    out = []
    for mode in ("none", "mg", "mg_sym"):
        if mode == "none" and np.prod(shape) > 3e5 and len([n for n in shape if n > 1]) == 3: 
            pass
        t = time.time(); it, rr = pcg(shape, mode); out.append("%s: %4d it (%.1e, %.1fs)" % (mode, it, rr, time.time() - t))
    print(shape, " | ".join(out), flush=True)
