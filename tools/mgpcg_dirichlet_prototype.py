#!/usr/bin/env python
"""numpy / scipy experiment for DESIGN section 9 ("multigrid as the preconditioner of the reference's CG") -- CPU only,
second take.  tools/mgpcg_prototype.py preconditioned the reference's Krylov operator with the hierarchy of the
STATIONARY solvers (mirrored ghosts) and gained too little; this one builds the hierarchy for the Krylov operator
itself:

    A = -lap on the interior cells 1..n-2 of every axis, boundary cells of the vectors held at ZERO
    (linear_solver_cg.c:103-123: p, r, Ap are calloc'ed and only interior cells are ever written)

i.e. a homogeneous-Dirichlet Laplacian whose wall sits ON the boundary cell.  Coarsening keeps the even interior
cells of every axis (vertex-centred, spacing 2h); where n-1 is odd the last coarse cell ends up h instead of 2h away
from the wall.  Two coarse operators are compared:

  galerkin   A_c = P^T A P with tensor-product linear interpolation P by node POSITION (the wall value is 0), R = P^T
             -- symmetric by construction, 9-point (2-D) / 27-point (3-D) stencils on the coarse levels
  rediscr    5 / 7-point re-discretisation on the coarse nodes, the wall-distance row in its symmetric
             (finite-volume) form: 1-D operator rows  (u_i - u_{i-1}) / hl - (u_{i+1} - u_i) / hr  scaled by
             2 / (hl + hr) on BOTH sides, which is what a GPU kernel with per-axis coefficient tables would apply

Smoother: nu damped-Jacobi sweeps before and after (symmetric, so the V-cycle is a valid PCG preconditioner and
every pass is a plain stencil sweep).  Output: PCG iterations to 1e-6 relative residual for CG, Jacobi-PCG (the
reference's preconditioner) and one V(nu,nu) cycle per iteration, and the cycle's cost in fine-grid sweeps.

    python tools/mgpcg_dirichlet_prototype.py            # the table in DESIGN.md section 9
"""
import sys
import time

import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spla


def lap1d(x, wall_lo, wall_hi):
    """Symmetric 1-D operator -d2/dx2 on nodes x (interior), zero Dirichlet walls at wall_lo / wall_hi: the
    finite-volume form  (u_i - u_{i-1}) / hl + (u_i - u_{i+1}) / hr  and the diagonal volume matrix (hl + hr) / 2."""
    xe = np.concatenate([[wall_lo], x, [wall_hi]])
    hl, hr = xe[1:-1] - xe[:-2], xe[2:] - xe[1:-1]
    main = 1.0 / hl + 1.0 / hr
    off = -1.0 / hr[:-1]
    K = sp.diags([off, main, off], [-1, 0, 1], format="csr")
    M = sp.diags((hl + hr) * 0.5, format="csr")
    return K, M


def interp1d(xf, xc, wall_lo, wall_hi):
    """Linear interpolation from coarse nodes xc (plus the zero walls) to fine nodes xf, by position."""
    xe = np.concatenate([[wall_lo], xc, [wall_hi]])
    rows, cols, vals = [], [], []
    for i, x in enumerate(xf):
        k = np.searchsorted(xe, x, side="right") - 1
        k = min(max(k, 0), len(xe) - 2)
        t = (x - xe[k]) / (xe[k + 1] - xe[k])
        for kk, w in ((k, 1.0 - t), (k + 1, t)):
            if 1 <= kk <= len(xc) and w != 0.0:
                rows.append(i)
                cols.append(kk - 1)
                vals.append(w)
    return sp.csr_matrix((vals, (rows, cols)), shape=(len(xf), len(xc)))


def kron_all(mats):
    out = mats[0]
    for m in mats[1:]:
        out = sp.kron(out, m, format="csr")
    return out


class Level:
    pass


def build(ns, kind, min_nodes=3, coarse_sweeps=0, coupling=0.0):
    """Hierarchy for a box of ns = (nx, ny[, nz]) cells incl. boundary cells, unit box, h = 1 / (n - 1) per axis.
    The reference operator is the point-wise stencil (sum over axes of (2u - u- - u+) / h^2) = M^-1 K per axis on the
    uniform fine grid; the symmetric form K_total = sum_axes K_a (x) M_others is what the hierarchy works with
    (on the fine level it is the reference operator times the constant cell volume)."""
    dims = len(ns)
    xs = [np.arange(1, n - 1) / (n - 1.0) for n in ns]   # fine interior node positions
    levels = []
    while True:
        L = Level()
        L.xs = xs
        KM = [lap1d(x, 0.0, 1.0) for x in xs]
        L.K1 = [k for k, _ in KM]
        L.M1 = [m for _, m in KM]
        if kind == "rediscr" or not levels:
            terms = []
            for a in range(dims):
                terms.append(kron_all([L.K1[b] if b == a else L.M1[b] for b in range(dims)]))
            L.A = sum(terms[1:], terms[0]).tocsr()
        levels.append(L)
        if max(len(x) for x in xs) <= min_nodes:
            break
        # keep the even cells (index 2, 4, ... of 1..n-2) of every axis that still has more than min_nodes of them --
        # with coupling > 0 (semi-coarsening, what csrc/mgpc_cell.h does with 0.5) only of the axes whose 1 / H^2 is
        # at least that share of the strongest coarsenable one
        can = [len(x) > min_nodes for x in xs]
        strength = [1.0 / (x[1] - x[0]) ** 2 if len(x) > 1 else 0.0 for x in xs]
        cmax = max(s_ for s_, c in zip(strength, can) if c)
        xc = [x[1::2] if (c and s_ >= coupling * cmax) else x for x, c, s_ in zip(xs, can, strength)]
        L.P = kron_all([interp1d(xs[a], xc[a], 0.0, 1.0) if len(xc[a]) != len(xs[a]) else sp.identity(len(xs[a]), format="csr")
                        for a in range(dims)])
        xs = xc
    if kind == "galerkin":
        for l in range(1, len(levels)):
            P = levels[l - 1].P
            levels[l].A = (P.T @ levels[l - 1].A @ P).tocsr()
    for L in levels:
        L.dinv = 1.0 / L.A.diagonal()
    if coarse_sweeps:   # what a GPU block would do: a fixed number of damped-Jacobi sweeps (a symmetric operator)
        Lc = levels[-1]

        def sweeps(b, Lc=Lc):
            x = 0.85 * Lc.dinv * b
            for _ in range(coarse_sweeps - 1):
                x += 0.85 * Lc.dinv * (b - Lc.A @ x)
            return x
        Lc.solve = sweeps
    else:
        levels[-1].solve = spla.factorized(levels[-1].A.tocsc())
    return levels


def vcycle(levels, l, b, nu, omega):
    L = levels[l]
    if l == len(levels) - 1:
        return L.solve(b)
    x = omega * L.dinv * b
    for _ in range(nu - 1):
        x += omega * L.dinv * (b - L.A @ x)
    r = b - L.A @ x
    x += L.P @ vcycle(levels, l + 1, L.P.T @ r, nu, omega)
    for _ in range(nu):
        x += omega * L.dinv * (b - L.A @ x)
    return x


def pcg(A, b, prec, tol=1e-6, cap=20000):
    x = np.zeros_like(b)
    r = b.copy()
    z = prec(r)
    p = z.copy()
    rz = r @ z
    r0 = np.linalg.norm(r)
    for it in range(1, cap + 1):
        Ap = A @ p
        alpha = rz / (p @ Ap)
        x += alpha * p
        r -= alpha * Ap
        if np.linalg.norm(r) < tol * r0:
            return it
        z = prec(r)
        rz_new = r @ z
        p = z + (rz_new / rz) * p
        rz = rz_new
    return cap


def cycle_cost(levels, nu):
    """Fine-grid-sweep equivalents of one V(nu, nu) cycle: (2 nu smoothing + 1 residual + ~1 for the transfers) per
    level, weighted by the level's unknowns and by its stencil size relative to the fine one."""
    n0, nnz0 = levels[0].A.shape[0], levels[0].A.nnz / levels[0].A.shape[0]
    return sum((2 * nu + 2) * (L.A.shape[0] / n0) * ((L.A.nnz / L.A.shape[0]) / nnz0) for L in levels[:-1])


def main():
    shapes = [(129, 129), (130, 130), (513, 513), (514, 514), (1026, 1026), (65, 65, 65), (66, 66, 66), (130, 130, 130)]
    if len(sys.argv) > 1:
        shapes = [tuple(int(v) for v in a.split(",")) for a in sys.argv[1:]]
    rng = np.random.default_rng(3)
    print("%-16s %6s %6s | %-9s %5s %6s %6s | %s" % ("grid", "CG", "J-PCG", "coarse op", "nu", "iters", "cost", "CG-iteration equivalents"))
    for ns in shapes:
        t0 = time.time()
        fine = build(ns, "rediscr")[0]
        A = fine.A
        b = rng.standard_normal(A.shape[0])
        it_cg = pcg(A, b, lambda r: r)
        dinv = 1.0 / A.diagonal()
        it_j = pcg(A, b, lambda r: dinv * r)
        for kind in ("galerkin", "rediscr"):
            levels = build(ns, kind)
            for nu in (1, 2):
                it = pcg(A, b, lambda r: vcycle(levels, 0, r, nu, 0.8 if len(ns) == 2 else 0.85))
                cost = cycle_cost(levels, nu)
                # a CG iteration moves ~80 B/cell (dir + update kernels), a stencil sweep of the cycle ~24 B/cell
                equiv = it * (1.0 + cost * 24.0 / 80.0)
                print("%-16s %6d %6d | %-9s %5d %6d %6.1f | %.0f   (%.1fx fewer bytes than CG)" %
                      ("x".join(map(str, ns)), it_cg, it_j, kind, nu, it, cost, equiv, it_cg / equiv))
        print("    (%.1f s)" % (time.time() - t0), flush=True)


if __name__ == "__main__":
    main()
