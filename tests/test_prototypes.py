"""The CPU prototypes behind DESIGN.md section 9 keep producing the numbers the design decisions quote."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tools"))


def test_mgpcg_on_the_krylov_operator_needs_a_handful_of_iterations():
    """Multigrid built for the reference's Krylov operator (boundary cells of the vectors held at zero,
    linear_solver_cg.c:103-123) as the PCG preconditioner: <= 8 iterations to 1e-6 on odd AND even grids, 2-D and
    3-D, with the re-discretised (5/7-point, wall-distance row) and the Galerkin coarse operators; plain CG (= the
    reference's Jacobi-PCG: constant diagonal) needs hundreds."""
    import mgpcg_dirichlet_prototype as mp
    rng = np.random.default_rng(3)
    for ns, cg_min in (((65, 65), 120), ((66, 66), 120), ((33, 33, 33), 60), ((34, 34, 34), 60)):
        A = mp.build(ns, "rediscr")[0].A
        b = rng.standard_normal(A.shape[0])
        it_cg = mp.pcg(A, b, lambda r: r)
        dinv = 1.0 / A.diagonal()
        assert mp.pcg(A, b, lambda r: dinv * r) == it_cg  # constant diagonal: Jacobi-PCG is CG
        assert it_cg >= cg_min, (ns, it_cg)
        for kind in ("rediscr", "galerkin"):
            levels = mp.build(ns, kind)
            it = mp.pcg(A, b, lambda r: mp.vcycle(levels, 0, r, 1, 0.8 if len(ns) == 2 else 0.85))
            assert it <= 8, (ns, kind, it)
            assert mp.cycle_cost(levels, 1) < 7.0


def test_the_fine_operator_of_the_prototype_is_the_reference_stencil():
    """K_total = cell volume x (the reference's -lap on interior cells with zero boundary cells)."""
    import mgpcg_dirichlet_prototype as mp
    nx, ny = 9, 7
    A = mp.build((nx, ny), "rediscr")[0].A.toarray()
    hx, hy = 1.0 / (nx - 1), 1.0 / (ny - 1)
    rng = np.random.default_rng(0)
    p = np.zeros((nx, ny))
    p[1:-1, 1:-1] = rng.standard_normal((nx - 2, ny - 2))
    ref = (2 * p[1:-1, 1:-1] - p[:-2, 1:-1] - p[2:, 1:-1]) / hx ** 2 + (2 * p[1:-1, 1:-1] - p[1:-1, :-2] - p[1:-1, 2:]) / hy ** 2
    got = (A @ p[1:-1, 1:-1].ravel()).reshape(nx - 2, ny - 2) / (hx * hy)
    np.testing.assert_allclose(got, ref, rtol=1e-12, atol=1e-9)
