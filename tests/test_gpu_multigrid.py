"""Geometric multigrid (POISSON_METHOD_MULTIGRID) on the GPU.

The reference reserves the method name without an implementation (poisson_solver.h:60), so there is
no reference output: parity here is against the CPU statement of the same algorithm in
oracle/oracle_mg.inc (bit for bit -- the only reduction is a maximum), and the ANSWER is anchored to
the reference through the pinned Red-Black SOR restatement, which solves the same discrete system
(tests/test_oracle.py::test_multigrid_solves_the_reference_system does the same on the CPU).
"""
import numpy as np
import pytest

import cases
from cfd_b200 import abi, api

pytestmark = pytest.mark.gpu

MG = abi.POISSON_METHOD_MULTIGRID


def mg_params(**kw):
    pp = abi.poisson_params_default()
    pp.max_iterations = 100
    for k, v in kw.items():
        setattr(pp, k, v)
    return pp


def gpu_mg(g, x, rhs, pp):
    dx, dy, dz = g.spacing
    s = api.PoissonGpuSolver(MG, g.nx, g.ny, g.nz, dx, dy, dz, pp)
    try:
        return s.solve(x, rhs, np.zeros_like(x))
    finally:
        s.close()


# odd and even unknown counts per axis (vertex / cell coarsening), mixed, 2-D and 3-D, grids too small to
# coarsen, a coarsest level above the one-block limit ((130, 130, 6): z stops the hierarchy at 64 x 64 x 2)
SHAPES = [(129, 129, 1), (130, 130, 1), (100, 64, 1), (67, 41, 1), (33, 33, 33), (34, 34, 34), (35, 50, 21),
          (3, 3, 1), (5, 4, 1), (4, 4, 4), (9, 9, 9), (130, 130, 6), (258, 258, 1), (66, 66, 66), (75, 21, 13)]


@pytest.mark.parametrize("path", ["default", "no_graph", "two_pass_smoother", "one_pass_3d", "no_march",
                                  "coarsest_in_global_memory"])
@pytest.mark.parametrize("shape", SHAPES)
def test_multigrid_matches_cpu_statement_bitwise(orc, shape, path, monkeypatch):
    """default: one-pass two-colour sweeps on 2-D levels, one kernel per colour on 3-D ones, one V-cycle +
    residual check replayed as a CUDA graph; no_graph: the same kernels launched one by one;
    two_pass_smoother: one kernel per colour everywhere; one_pass_3d: the one-pass sweep on 3-D levels too;
    no_march: prolongation with one cell per thread on 3-D levels (default: z-marching threads)."""
    if path == "no_graph":
        monkeypatch.setenv("CFD_B200_NO_MG_GRAPH", "1")
    if path == "two_pass_smoother":
        monkeypatch.setenv("CFD_B200_NO_MG_FUSED", "1")
    if path in ("one_pass_3d", "no_march") and shape[2] == 1:
        pytest.skip("2-D grid: same kernels as the default path")
    if path == "one_pass_3d":
        monkeypatch.setenv("CFD_B200_MG_FUSED3D", "1")
    if path == "no_march":
        monkeypatch.setenv("CFD_B200_NO_MG_MARCH", "1")
    if path == "coarsest_in_global_memory":  # default: the coarsest level is relaxed in shared memory when it fits
        monkeypatch.setenv("CFD_B200_NO_MG_COARSE_SMEM", "1")
    nx, ny, nz = shape
    g, x, rhs = cases.poisson_problem(nx, ny, nz, seed=11)
    pp = mg_params()
    dx, dy, dz = g.spacing
    xo = x.copy()
    rc_o, st_o = orc.poisson(MG, nx, ny, nz, dx, dy, dz, pp, xo, rhs)
    rc_g, st_g = gpu_mg(g, x, rhs, pp)
    assert rc_g == rc_o
    assert (st_g.status, st_g.iterations) == (st_o.status, st_o.iterations)
    assert st_g.initial_residual == st_o.initial_residual
    assert st_g.final_residual == st_o.final_residual
    assert np.array_equal(x, xo)


@pytest.mark.parametrize("shape", [(129, 129, 1), (66, 50, 1), (33, 33, 33), (40, 36, 34)])
def test_multigrid_answer_is_the_stationary_solvers_answer(orc, shape):
    """Same discrete system as the reference's Jacobi / Red-Black SOR (mirrored boundary cells): the two
    solutions differ by a constant (pure Neumann problem) and by the solvers' tolerances only."""
    nx, ny, nz = shape
    g, x, rhs = cases.poisson_problem(nx, ny, nz, seed=3)
    inner = (slice(1, -1) if nz > 1 else slice(None), slice(1, -1), slice(1, -1))
    rhs[inner] -= rhs[inner].mean()  # compatible right-hand side
    pp = mg_params(tolerance=1e-10, absolute_tolerance=1e-14)
    rc, st = gpu_mg(g, x, rhs, pp)
    assert rc == abi.CFD_SUCCESS and st.iterations <= 25
    dx, dy, dz = g.spacing
    xr = np.zeros_like(x)
    pr = abi.poisson_params_default()
    pr.max_iterations, pr.tolerance, pr.absolute_tolerance = 200000, 1e-10, 1e-14
    rc_r, st_r = orc.poisson(abi.POISSON_METHOD_REDBLACK_SOR, nx, ny, nz, dx, dy, dz, pr, xr, rhs)
    assert rc_r == abi.CFD_SUCCESS
    d = (x[inner] - x[inner].mean()) - (xr[inner] - xr[inner].mean())
    scale = np.abs(xr[inner] - xr[inner].mean()).max()
    assert np.abs(d).max() <= 1e-6 * scale
    # boundary cells are left as the other methods leave them: nearest interior value
    xb = x.copy()
    orc.neumann(xb)
    assert np.array_equal(xb, x)


def test_multigrid_needs_far_fewer_passes_than_the_reference_methods(orc):
    """129 x 129 (config C1): V-cycles vs Red-Black SOR iterations to the same residual."""
    g, x, rhs = cases.poisson_problem(129, 129, 1, seed=2)
    rhs[0, 1:-1, 1:-1] -= rhs[0, 1:-1, 1:-1].mean()
    pp = mg_params(tolerance=1e-8)
    rc, st = gpu_mg(g, x, rhs, pp)
    assert rc == abi.CFD_SUCCESS
    assert st.iterations <= 16
    assert st.final_residual < 1e-8 * st.initial_residual


def test_multigrid_cycle_cap_and_early_exit(orc):
    g, x, rhs = cases.poisson_problem(65, 65, 1, seed=4)
    pp = mg_params(max_iterations=2, tolerance=1e-12)
    xo = x.copy()
    dx, dy, dz = g.spacing
    rc_o, st_o = orc.poisson(MG, 65, 65, 1, dx, dy, dz, pp, xo, rhs)
    rc, st = gpu_mg(g, x, rhs, pp)
    assert rc == rc_o == abi.CFD_ERROR_MAX_ITER
    assert st.status == abi.POISSON_MAX_ITER and st.iterations == st_o.iterations == 2
    assert np.array_equal(x, xo)
    # zero right-hand side, zero start: converged before the first cycle
    z = np.zeros_like(x)
    rc, st = gpu_mg(g, z, np.zeros_like(x), mg_params())
    assert rc == abi.CFD_SUCCESS and st.iterations == 0 and not z.any()


@pytest.mark.parametrize("case", ["cavity2d", "box3d"])
def test_projection_step_with_multigrid_is_bitwise_the_cpu_statement(orc, case):
    """The whole projection step with the multigrid solve inside (warm start from the previous pressure).
    Fixed number of cycles, "cap reached => go on" (the reference GPU path's policy,
    solver_projection_gpu.cu:724-733): the right-hand side of a cavity step is not compatible with the
    all-Neumann system, so -- exactly like the reference's Jacobi / Red-Black SOR on these steps -- the
    L-infinity residual stalls at the incompatible part and the relative test is never met."""
    if case == "cavity2d":
        g, f, p = cases.cavity(65, 1)
    else:
        g, f = cases.random_field(34, 30, 26, seed=6)
        p = cases.params(dt=1e-3)
    pc = abi.poisson_config("multigrid", max_iterations=5, on_nonconvergence=abi.CFD_B200_NONCONV_CONTINUE)
    fo = f.copy()
    rc_o, _ = orc.projection(fo, g, p, pc, nsteps=3)
    api.set_poisson(pc)
    try:
        p.max_iter = 3
        rc = api.solve_method_gpu("projection", f, g, p)
    finally:
        api.set_poisson("cg")
    assert rc == rc_o == abi.CFD_SUCCESS
    for name in ("u", "v", "w", "p"):
        assert np.array_equal(getattr(f, name), getattr(fo, name)), name


def test_multigrid_solver_object_reuses_its_hierarchy():
    """Two solves on one solver object (the levels are built on the first one) give the same answer."""
    g, x, rhs = cases.poisson_problem(130, 130, 1, seed=9)
    dx, dy, dz = g.spacing
    s = api.PoissonGpuSolver(MG, g.nx, g.ny, g.nz, dx, dy, dz, mg_params())
    try:
        x1 = x.copy()
        rc1, st1 = s.solve(x1, rhs, None)
        x2 = x.copy()
        rc2, st2 = s.solve(x2, rhs, None)
    finally:
        s.close()
    assert rc1 == rc2 == abi.CFD_SUCCESS and st1.iterations == st2.iterations
    assert np.array_equal(x1, x2)
