"""The "fast" arithmetic build (csrc/common.cuh: FMA contraction on, reciprocal multiplies instead of the
reference's divisions; cfd_b200_set_arithmetic(1) / CFD_B200_ARITH=fast) against the oracle at the north-star
tolerances: rel L2 / L-inf <= 1e-10 PER STEP for Jacobi / Red-Black SOR / RK / Euler / predictor / corrector,
with EQUAL iteration counts on the reference's L-inf stopping rule; <= 1e-8 for the Krylov methods.  The exact
build (default) stays bit-identical -- every other test file asserts that."""
import numpy as np
import pytest

import cases
from cfd_b200 import abi, api

pytestmark = pytest.mark.gpu
FIELDS = ("u", "v", "w", "p", "T")


@pytest.fixture
def fast():
    api.set_arithmetic(1)
    api.load().cfd_b200_release_cached_sessions()
    yield
    api.set_arithmetic(-1)
    api.set_poisson("cg")
    api.load().cfd_b200_release_cached_sessions()


def close(a, b, tol, what):
    worst = 0.0
    for k in FIELDS:
        x, y = getattr(a, k), getattr(b, k)
        if np.max(np.abs(y)) < 1e-12:
            assert np.max(np.abs(x - y)) <= tol, (what, k)
            continue
        l2, li = cases.rel_err(x, y)
        worst = max(worst, l2, li)
        assert l2 <= tol and li <= tol, (what, k, l2, li)
    return worst


def test_fast_build_is_selected_and_differs_in_the_last_bits(fast):
    g, f = cases.random_field(40, 30, 20, seed=2, scale=0.2)
    s = api.Session.for_grid(g)
    assert api.load().cfd_b200_session_arithmetic(s.h) == 1
    s.close()
    p = cases.params(dt=1e-3, amp=(0.1, 0.05))
    a, b = f.copy(), f.copy()
    slv = api.GpuSolver("rk4_gpu")
    assert slv.step(a, g, p)[0] == 0
    slv.close()
    api.set_arithmetic(0)
    slv = api.GpuSolver("rk4_gpu")
    assert slv.step(b, g, p)[0] == 0
    slv.close()
    assert not np.array_equal(a.u, b.u)           # a different instruction stream really ran
    assert np.max(np.abs(a.u - b.u)) <= 1e-12 * np.max(np.abs(b.u))


@pytest.mark.parametrize("method,pkw,tol", [("jacobi", dict(max_iterations=200), 1e-10), ("rbsor", dict(max_iterations=100), 1e-10),
                                            ("cg", {}, 1e-8), ("pcg", {}, 1e-8)])
@pytest.mark.parametrize("case", ["cavity3d_66", "natconv2d_130", "random3d"])
def test_fast_projection_per_step(fast, orc, method, pkw, tol, case):
    if case == "cavity3d_66":
        g, f, p = cases.cavity(66, 34, dt=5e-4)
    elif case == "natconv2d_130":
        g, f, p = cases.natural_convection(130, 1, dt=2e-4)
    else:
        g, f = cases.random_field(68, 36, 21, seed=3, scale=0.05)
        p = cases.params(dt=1e-3, amp=(0.1, 0.05))
    pc = abi.poisson_config(method, on_nonconvergence=abi.CFD_B200_NONCONV_CONTINUE, **pkw)
    api.set_poisson(pc)
    fo = f.copy()
    slv = api.GpuSolver("projection_gpu")
    try:
        for step in range(4):
            if case.startswith("cavity"):
                cases.apply_cavity_bc(f)
                cases.apply_cavity_bc(fo)
            rc, _ = slv.step(f, g, p)
            rco, _ = orc.projection(fo, g, p, pc, 1)
            assert rc == rco == 0
            close(f, fo, tol, "fast %s %s step %d" % (case, method, step))
            if tol == 1e-10:  # per-step bar: restart both sides from the same state, as "per step" means
                for k in FIELDS:
                    getattr(f, k)[...] = getattr(fo, k)
    finally:
        slv.close()


@pytest.mark.parametrize("name,order", [("rk4_gpu", 4), ("rk2_gpu", 2), ("explicit_euler_gpu", 1)])
@pytest.mark.parametrize("shape", [(66, 34, 18), (130, 66, 1)])
def test_fast_explicit_steppers_per_step(fast, orc, name, order, shape):
    g, f = cases.random_field(*shape, seed=9, scale=0.2)
    f.rho[...] = 1.0 + 0.1 * np.random.default_rng(3).random(f.rho.shape)
    p = cases.params(dt=1e-3, amp=(0.1, 0.05))
    fo = f.copy()
    slv = api.GpuSolver(name)
    try:
        for step in range(5):
            assert slv.step(f, g, p)[0] == 0
            assert (orc.explicit_euler(fo, g, p, 1) if order == 1 else orc.rk(fo, g, p, order, 1)) == 0
            close(f, fo, 1e-10, "fast %s step %d" % (name, step))
    finally:
        slv.close()


def test_fast_taylor_green_128_cubed_rk4_twenty_steps_without_resync(fast, orc):
    """C2 at its own size, 20 steps, the fast build running FREE (no per-step resynchronisation): the drift
    against the oracle stays below 1e-10 after all 20 steps."""
    g, f, p = cases.taylor_green(128, nu=0.01, dt=1e-3)
    fo = f.copy()
    p.max_iter = 20
    slv = api.GpuSolver("rk4_gpu")
    try:
        assert slv.solve(f, g, p)[0] == 0
    finally:
        slv.close()
    assert orc.rk(fo, g, p, 4, 20) == 0
    close(f, fo, 1e-10, "fast C2 rk4 x20")


@pytest.mark.parametrize("method", ["jacobi", "rbsor"])
@pytest.mark.parametrize("shape", [(66, 66, 1), (34, 30, 26)])
def test_fast_stationary_solvers_stop_at_the_same_iteration(fast, orc, method, shape):
    """The reference's stopping rule (L-inf residual, tol 1e-6 relative) on a compatible problem: the fast build
    converges at the SAME iteration as the oracle and to the same solution (<= 1e-10)."""
    g, x, rhs = cases.poisson_problem(*shape, seed=6)
    dx, dy, dz = g.spacing
    m = abi.POISSON_METHOD_JACOBI if method == "jacobi" else abi.POISSON_METHOD_REDBLACK_SOR
    pp = abi.poisson_config(method, max_iterations=20000).params
    xo = x.copy()
    xt = np.zeros_like(x)
    rco, sto = orc.poisson(m, g.nx, g.ny, g.nz, dx, dy, dz, pp, xo, rhs, xt)
    slv = api.PoissonGpuSolver(m, g.nx, g.ny, g.nz, dx, dy, dz, pp)
    rc, st = slv.solve(x, rhs, None)
    slv.close()
    assert rc == rco == 0
    assert st.iterations == sto.iterations, (st.iterations, sto.iterations)
    assert max(cases.rel_err(x, xo)) <= 1e-10
