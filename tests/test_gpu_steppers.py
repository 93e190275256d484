"""GPU parity of the time steppers against the oracle, through the C-ABI:
the registry-style vtables (projection_gpu / explicit_euler_gpu / rk2_gpu / rk4_gpu step + solve) and
the stateless link-time symbols solve_{projection,explicit_euler,rk2,rk4}_method_gpu.

Tolerances (north_star): rel L2 / Linf <= 1e-10 per step for projection with Jacobi / RB-SOR and
for RK2 / RK4 (no reductions); <= 1e-8 on the final fields for projection with CG / PCG / BiCGSTAB.
"""
import os

import numpy as np
import pytest

import cases
from cfd_b200 import abi, api

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(__file__), "golden")
FIELDS = ("u", "v", "w", "p", "T")


def assert_close(fg, fo, tol, what=""):
    for k in FIELDS:
        a, b = getattr(fg, k), getattr(fo, k)
        scale = max(np.max(np.abs(b)), 1e-300)
        if np.max(np.abs(b)) < 1e-12:  # reference field is (numerically) zero: absolute check
            assert np.max(np.abs(a - b)) <= tol, (what, k, np.max(np.abs(a - b)))
            continue
        l2, li = cases.rel_err(a, b)
        assert l2 <= tol and li <= tol, (what, k, l2, li, scale)


@pytest.fixture
def poisson_cfg():
    """Restore the process-wide Poisson configuration after each test."""
    yield
    api.set_poisson("cg")


@pytest.mark.parametrize("method,tol", [("jacobi", 1e-10), ("rbsor", 1e-10), ("cg", 1e-8), ("pcg", 1e-8),
                                        ("bicgstab", 1e-8)])
@pytest.mark.parametrize("case", ["cavity2d", "cavity3d", "random3d"])
def test_projection_step_by_step(orc, poisson_cfg, method, tol, case):
    if case == "cavity2d":
        g, f, p = cases.cavity(33, 1, dt=5e-4)
    elif case == "cavity3d":
        g, f, p = cases.cavity(17, 13, dt=5e-4)
    else:
        g, f = cases.random_field(20, 14, 11, seed=3, scale=0.05)
        p = cases.params(dt=1e-3, amp=(0.1, 0.05))
    # early cavity steps do not converge within Jacobi's 2000-iteration cap: use the documented
    # "continue" policy on both sides
    # BiCGSTAB's <= 1e-8 gate is taken at Poisson tolerance 1e-13: at the default 1e-6 the
    # reference's own iterates are chaotic under reduction re-rounding (test_oracle.py,
    # test_bicgstab_is_sensitive_to_reduction_rounding_but_cg_is_not)
    kw = dict(tolerance=1e-13, absolute_tolerance=1e-15) if method == "bicgstab" else {}
    pc = abi.poisson_config(method, on_nonconvergence=abi.CFD_B200_NONCONV_CONTINUE, **kw)
    api.set_poisson(pc)
    fo = f.copy()
    slv = api.GpuSolver("projection_gpu")
    try:
        for step in range(4):
            if case.startswith("cavity"):
                cases.apply_cavity_bc(f)
                cases.apply_cavity_bc(fo)
            rc_g, st = slv.step(f, g, p)
            rc_o, pst = orc.projection(fo, g, p, pc, 1)
            assert rc_g == rc_o == 0
            assert_close(f, fo, tol, "%s %s step %d" % (case, method, step))
            mv, mp, mt = (np.zeros(1) for _ in range(3))
            orc.lib.orc_field_stats(fo.c, abi._ptr(mv), abi._ptr(mp), abi._ptr(mt))
            assert abs(st.max_velocity - mv[0]) <= 1e-8 * max(mv[0], 1e-30)
            assert abs(st.max_pressure - mp[0]) <= 1e-8 * max(mp[0], 1e-30)
            assert abs(st.max_temperature - mt[0]) <= 1e-12 * max(abs(mt[0]), 1.0)
    finally:
        slv.close()


@pytest.mark.parametrize("shape", [(34, 20, 1), (20, 18, 12), (21, 17, 9)])
def test_projection_launders_non_finite_velocities_like_the_reference(orc, poisson_cfg, shape):
    """NaN in INTERIOR velocity cells: the reference's velocity clamp turns it into +100
    (solver_projection.c:180-182) and the step goes on.  Bitwise against the oracle (Jacobi pressure
    solve, capped).  (+-inf inputs are outside the parity claim: the 3-instruction quotient turns an
    infinite numerator into NaN, so a clamp can return +100 where the reference returns -100.)"""
    g, f = cases.random_field(*shape, seed=13, scale=0.05)
    kz = shape[2] // 2
    f.u[kz, 5, 7] = np.nan
    f.v[kz, 9, 4] = -np.nan
    f.u[kz, 8, 11] = np.frombuffer(np.uint64(0xFFF0000000000001).tobytes(), dtype=np.float64)[0]
    if shape[2] > 1:
        f.w[kz, 7, 13] = np.nan
    p = cases.params(dt=1e-3, amp=(0.1, 0.05))
    pc = abi.poisson_config("jacobi", on_nonconvergence=abi.CFD_B200_NONCONV_CONTINUE, max_iterations=30)
    api.set_poisson(pc)
    fo = f.copy()
    slv = api.GpuSolver("projection_gpu")
    try:
        for step in range(2):
            rc_g, _ = slv.step(f, g, p)
            rc_o, _ = orc.projection(fo, g, p, pc, 1)
            assert rc_g == rc_o
            for name in ("u", "v", "w", "p"):
                np.testing.assert_array_equal(getattr(f, name), getattr(fo, name), err_msg="%s step %d" % (name, step))
    finally:
        slv.close()


@pytest.mark.parametrize("order", [2, 4])
@pytest.mark.parametrize("shape", [(16, 16, 1), (12, 10, 9), (3, 3, 3), (33, 20, 1)])
def test_rk_step_by_step(orc, order, shape):
    g, f = cases.random_field(*shape, seed=8, scale=0.2)
    f.rho[...] = 1.0 + 0.1 * np.random.default_rng(1).random(f.rho.shape)
    p = cases.params(dt=1e-3, amp=(0.1, 0.05))
    fo = f.copy()
    slv = api.GpuSolver("rk%d_gpu" % order)
    try:
        for step in range(3):
            rc_g, _ = slv.step(f, g, p)
            rc_o = orc.rk(fo, g, p, order, 1)
            assert rc_g == rc_o == 0
            assert_close(f, fo, 1e-10, "rk%d step %d" % (order, step))
            np.testing.assert_array_equal(f.rho, fo.rho)  # periodic wrap of the density ghosts
    finally:
        slv.close()


@pytest.mark.parametrize("shape", [(16, 16, 1), (12, 10, 9), (3, 3, 3), (33, 20, 1), (130, 20, 11), (66, 40, 1)])
def test_explicit_euler_step_by_step(orc, shape):
    """explicit_euler_gpu against cpu/solver_explicit_euler.c:337-582, bit for bit: plain (unwrapped)
    neighbours, +-1 increment limiter, dt capped at 1e-4, caller boundary velocities preserved,
    periodic images on p / rho / T (and on w when nz == 1), a rho <= 1e-10 cell keeps its state."""
    g, f = cases.random_field(*shape, seed=9, scale=0.2)
    f.rho[...] = 1.0 + 0.1 * np.random.default_rng(1).random(f.rho.shape)
    if shape[0] > 4:
        f.rho[shape[2] // 2, 2, 3] = 0.0
    if shape[2] == 1:
        f.w[...] = 0.05 * np.random.default_rng(2).standard_normal(f.w.shape)
    p = cases.params(dt=1e-3, amp=(0.1, 0.05))
    fo = f.copy()
    slv = api.GpuSolver("explicit_euler_gpu")
    try:
        for step in range(3):
            rc_g, _ = slv.step(f, g, p)
            rc_o = orc.explicit_euler(fo, g, p, 1)
            assert rc_g == rc_o == 0
            for name in FIELDS + ("rho",):
                np.testing.assert_array_equal(getattr(f, name), getattr(fo, name), err_msg="%s step %d" % (name, step))
    finally:
        slv.close()


@pytest.mark.parametrize("poison", [False, True])
@pytest.mark.parametrize("shape", [(20, 18, 12), (130, 20, 11), (34, 20, 1)])
def test_explicit_euler_limiters_engage_like_the_reference(orc, shape, poison):
    """Large amplitudes: derivative clamps, the +-1 increment limiter, the +-10 divergence clamp and
    the +-100 velocity clamp all engage; NaNs are laundered exactly like fmax(-L, fmin(L, v))."""
    g, f = cases.random_field(*shape, seed=22, scale=40.0)
    f.rho[...] = 1.0 + 0.1 * np.random.default_rng(2).random(f.rho.shape)
    if poison:
        kz = shape[2] // 2
        f.u[kz, 5, 7] = np.nan
        f.v[kz, 3, 3] = np.frombuffer(np.uint64(0xFFF0000000000001).tobytes(), dtype=np.float64)[0]
        f.w[kz, 7, 11] = -np.nan
    p = cases.params(dt=5e-2, amp=(0.1, 0.05))  # > DT_CONSERVATIVE_LIMIT: the cap engages too
    fo = f.copy()
    slv = api.GpuSolver("explicit_euler_gpu")
    try:
        for step in range(2):
            rc_g, _ = slv.step(f, g, p)
            rc_o = orc.explicit_euler(fo, g, p, 1)
            assert rc_g == rc_o
            for name in ("u", "v", "w", "p"):
                np.testing.assert_array_equal(getattr(f, name), getattr(fo, name), err_msg="%s step %d" % (name, step))
    finally:
        slv.close()


def test_explicit_euler_multi_step_with_energy(orc):
    """solve(max_iter = N): buoyancy, energy step at the capped dt, thermal BCs, decaying source,
    later iterations reading the periodic pressure images left by the earlier ones."""
    for nz in (1, 6):
        g, f, p = cases.natural_convection(22, nz, dt=5e-5)
        rng = np.random.default_rng(3)
        for a in (f.u, f.v, f.w, f.p):
            a[...] = 0.05 * rng.standard_normal(a.shape)
        p.source_amplitude_u, p.source_amplitude_v = 0.1, 0.05
        p.max_iter = 4
        fo = f.copy()
        assert api.solve_method_gpu("explicit_euler", f, g, p) == 0
        assert orc.explicit_euler(fo, g, p, 4) == 0
        for name in FIELDS + ("rho",):
            np.testing.assert_array_equal(getattr(f, name), getattr(fo, name), err_msg="nz=%d %s" % (nz, name))


@pytest.mark.parametrize("poison", [False, True])
@pytest.mark.parametrize("order", [2, 4])
@pytest.mark.parametrize("shape", [(20, 18, 12), (130, 20, 11), (34, 20, 1), (21, 17, 9)])
def test_rk_clamps_engage_like_the_reference(orc, order, shape, poison):
    """Large-amplitude fields: the derivative clamps (+-100, +-1000), the divergence clamp (+-10) and
    the velocity clamp (+-100) of ns_momentum_rhs_scalar.h all engage.  The kernels take an
    exponent-word fast path when nothing is near a limit; this drives the exact-clamp path."""
    g, f = cases.random_field(*shape, seed=21, scale=40.0)
    f.rho[...] = 1.0 + 0.1 * np.random.default_rng(2).random(f.rho.shape)
    if poison:  # NaN -> +limit in every clamp (fmax(-L, fmin(L, v)) semantics): both signs, odd payloads
        kz = shape[2] // 2
        f.u[kz, 5, 7] = np.nan
        f.u[kz, 9, 4] = -np.nan
        f.v[kz, 3, 3] = np.frombuffer(np.uint64(0xFFF0000000000001).tobytes(), dtype=np.float64)[0]
        f.w[kz, 7, 11] = np.frombuffer(np.uint64(0x7FF0000000000001).tobytes(), dtype=np.float64)[0]
        f.p[kz, 6, 9] = np.nan
    p = cases.params(dt=1e-3, amp=(0.1, 0.05))
    fo = f.copy()
    slv = api.GpuSolver("rk%d_gpu" % order)
    try:
        for step in range(2):
            rc_g, _ = slv.step(f, g, p)
            rc_o = orc.rk(fo, g, p, order, 1)
            assert rc_g == rc_o
            for name in ("u", "v", "w", "p"):
                np.testing.assert_array_equal(getattr(f, name), getattr(fo, name), err_msg="%s step %d" % (name, step))
        assert np.nanmax(np.abs(f.u)) == 100.0  # the velocity clamp really engaged
    finally:
        slv.close()


@pytest.mark.parametrize("nz", [1, 9])
@pytest.mark.parametrize("kind", ["projection", "rk4"])
def test_energy_buoyancy_and_thermal_bcs(orc, poisson_cfg, nz, kind):
    g, f, p = cases.natural_convection(21, nz, dt=1e-3)
    p.thermal_bc.top = abi.BC_TYPE_PERIODIC  # mixed face types: corner precedence matters
    api.set_poisson("cg")
    fo = f.copy()
    slv = api.GpuSolver("projection_gpu" if kind == "projection" else "rk4_gpu")
    try:
        for step in range(3):
            rc_g, _ = slv.step(f, g, p)
            if kind == "projection":
                rc_o, _ = orc.projection(fo, g, p, abi.poisson_config("cg"), 1)
            else:
                rc_o = orc.rk(fo, g, p, 4, 1)
            assert rc_g == rc_o == 0
            assert_close(f, fo, 1e-8 if kind == "projection" else 1e-10, "%s nz=%d step %d" % (kind, nz, step))
    finally:
        slv.close()


@pytest.mark.parametrize("kind,order", [("projection", 0), ("rk2", 2), ("rk4", 4)])
def test_stateless_entry_points_multi_step(orc, poisson_cfg, kind, order):
    """solve_*_method_gpu(max_iter = N): N device-resident steps, source term decays with the
    iteration index (solver_explicit_euler.c:326-331)."""
    g, f = cases.random_field(18, 15, 7, seed=12, scale=0.05)
    p = cases.params(dt=1e-3, amp=(0.1, 0.05))
    p.max_iter = 5
    api.set_poisson("cg")
    fo = f.copy()
    rc_g = api.solve_method_gpu(kind, f, g, p)
    if kind == "projection":
        rc_o, _ = orc.projection(fo, g, p, abi.poisson_config("cg"), 5)
    else:
        rc_o = orc.rk(fo, g, p, order, 5)
    assert rc_g == rc_o == 0
    assert_close(f, fo, 1e-8 if kind == "projection" else 1e-10, kind)


@pytest.mark.parametrize("tag,case,name", [
    ("golden16_projection", lambda: cases.golden16(), "projection_gpu"),
    ("golden16_rk2", lambda: cases.golden16(), "rk2_gpu"),
    ("golden16_rk4", lambda: cases.golden16(), "rk4_gpu"),
    ("tg3d_rk4", lambda: cases.taylor_green(12, 10, dt=1e-3), "rk4_gpu"),
    ("golden16_explicit_euler", lambda: cases.golden16(), "explicit_euler_gpu"),
    ("tg3d_explicit_euler", lambda: cases.taylor_green(12, 10, dt=1e-3), "explicit_euler_gpu"),
    ("natconv2d_explicit_euler", lambda: cases.natural_convection(17, 1, dt=5e-5), "explicit_euler_gpu"),
    ("tg3d_projection", lambda: cases.taylor_green(12, 10, dt=1e-3), "projection_gpu"),
    ("natconv2d_projection", lambda: cases.natural_convection(17, 1, dt=2e-3), "projection_gpu"),
])
def test_against_reference_golden_vectors(poisson_cfg, tag, case, name):
    z = np.load(os.path.join(GOLD, tag + ".npz"))
    g, f, p = case()
    api.set_poisson("cg")
    slv = api.GpuSolver(name)
    try:
        for _ in range(int(z["nsteps"])):
            rc, _ = slv.step(f, g, p)
            assert rc == 0
    finally:
        slv.close()
    tol = 1e-8 if "projection" in tag else 1e-10
    for k in FIELDS:
        ref = z["out_" + k]
        if np.max(np.abs(ref)) < 1e-12:
            assert np.max(np.abs(getattr(f, k) - ref)) <= tol
        else:
            assert max(cases.rel_err(getattr(f, k), ref)) <= tol, (tag, k)


def test_reference_golden_l2_norms_on_gpu(poisson_cfg):
    """tests/solvers/navier_stokes/cpu/test_ns_solver_3d.c:343-367 (tolerance 1e-12 there; the GPU
    path reorders CG reductions, so 1e-9 here for projection and 1e-12 for RK)."""
    gold = {"projection_gpu": (6.84647639323831686e-02, 3.42315494726977212e-02, 1.00000039251590289e+00, 1e-9),
            "rk2_gpu": (6.88584742267375205e-02, 3.49775875753182836e-02, 1.00000000000000044e+00, 1e-12),
            "rk4_gpu": (6.88584742267390471e-02, 3.49775875752817156e-02, 1.00000000000000000e+00, 1e-12),
            "explicit_euler_gpu": (6.84647305901105868e-02, 3.42314945425199885e-02, 1.00000000000000000e+00, 1e-12)}
    api.set_poisson("cg")
    for name, (gu, gv, gp, tol) in gold.items():
        g, f, p = cases.golden16()
        slv = api.GpuSolver(name)
        rc, _ = slv.step(f, g, p)
        slv.close()
        assert rc == 0
        l2 = lambda a: np.sqrt(np.sum(a.ravel() ** 2) / a.size)
        assert abs(l2(f.u) - gu) < tol and abs(l2(f.v) - gv) < tol and abs(l2(f.p) - gp) < tol
        assert np.max(np.abs(f.w)) == 0.0


def test_error_behaviour(poisson_cfg):
    g, f, p = cases.cavity(17)
    p.max_iter = 1
    # gpu_should_use() false -> CFD_ERROR (solver_projection_gpu.cu:623-624)
    cfg = api.load().gpu_config_default()
    assert api.solve_method_gpu("projection", f, g, p, cfg) == abi.CFD_ERROR  # min_grid_size 10000
    # host callbacks cannot run on the device
    p2 = cases.params()
    p2.max_iter = 1
    p2.source_func = 1
    assert api.solve_method_gpu("projection", f, g, p2) == abi.CFD_ERROR_UNSUPPORTED
    p3 = cases.params()
    p3.max_iter = 1
    p3.alpha, p3.heat_source_func = 0.1, 1
    assert api.solve_method_gpu("rk4", f, g, p3) == abi.CFD_ERROR_UNSUPPORTED
    # unsupported thermal BC type
    p4 = cases.params(alpha=0.1)
    p4.max_iter = 1
    p4.thermal_bc.left = abi.BC_TYPE_NOSLIP
    assert api.solve_method_gpu("projection", f, g, p4) == abi.CFD_ERROR_INVALID
    # non-uniform spacing
    g2 = abi.HostGrid(17, 17, 1)
    g2.dx[3] *= 1.5
    assert api.solve_method_gpu("projection", f, g2, p) == abi.CFD_ERROR_UNSUPPORTED
    # Poisson hits its cap -> CFD_ERROR_MAX_ITER and the host field is left untouched
    api.set_poisson("jacobi", max_iterations=3)
    cases.apply_cavity_bc(f)
    before = f.copy()
    assert api.solve_method_gpu("projection", f, g, p) == abi.CFD_ERROR_MAX_ITER
    for k in FIELDS:
        np.testing.assert_array_equal(getattr(f, k), getattr(before, k))
    # NaN handling is the reference's, oddities included: a NaN in an INTERIOR velocity cell is
    # laundered to +100 by the fmax(-100, fmin(100, .)) clamp (solver_projection.c:180-182) and the
    # step reports success; a NaN on a boundary cell or in p survives to the scan (:281-289).
    api.set_poisson("cg", on_nonconvergence=abi.CFD_B200_NONCONV_CONTINUE)
    f1 = before.copy()
    f1.u[0, 5, 5] = np.nan
    assert api.solve_method_gpu("projection", f1, g, p) == abi.CFD_SUCCESS
    assert np.all(np.isfinite(f1.u))
    f2 = before.copy()
    f2.u[0, 0, 5] = np.nan
    assert api.solve_method_gpu("projection", f2, g, p) == abi.CFD_ERROR_DIVERGED
    f3 = before.copy()
    f3.p[0, 5, 5] = np.nan
    assert api.solve_method_gpu("projection", f3, g, p) == abi.CFD_ERROR_DIVERGED
    # explicit Euler: same rules (interior NaN velocity laundered by the clamps, boundary NaN reported)
    f4 = before.copy()
    f4.u[0, 5, 5] = np.nan
    assert api.solve_method_gpu("explicit_euler", f4, g, p) == abi.CFD_SUCCESS
    f5 = before.copy()
    f5.v[0, 0, 5] = np.nan
    assert api.solve_method_gpu("explicit_euler", f5, g, p) == abi.CFD_ERROR_DIVERGED
    p5 = cases.params()
    p5.max_iter, p5.source_func = 1, 1
    assert api.solve_method_gpu("explicit_euler", before.copy(), g, p5) == abi.CFD_ERROR_UNSUPPORTED


def test_velocity_boundaries_are_preserved(poisson_cfg):
    """Caller-set boundary velocities survive the step (boundary_copy_utils.h:93-148)."""
    g, f, p = cases.cavity(21, 9)
    rng = np.random.default_rng(0)
    for a in (f.u, f.v, f.w):
        a[...] = 0.01 * rng.standard_normal(a.shape)
    before = f.copy()
    api.set_poisson("cg")
    slv = api.GpuSolver("projection_gpu")
    rc, _ = slv.step(f, g, p)
    slv.close()
    assert rc == 0
    for name in ("u", "v", "w"):
        a, b = getattr(f, name), getattr(before, name)
        for sl in ((0,), (-1,), (slice(None), 0), (slice(None), -1), (slice(None), slice(None), 0),
                   (slice(None), slice(None), -1)):
            np.testing.assert_array_equal(a[sl], b[sl])


def test_divergence_drops_after_projection(poisson_cfg):
    """Property at a size the oracle is not run at: the corrected field is (nearly)
    divergence-free in the discrete sense used by the RHS (solver_projection.c:195-214)."""
    g, f, p = cases.taylor_green(128, 1, dt=1e-3)
    api.set_poisson("cg", tolerance=1e-10)
    dx, dy, _ = g.spacing

    def div(F):
        return ((F.u[0, 1:-1, 2:] - F.u[0, 1:-1, :-2]) / (2 * dx) + (F.v[0, 2:, 1:-1] - F.v[0, :-2, 1:-1]) / (2 * dy))

    slv = api.GpuSolver("projection_gpu")
    rc, _ = slv.step(f, g, p)
    slv.close()
    assert rc == 0 and np.all(np.isfinite(f.u))
    assert np.max(np.abs(div(f)[2:-2, 2:-2])) < 0.25  # reference gate (test_solver_projection_gpu.c:335)


def test_taylor_green_rk4_decay(orc):
    """Taylor-Green 3-D decay (tests/validation/taylor_green_3d_reference.h): velocity decays like
    exp(-3 nu t) within 15 %, and GPU == oracle to 1e-10."""
    g, f, p = cases.taylor_green(16, 16, nu=0.01, dt=1e-3)
    fo = f.copy()
    u0 = np.max(np.abs(f.u))
    slv = api.GpuSolver("rk4_gpu")
    p.max_iter = 20
    rc, _ = slv.solve(f, g, p)
    slv.close()
    assert rc == 0 and orc.rk(fo, g, p, 4, 20) == 0
    assert_close(f, fo, 1e-10, "tg3d")
    decay = np.max(np.abs(f.u)) / u0
    assert abs(decay - np.exp(-3 * 0.01 * 20e-3)) < 0.15


def test_legacy_context_api(orc, poisson_cfg):
    """gpu_solver_create / upload / step / download (gpu_device.h:145-194)."""
    import ctypes as C
    L = api.load()
    g, f, p = cases.cavity(17, 1)
    cases.apply_cavity_bc(f)
    fo = f.copy()
    api.set_poisson("cg")
    ctx = L.gpu_solver_create(17, 17, 1, None)
    assert ctx
    assert L.gpu_solver_upload(ctx, C.byref(f.c)) == 0
    st = abi.GpuSolverStats()
    for _ in range(2):
        assert L.gpu_solver_step(ctx, C.byref(g.c), C.byref(p), C.byref(st)) == 0
    assert L.gpu_solver_download(ctx, C.byref(f.c)) == 0
    L.gpu_solver_destroy(ctx)
    rc, _ = orc.projection(fo, g, p, abi.poisson_config("cg"), 1)
    rc, _ = orc.projection(fo, g, p, abi.poisson_config("cg"), 1)
    assert_close(f, fo, 1e-8, "legacy")
    assert st.kernels_launched > 0 and st.poisson_iterations > 0


@pytest.mark.parametrize("name,kind", [("projection_gpu", "projection"), ("rk4_gpu", "rk4"), ("rk2_gpu", "rk2"),
                                       ("explicit_euler_gpu", "euler")])
def test_dropin_through_the_reference_registry(orc, poisson_cfg, name, kind):
    """Link-time drop-in: the UNMODIFIED reference core/api, compiled with
    CFD_HAS_CUDA=1 and linked against libcfd_b200.so instead of libcfd_cuda.a
    (oracle/_ref/libcfd_ref_dropin.so), creates "<name>" from its own registry and steps it; the
    reference's wrappers (solver_registry.c:1121-1245) call our solve_*_method_gpu symbols."""
    import oracle
    try:
        ref = oracle.Ref("dropin")
    except (FileNotFoundError, OSError) as e:
        pytest.skip(str(e))
    api.set_poisson("cg")
    g, f = cases.random_field(18, 14, 9, seed=4, scale=0.05)
    p = cases.params(dt=1e-3, amp=(0.1, 0.05))
    fo = f.copy()
    n0 = api.load().cfd_b200_kernel_launch_count()
    rc, st = ref.run_solver(name, f, g, p, nsteps=2)
    assert rc == 0
    assert api.load().cfd_b200_kernel_launch_count() > n0  # our kernels did the work
    for _ in range(2):
        if kind == "projection":
            rco, _ = orc.projection(fo, g, p, abi.poisson_config("cg"), 1)
        elif kind == "euler":
            rco = orc.explicit_euler(fo, g, p, 1)
        else:
            rco = orc.rk(fo, g, p, int(kind[2]), 1)
        assert rco == 0
    assert_close(f, fo, 1e-8 if kind == "projection" else 1e-10, "dropin " + name)


# Ghia, Ghia & Shin (1982), J. Comput. Phys. 48, 387-411, Re = 100: u on the vertical and v on the
# horizontal centreline (the tables the reference validates against, tests/validation/cavity_reference_data.h)
GHIA_Y = [0.0000, 0.0547, 0.0625, 0.0703, 0.1016, 0.1719, 0.2813, 0.4531, 0.5000, 0.6172, 0.7344, 0.8516, 0.9531,
          0.9609, 0.9688, 0.9766, 1.0000]
GHIA_U = [0.00000, -0.03717, -0.04192, -0.04775, -0.06434, -0.10150, -0.15662, -0.21090, -0.20581, -0.13641, 0.00332,
          0.23151, 0.68717, 0.73722, 0.78871, 0.84123, 1.00000]
GHIA_X = [0.0000, 0.0625, 0.0703, 0.0781, 0.0938, 0.1563, 0.2266, 0.2344, 0.5000, 0.8047, 0.8594, 0.9063, 0.9453,
          0.9531, 0.9609, 0.9688, 1.0000]
GHIA_V = [0.00000, 0.09233, 0.10091, 0.10890, 0.12317, 0.16077, 0.17507, 0.17527, 0.05454, -0.24533, -0.22445,
          -0.16914, -0.10313, -0.08864, -0.07391, -0.05906, 0.00000]


def _ghia_rms(f, n):
    """RMS against the tables by linear interpolation of the computed centrelines
    (tests/validation/cavity_validation_utils.h:36-65)."""
    c = np.linspace(0.0, 1.0, n)
    u_mid, v_mid = f.u[0, :, n // 2], f.v[0, n // 2, :]
    ru = np.sqrt(np.mean((np.interp(GHIA_Y, c, u_mid) - np.array(GHIA_U)) ** 2))
    rv = np.sqrt(np.mean((np.interp(GHIA_X, c, v_mid) - np.array(GHIA_V)) ** 2))
    return ru, rv


def test_ghia_lid_driven_cavity_re100(orc, poisson_cfg):
    """The reference's validation run (tests/validation/test_ghia_projection_gpu.c: 33 x 33, 5000 steps,
    dt = 5e-4, host-side cavity BCs before every step, gate RMS_u < 0.10; the reference documents
    RMS_u 0.0382 / RMS_v 0.0440 for its CPU projection, docs/validation/cavity-backends-validation.md).
    Run through the persistent projection_gpu solver; the first 300 steps are also taken on the oracle
    and compared (<= 1e-8, CG reorders reductions)."""
    n = 33
    g, f, p = cases.cavity(n, 1, dt=5e-4)
    api.set_poisson(abi.poisson_config("cg", on_nonconvergence=abi.CFD_B200_NONCONV_CONTINUE))
    fo = f.copy()
    pc = abi.poisson_config("cg", on_nonconvergence=abi.CFD_B200_NONCONV_CONTINUE)
    slv = api.GpuSolver("projection_gpu")
    try:
        for step in range(5000):
            cases.apply_cavity_bc(f)
            rc, _ = slv.step(f, g, p)
            assert rc == 0, step
            if step < 300:
                cases.apply_cavity_bc(fo)
                assert orc.projection(fo, g, p, pc, 1)[0] == 0
                if step == 299:
                    assert_close(f, fo, 1e-8, "cavity after 300 steps")
    finally:
        slv.close()
    ru, rv = _ghia_rms(f, n)
    assert ru < 0.10 and rv < 0.10, (ru, rv)                  # the reference's gate
    assert abs(ru - 0.0382) < 0.004 and abs(rv - 0.0440) < 0.004, (ru, rv)  # its documented result


def test_c1_ghia_129_jacobi_per_step_bitwise(orc):
    """BASELINE config C1: 129 x 129 lid-driven cavity, Re = 100, dt = 2e-4, projection with the Jacobi
    pressure solve under the reference rule (1e-6 relative, cap 2000, check every 10 sweeps; early steps
    hit the cap and go on, as the reference GPU path does).  Odd nx: no TMA; the pressure solve is the
    whole-solve cluster kernel.  Every step equals the oracle bit for bit, sweep counts included."""
    n = 129
    g, f, p = cases.cavity(n, 1, dt=2e-4)
    pc = abi.poisson_config("jacobi", on_nonconvergence=abi.CFD_B200_NONCONV_CONTINUE)
    api.set_poisson(pc)
    fo = f.copy()
    slv = api.GpuSolver("projection_gpu")
    try:
        for step in range(40):
            cases.apply_cavity_bc(f)
            cases.apply_cavity_bc(fo)
            rc, _ = slv.step(f, g, p)
            rco, sto = orc.projection(fo, g, p, pc, 1)
            assert rc == rco == 0, step
            for k in ("u", "v", "p"):
                np.testing.assert_array_equal(getattr(f, k), getattr(fo, k), err_msg="%s step %d" % (k, step))
    finally:
        slv.close()
        api.set_poisson("cg")
