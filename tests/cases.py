"""Deterministic set-ups shared by the tests (restated from the reference's own tests)."""
import numpy as np

from cfd_b200 import abi


def params(dt=1e-4, mu=0.01, amp=(0.0, 0.0), **kw):
    p = abi.ns_params_default()
    p.dt, p.mu = dt, mu
    p.source_amplitude_u, p.source_amplitude_v = amp
    for k, v in kw.items():
        if k == "gravity":
            p.gravity[0], p.gravity[1], p.gravity[2] = v
        else:
            setattr(p, k, v)
    return p


def golden16():
    """tests/solvers/navier_stokes/cpu/test_ns_solver_3d.c:263-300 (16x16, nz=1)."""
    g = abi.HostGrid(16, 16, 1, 0.0, 1.0, 0.0, 1.0)
    f = abi.HostField(16, 16, 1)
    X, Y = np.meshgrid(g.x, g.y)
    f.u[0] = 0.1 * np.sin(np.pi * Y)
    f.v[0] = 0.05 * np.sin(2.0 * np.pi * X)
    f.p[...] = 1.0
    f.T[...] = 300.0
    return g, f, params(dt=1e-4)


def cavity(n=33, nz=1, dt=5e-4, lid=1.0):
    """Lid-driven cavity (tests/validation/lid_driven_cavity_common.h:98-148): quiescent,
    u = lid on j = ny-1, no-slip elsewhere.  3-D: lid on the whole j = ny-1 face."""
    g = abi.HostGrid(n, n, nz, 0.0, 1.0, 0.0, 1.0, 0.0, 1.0 if nz > 1 else 0.0)
    f = abi.HostField(n, n, nz)
    f.u[:, -1, :] = lid
    return g, f, params(dt=dt, mu=0.01)


def apply_cavity_bc(f, lid=1.0):
    """Host-side BCs the reference drivers re-apply before every step
    (lid_driven_cavity_common.h:120-148): no-slip walls, moving lid, Neumann pressure."""
    for a in (f.u, f.v, f.w):
        a[:, 0, :] = 0.0
        a[:, -1, :] = 0.0
        a[:, :, 0] = 0.0
        a[:, :, -1] = 0.0
        if f.nz > 1:
            a[0] = 0.0
            a[-1] = 0.0
    f.u[:, -1, :] = lid
    p = f.p
    if f.nz > 1:
        p[0] = p[1]
        p[-1] = p[-2]
    p[:, :, 0] = p[:, :, 1]
    p[:, :, -1] = p[:, :, -2]
    p[:, 0, :] = p[:, 1, :]
    p[:, -1, :] = p[:, -2, :]


def taylor_green(n=16, nz=None, nu=0.01, dt=1e-3):
    """Taylor-Green vortex on [0,2pi]^d (tests/validation/taylor_green_3d_reference.h:64-70)."""
    nz = n if nz is None else nz
    L = 2.0 * np.pi
    g = abi.HostGrid(n, n, nz, 0.0, L, 0.0, L, 0.0, L if nz > 1 else 0.0)
    f = abi.HostField(n, n, nz)
    if nz > 1:
        Z, Y, X = np.meshgrid(g.z, g.y, g.x, indexing="ij")
        f.u[...] = np.cos(X) * np.sin(Y) * np.cos(Z)
        f.v[...] = -np.sin(X) * np.cos(Y) * np.cos(Z)
    else:
        Y, X = np.meshgrid(g.y, g.x, indexing="ij")
        f.u[0] = np.cos(X) * np.sin(Y)
        f.v[0] = -np.sin(X) * np.cos(Y)
    f.T[...] = 300.0
    return g, f, params(dt=dt, mu=nu)


def natural_convection(n=41, nz=1, dt=2e-3):
    """de Vahl Davis Ra=1e3 (tests/validation/test_natural_convection.c:49-58, 139-200)."""
    g = abi.HostGrid(n, n, nz, 0.0, 1.0, 0.0, 1.0, 0.0, 1.0 if nz > 1 else 0.0)
    f = abi.HostField(n, n, nz)
    f.T[...] = 310.0 - 20.0 * g.x[None, None, :]
    Pr, Ra, beta, gmag, dT = 0.71, 1e3, 0.003333, 9.81, 20.0
    nu = np.sqrt(gmag * beta * dT * Pr / Ra)
    alpha = nu / Pr
    p = params(dt=dt, mu=nu, alpha=alpha, beta=beta, T_ref=300.0, gravity=(0.0, -gmag, 0.0))
    tb = p.thermal_bc
    tb.left, tb.right = abi.BC_TYPE_DIRICHLET, abi.BC_TYPE_DIRICHLET
    tb.bottom, tb.top = abi.BC_TYPE_NEUMANN, abi.BC_TYPE_NEUMANN
    tb.front, tb.back = abi.BC_TYPE_NEUMANN, abi.BC_TYPE_NEUMANN
    tb.dirichlet_values.left, tb.dirichlet_values.right = 310.0, 290.0
    return g, f, p


def random_field(nx, ny, nz, seed=0, scale=0.1):
    rng = np.random.default_rng(seed)
    g = abi.HostGrid(nx, ny, nz, 0.0, 1.0, 0.0, 1.3, 0.0, 0.7 if nz > 1 else 0.0)
    f = abi.HostField(nx, ny, nz)
    for name in ("u", "v", "w", "p"):
        getattr(f, name)[...] = scale * rng.standard_normal((nz, ny, nx))
    if nz == 1:
        f.w[...] = 0.0
    f.T[...] = 300.0 + rng.standard_normal((nz, ny, nx))
    return g, f


def poisson_problem(nx, ny, nz, seed=0):
    """Zero-mean random rhs (compatible with the all-Neumann problem) and a random start."""
    rng = np.random.default_rng(seed)
    g = abi.HostGrid(nx, ny, nz, 0.0, 1.0, 0.0, 1.0, 0.0, 1.0 if nz > 1 else 0.0)
    rhs = abi.aligned_zeros(nx * ny * nz).reshape(nz, ny, nx)
    sl = (slice(1, -1) if nz > 1 else slice(None), slice(1, -1), slice(1, -1))
    r = rng.standard_normal(rhs[sl].shape)
    rhs[sl] = r - r.mean()
    x = abi.aligned_zeros(nx * ny * nz).reshape(nz, ny, nx)
    x[...] = 0.01 * rng.standard_normal(x.shape)
    return g, x, rhs


def rel_err(a, b):
    """(rel-L2, rel-Linf) of a against reference b (absolute when b is identically zero)."""
    a, b = np.asarray(a, float).ravel(), np.asarray(b, float).ravel()
    d = a - b
    nb, mb = np.linalg.norm(b), np.max(np.abs(b))
    l2 = np.linalg.norm(d) / nb if nb > 0 else np.linalg.norm(d)
    li = np.max(np.abs(d)) / mb if mb > 0 else np.max(np.abs(d))
    return l2, li
