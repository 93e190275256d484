"""CG preconditioned by the multigrid cycle for the Krylov operator (csrc/poisson_mgpc.cu; POISSON_METHOD_CG with
params.preconditioner = CFD_B200_PRECOND_MULTIGRID): the GPU solve against its CPU restatement
(oracle/mgpc_host.cpp -- host loops over the same per-cell functions) and against the reference's CG (the oracle)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import cases
from cfd_b200 import abi, api

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DP = C.POINTER(C.c_double)


def ptr(a):
    return a.ctypes.data_as(DP)


@pytest.fixture(scope="module")
def host():
    path = os.path.join(ROOT, "oracle", "libmgpc_host.so")
    if not os.path.exists(path):
        r = subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "libmgpc_host.so"], capture_output=True, text=True)
        assert r.returncode == 0, r.stdout + r.stderr
    L = C.CDLL(path)
    L.mgpc_host_solve.argtypes = [C.c_int] * 3 + [C.c_double] * 3 + [DP, DP, C.c_double, C.c_double, C.c_int, C.c_int, DP]
    L.mgpc_host_solve.restype = C.c_int
    return L


def gpu_solve(g, x, rhs, pp):
    dx, dy, dz = g.spacing
    s = api.PoissonGpuSolver(abi.POISSON_METHOD_CG, g.nx, g.ny, g.nz, dx, dy, dz, pp)
    try:
        return s.solve(x, rhs, np.zeros_like(x))
    finally:
        s.close()


@pytest.mark.parametrize("shape", [(130, 130, 1), (129, 129, 1), (64, 40, 1), (258, 130, 1), (34, 20, 18), (66, 66, 66),
                                   (65, 33, 17)])
def test_gpu_solve_equals_the_cpu_restatement_and_the_reference_cg(orc, host, shape):
    nx, ny, nz = shape
    g, x, rhs = cases.poisson_problem(nx, ny, nz, seed=5)
    dx, dy, dz = g.spacing
    for tol in (1e-6, 1e-12):
        pp = abi.poisson_params_default()
        pp.tolerance, pp.absolute_tolerance, pp.max_iterations = tol, 1e-14, 200
        pp.preconditioner = abi.CFD_B200_PRECOND_MULTIGRID
        xg = x.copy()
        rc, st = gpu_solve(g, xg, rhs, pp)
        xh = x.copy()
        out = np.zeros(3)
        status = host.mgpc_host_solve(nx, ny, nz, dx, dy, dz if nz > 1 else 0.0, ptr(xh), ptr(rhs), tol, 1e-14, 200, 1, ptr(out))
        assert rc == 0 and status == 0 and st.status == abi.POISSON_CONVERGED
        assert abs(st.initial_residual - out[1]) <= 1e-11 * out[1]
        assert abs(st.iterations - int(out[0])) <= 2, (st.iterations, out[0])  # (dot products are summed in another order)
        assert st.final_residual < tol * st.initial_residual
        if st.iterations == int(out[0]):
            l2, li = cases.rel_err(xg, xh)
            assert l2 <= 1e-9 and li <= 1e-9, (l2, li)
        if tol == 1e-12:  # pinned by the reference CG's own answer
            pc = abi.poisson_params_default()
            pc.tolerance, pc.absolute_tolerance, pc.max_iterations = tol, 1e-14, 20000
            xo = x.copy()
            rc_o, st_o = orc.poisson(abi.POISSON_METHOD_CG, nx, ny, nz, dx, dy, dz, pc, xo, rhs, np.zeros_like(x))
            assert rc_o == 0
            l2, li = cases.rel_err(xg, xo)
            assert l2 <= 1e-8 and li <= 1e-8, (l2, li)
            assert st_o.iterations >= 4 * st.iterations


def test_projection_step_with_the_multigrid_preconditioner(orc):
    """The projection step with mgpcg at a tight Poisson tolerance against the oracle's step with the reference's CG."""
    g, f = cases.random_field(66, 34, 1, seed=3, scale=0.05)
    f.rho[...] = 1.0
    p = cases.params(dt=1e-3, amp=(0.1, 0.05))
    pc_g = abi.poisson_config("mgpcg", tolerance=1e-12, absolute_tolerance=1e-14, max_iterations=200)
    pc_o = abi.poisson_config("cg", tolerance=1e-12, absolute_tolerance=1e-14, max_iterations=20000)
    fo = f.copy()
    api.set_poisson(pc_g)
    slv = api.GpuSolver("projection_gpu")
    try:
        for _ in range(2):
            rc, _ = slv.step(f, g, p)
            rco, _ = orc.projection(fo, g, p, pc_o, 1)
            assert rc == rco == 0, api.last_error()
    finally:
        slv.close()
        api.set_poisson("cg")
    for k in ("u", "v", "p"):
        l2, li = cases.rel_err(getattr(f, k), getattr(fo, k))
        assert l2 <= 1e-7 and li <= 1e-7, (k, l2, li)
