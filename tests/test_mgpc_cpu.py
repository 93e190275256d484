"""CPU restatement of the multigrid-preconditioned CG (oracle/mgpc_host.cpp: host loops over the per-cell functions of
cfd_b200/csrc/mgpc_cell.h, the same functions the CUDA kernels loop over) against the numpy prototype of the method
and against the reference's own CG (the oracle)."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest

import cases
from cfd_b200 import abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
DP = C.POINTER(C.c_double)


def ptr(a):
    return a.ctypes.data_as(DP)


@pytest.fixture(scope="module")
def host():
    r = subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "libmgpc_host.so"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    L = C.CDLL(os.path.join(ROOT, "oracle", "libmgpc_host.so"))
    L.mgpc_host_apply.argtypes = [C.c_int] * 3 + [C.c_double] * 3 + [DP, DP]
    L.mgpc_host_apply.restype = C.c_int
    L.mgpc_host_solve.argtypes = [C.c_int] * 3 + [C.c_double] * 3 + [DP, DP, C.c_double, C.c_double, C.c_int, C.c_int, DP]
    L.mgpc_host_solve.restype = C.c_int
    return L


@pytest.mark.parametrize("ns", [(9, 7), (33, 33), (34, 34), (40, 23), (17, 17, 17), (18, 18, 18), (20, 14, 12), (33, 34, 35)])
def test_cycle_equals_the_numpy_prototype(host, ns):
    """One V(1,1) cycle z = M^-1 r: the C++ per-cell functions (operator rows with the wall-distance coefficient,
    position-based restriction / prolongation, level plan) against scipy sparse matrices built from node positions."""
    import mgpcg_dirichlet_prototype as mp
    d3 = len(ns) == 3
    nx, ny = ns[0], ns[1]
    nz = ns[2] if d3 else 1
    h = [1.0 / (n - 1) for n in ns]
    rng = np.random.default_rng(1)
    r = np.zeros((nz, ny, nx))
    inner = (slice(1, -1) if d3 else slice(None), slice(1, -1), slice(1, -1))
    rin = rng.standard_normal(r[inner].shape)
    r[inner] = rin
    z = np.full((nz, ny, nx), 7.0)
    nl = host.mgpc_host_apply(nx, ny, nz, h[0], h[1], h[2] if d3 else 0.0, ptr(r), ptr(z))
    levels = mp.build(ns, "rediscr", coarse_sweeps=40, coupling=0.5)
    assert nl == len(levels)
    vec = np.transpose(rin, (2, 1, 0)).ravel() if d3 else rin[0].T.ravel()  # the prototype orders (x, y[, z])
    zp = mp.vcycle(levels, 0, vec, 1, 0.85 if d3 else 0.8)
    zp = np.transpose(zp.reshape(nx - 2, ny - 2, nz - 2), (2, 1, 0)) if d3 else zp.reshape(nx - 2, ny - 2).T[None]
    # (the prototype relaxes its coarsest level with omega 0.85 in 2-D too: 1e-5 of the cycle at most)
    assert np.max(np.abs(z[inner] - zp)) <= 2e-5 * np.max(np.abs(zp))
    zb = z.copy()
    zb[inner] = 0.0
    assert not zb.any()  # boundary cells of the result are zero


@pytest.mark.parametrize("shape", [(33, 33, 1), (64, 40, 1), (129, 129, 1), (130, 130, 1), (258, 66, 1), (34, 20, 18),
                                   (33, 33, 33), (34, 34, 34), (66, 18, 21)])
def test_solve_against_the_reference_cg(host, orc, shape):
    """Same system, same stopping rule: the initial residual is the reference's, a tight solve lands on the reference
    CG's answer (<= 1e-8), and the iteration count is a handful where the reference needs tens to hundreds."""
    nx, ny, nz = shape
    g, x, rhs = cases.poisson_problem(nx, ny, nz, seed=5)
    dx, dy, dz = g.spacing
    for tol, gate in ((1e-6, 5e-3), (1e-11, 1e-8)):
        pp = abi.poisson_params_default()
        pp.tolerance, pp.absolute_tolerance, pp.max_iterations = tol, 1e-14, 20000
        xo = x.copy()
        rc_o, st_o = orc.poisson(abi.POISSON_METHOD_CG, nx, ny, nz, dx, dy, dz, pp, xo, rhs, np.zeros_like(x))
        assert rc_o == 0
        xm = x.copy()
        out = np.zeros(3)
        status = host.mgpc_host_solve(nx, ny, nz, dx, dy, dz if nz > 1 else 0.0, ptr(xm), ptr(rhs), tol, 1e-14, 200, 1, ptr(out))
        assert status == 0
        assert abs(out[1] - st_o.initial_residual) <= 1e-12 * st_o.initial_residual
        assert out[2] < tol * out[1]
        l2, li = cases.rel_err(xm, xo)
        assert l2 <= gate and li <= gate, (tol, l2, li)
        if tol == 1e-6:  # isotropic or not (semi-coarsening by coupling strength)
            assert out[0] <= 9, out[0]
            assert st_o.iterations >= 4 * out[0]


def test_level_plan():
    """Odd node counts move the wall away from the last coarse node; axes stop at three nodes."""
    r = subprocess.run(["g++", "-std=c++17", "-x", "c++", "-", "-o", os.path.join(ROOT, "oracle", "_mgpc_plan_probe")],
                       input='#include "%s/cfd_b200/csrc/mgpc_cell.h"\n#include <stdio.h>\nint main(){cfdb::MgpcLv lv[32];'
                             'int nl=cfdb::mgpc_plan(131,66,1,0.5,0.25,0.0,lv);for(int l=0;l<nl;l++)'
                             'printf("%%d %%d %%d %%g %%g %%g %%g %%d %%d\\n",lv[l].n[0],lv[l].n[1],lv[l].n[2],lv[l].H[0],lv[l].d[0],'
                             'lv[l].H[1],lv[l].d[1],lv[l].coarsen[0],lv[l].coarsen[1]);}' % ROOT,
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    out = subprocess.run([os.path.join(ROOT, "oracle", "_mgpc_plan_probe")], capture_output=True, text=True).stdout
    os.remove(os.path.join(ROOT, "oracle", "_mgpc_plan_probe"))
    rows = [[float(v) for v in line.split()] for line in out.strip().splitlines()]
    # y (h = 0.25) is coupled 4 x more strongly than x (h = 0.5): it is halved alone until the spacings meet
    # x: 129 nodes: 129, 129 -> 64 (odd: d = h + h = 1.0, H = 1.0) -> 32 -> 16 -> 8 -> 4 -> 2
    assert [int(r[0]) for r in rows] == [129, 129, 64, 32, 16, 8, 4, 2]
    assert rows[2][3] == 1.0 and rows[2][4] == 1.0 and rows[3][3] == 2.0 and rows[3][4] == 1.0
    # y: 64 nodes: 64 -> 32 -> 16 -> 8 -> 4 -> 2 -> 2 -> 2 (not coarsened below 3)
    assert [int(r[1]) for r in rows] == [64, 32, 16, 8, 4, 2, 2, 2]
    assert all(r[6] == 0.25 for r in rows)  # even counts keep the wall distance of the fine grid
    assert [int(r[2]) for r in rows] == [1] * 8
    assert (rows[0][7], rows[0][8]) == (0, 1) and (rows[1][7], rows[1][8]) == (1, 1)
    assert rows[-1][7] == 0 and rows[-1][8] == 0


@pytest.mark.parametrize("ns", [(34, 23), (33, 33), (18, 15, 13), (17, 17, 17)])
def test_cycle_is_a_symmetric_positive_operator(host, ns):
    """PCG needs M^-1 symmetric positive definite: <M^-1 a, b> = <a, M^-1 b> and <M^-1 a, a> > 0 -- which the
    damped-Jacobi pre / post smoothing, R = P^T and the symmetric wall rows are there for."""
    d3 = len(ns) == 3
    nx, ny = ns[0], ns[1]
    nz = ns[2] if d3 else 1
    h = [0.7 / (nx - 1), 1.0 / (ny - 1), (1.3 / (nz - 1)) if d3 else 0.0]
    rng = np.random.default_rng(4)
    inner = (slice(1, -1) if d3 else slice(None), slice(1, -1), slice(1, -1))

    def rand():
        a = np.zeros((nz, ny, nx))
        a[inner] = rng.standard_normal(a[inner].shape)
        return a

    def apply(a):
        z = np.zeros_like(a)
        host.mgpc_host_apply(nx, ny, nz, h[0], h[1], h[2], ptr(a), ptr(z))
        return z

    for _ in range(3):
        a, b = rand(), rand()
        za, zb = apply(a), apply(b)
        lhs, rhs = float(np.sum(za * b)), float(np.sum(a * zb))
        assert abs(lhs - rhs) <= 1e-11 * max(abs(lhs), abs(rhs), 1e-30), (lhs, rhs)
        assert float(np.sum(za * a)) > 0.0
