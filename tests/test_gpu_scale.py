"""Cross-checks at sizes the CPU oracle does not finish in seconds (many tiles, many z-chunks, many
waves of blocks): the TMA plane-pipeline kernels against the direct 2.5-D kernels of the same library
(CFD_B200_NO_TMA=1), and the one-pass Red-Black SOR against its two half-sweeps.  The two code paths
share the arithmetic but nothing of the scheduling (tiles, barriers, staging, prefetch), so bitwise
equality between them is the scale test for races; the small-size tests pin both to the oracle."""
import numpy as np
import pytest

from cfd_b200 import abi, api

import cases

pytestmark = pytest.mark.gpu

SHAPES = [(258, 130, 100), (1026, 514, 1), (130, 258, 37)]


def run(kind, method, shape, steps, monkeypatch, env):
    for k in ("CFD_B200_NO_TMA", "CFD_B200_NO_RB_FUSED"):
        monkeypatch.delenv(k, raising=False)
    for k in env:
        monkeypatch.setenv(k, "1")
    g, f = cases.random_field(*shape, seed=17, scale=0.3)
    f.rho[...] = 1.0 + 0.1 * np.random.default_rng(3).random(f.rho.shape)
    p = cases.params(dt=5e-4, amp=(0.1, 0.05))
    if kind == "projection":
        kw = dict(max_iterations=12) if method in ("jacobi", "rbsor") else dict(max_iterations=8)
        api.set_poisson(abi.poisson_config(method, on_nonconvergence=abi.CFD_B200_NONCONV_CONTINUE, tolerance=0.0,
                                           absolute_tolerance=0.0, **kw))
    slv = api.GpuSolver(kind + "_gpu")
    try:
        for _ in range(steps):
            rc, _ = slv.step(f, g, p)
            assert rc == 0
    finally:
        slv.close()
        api.set_poisson("cg")
    return f


@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("kind,method", [("projection", "jacobi"), ("projection", "rbsor"), ("rk4", None), ("rk2", None)])
def test_tma_and_direct_kernels_agree_bitwise(monkeypatch, shape, kind, method):
    a = run(kind, method, shape, 2, monkeypatch, [])
    b = run(kind, method, shape, 2, monkeypatch, ["CFD_B200_NO_TMA"])
    for name in ("u", "v", "w", "p"):
        np.testing.assert_array_equal(getattr(a, name), getattr(b, name), err_msg=name)


@pytest.mark.parametrize("shape", [(258, 130, 100), (514, 66, 300)])
def test_one_pass_rbsor_equals_half_sweeps_at_scale(monkeypatch, shape):
    a = run("projection", "rbsor", shape, 2, monkeypatch, [])
    b = run("projection", "rbsor", shape, 2, monkeypatch, ["CFD_B200_NO_RB_FUSED"])
    c = run("projection", "rbsor", shape, 2, monkeypatch, [])  # and run to run
    for name in ("u", "v", "w", "p"):
        np.testing.assert_array_equal(getattr(a, name), getattr(b, name), err_msg=name)
        np.testing.assert_array_equal(getattr(a, name), getattr(c, name), err_msg=name)


@pytest.mark.parametrize("shape", SHAPES[:2])
@pytest.mark.parametrize("method", ["cg", "pcg", "bicgstab"])
def test_krylov_tma_and_direct_agree(monkeypatch, shape, method):
    """Reductions are ordered per tile grid, so the two paths differ in rounding only: 8 fixed
    iterations keep them within 1e-11 of each other."""
    a = run("projection", method, shape, 1, monkeypatch, [])
    b = run("projection", method, shape, 1, monkeypatch, ["CFD_B200_NO_TMA"])
    for name in ("u", "v", "w", "p"):
        x, y = getattr(a, name), getattr(b, name)
        scale = max(np.max(np.abs(y)), 1e-30)
        assert np.max(np.abs(x - y)) <= 1e-11 * scale, name


@pytest.mark.parametrize("shape", [(200, 190, 140), (2049, 2049, 1)])
def test_pageable_host_fields_take_the_staged_transfer_path(shape):
    """Fields of 32 MB and more in PAGEABLE host memory (what the reference's allocator hands out; numpy
    arrays here) move through the multi-threaded pinned staging pipeline (csrc/xfer.cu): round trip of all
    six fields bit for bit -- chunk sizes that do not divide the field, several fields per call -- and a
    time step that goes through the stateless drop-in symbol equals the same step on a device-resident
    session fed from pinned memory (the direct path)."""
    nx, ny, nz = shape
    g, f = cases.random_field(nx, ny, nz, seed=23, scale=0.2)
    f.rho[...] = 1.0 + 0.1 * np.random.default_rng(4).random(f.rho.shape)
    assert f.u.nbytes >= 32 << 20
    back = abi.HostField(nx, ny, nz)
    s = api.Session.for_grid(g)
    try:
        s.upload(f)
        s.download(back)
    finally:
        s.close()
    for k in f.NAMES:
        np.testing.assert_array_equal(getattr(back, k), getattr(f, k), err_msg=k)
    # same RK2 step from pageable memory (stateless symbol) and from pinned memory (session)
    p = cases.params(dt=5e-4, amp=(0.1, 0.05))
    p.max_iter = 1
    a = f.copy()
    assert api.solve_method_gpu("rk2", a, g, p) == 0
    pin = api.pinned_field(nx, ny, nz)
    try:
        for k in f.NAMES:
            getattr(pin, k)[...] = getattr(f, k)
        s = api.Session.for_grid(g)
        try:
            s.set_coords(g.x, g.y)  # the grid's own coordinate arrays, as the drop-in symbol uses them
            s.upload(pin)
            rc, _ = s.rk(2, p, 1)
            assert rc == 0
            s.download(pin)
        finally:
            s.close()
        for k in ("u", "v", "w", "p", "T"):
            np.testing.assert_array_equal(getattr(a, k), getattr(pin, k), err_msg=k)
    finally:
        pin.release()
