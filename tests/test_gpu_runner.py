"""What stands behind one reference-facing call (csrc/gang.cu, csrc/host_sync.cu):

  * host-sync policies of the ns_solver_t objects: UPLOAD_BOUNDARY and DOWNLOAD_LAZY against
    the default FULL mode, bit for bit, on the reference's cavity driver loop
    (tests/validation/lid_driven_cavity_common.h:306-338);
  * the runner cache of the stateless solve_*_method_gpu symbols (alternating shapes, two threads);
  * the multi-GPU gang behind the SAME symbols (SURVEY 8e process model): cfd_b200_set_device_count(G) /
    CFD_B200_NGPU=G, even and uneven slabs, z and y decomposition, against the UNDECOMPOSED oracle -- bitwise for
    Jacobi / Red-Black SOR / RK / Euler, <= 1e-8 for the Krylov methods -- and through the UNMODIFIED reference
    registry.  Needs G visible GPUs, or virtual ranks on one GPU (CFD_B200_VIRTUAL_RANKS=1, peer transport).
"""
import os
import threading

import numpy as np
import pytest

import cases
from cfd_b200 import abi, api

pytestmark = pytest.mark.gpu
FIELDS = ("u", "v", "w", "p", "T")


@pytest.fixture
def clean():
    yield
    api.set_poisson("cg")
    api.set_device_count(0)
    api.load().cfd_b200_release_cached_sessions()


def bitwise(a, b, what):
    for k in FIELDS:
        np.testing.assert_array_equal(getattr(a, k), getattr(b, k), err_msg="%s: %s" % (what, k))


def close(a, b, tol, what):
    for k in FIELDS:
        x, y = getattr(a, k), getattr(b, k)
        if np.max(np.abs(y)) < 1e-12:
            assert np.max(np.abs(x - y)) <= tol, (what, k)
            continue
        l2, li = cases.rel_err(x, y)
        assert l2 <= tol and li <= tol, (what, k, l2, li)


# ------------------------------------------------------------------------------------------------
# host-sync policies
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("case", ["cavity2d_33", "cavity2d_66", "cavity3d", "natconv2d", "natconv3d"])
@pytest.mark.parametrize("method", ["jacobi", "cg"])
def test_upload_boundary_equals_full_mode(clean, case, method):
    """The driver loop of the reference's validation cases: BCs on the host, one step, repeat.  With
    UPLOAD_BOUNDARY only the six faces travel to the device after the first step; results are those of FULL mode
    bit for bit (both run the same kernels on the same data)."""
    def make():
        if case == "cavity2d_33":
            return cases.cavity(33, 1, dt=5e-4)
        if case == "cavity2d_66":
            return cases.cavity(66, 1, dt=2e-4)
        if case == "cavity3d":
            return cases.cavity(22, 19, dt=5e-4)
        if case == "natconv2d":
            return cases.natural_convection(41, 1, dt=2e-3)
        return cases.natural_convection(18, 14, dt=1e-3)
    api.set_poisson(abi.poisson_config(method, max_iterations=60, on_nonconvergence=abi.CFD_B200_NONCONV_CONTINUE))
    out = []
    for flags in (api.SYNC_FULL, api.SYNC_UPLOAD_BOUNDARY):
        g, f, p = make()
        slv = api.GpuSolver("projection_gpu")
        slv.set_host_sync(flags)
        try:
            for step in range(12):
                if case.startswith("cavity"):
                    cases.apply_cavity_bc(f, lid=1.0 + 0.01 * step)  # the boundary values really change
                rc, _ = slv.step(f, g, p)
                assert rc == 0
        finally:
            slv.close()
        out.append(f)
    bitwise(out[1], out[0], "%s %s UPLOAD_BOUNDARY vs FULL" % (case, method))
    assert np.max(np.abs(out[0].u)) > 0 or np.max(np.abs(out[0].v)) > 0


@pytest.mark.parametrize("name", ["projection_gpu", "rk4_gpu", "explicit_euler_gpu"])
def test_download_lazy_keeps_the_state_on_the_device(clean, orc, name):
    """DOWNLOAD_LAZY: ten step() calls move nothing but the first upload; cfd_b200_solver_sync_host() then
    delivers exactly what ten FULL-mode steps (and ten oracle steps) deliver."""
    api.set_poisson(abi.poisson_config("jacobi", max_iterations=40, on_nonconvergence=abi.CFD_B200_NONCONV_CONTINUE))
    g, f = cases.random_field(26, 22, 14, seed=8, scale=0.05)
    f.rho[...] = 1.0 + 0.1 * np.random.default_rng(3).random(f.rho.shape)
    p = cases.params(dt=1e-3)  # amplitudes 0: step() restarts the source decay at iter = 0 anyway
    f_full, f_lazy, fo = f.copy(), f.copy(), f.copy()
    a, b = api.GpuSolver(name), api.GpuSolver(name)
    b.set_host_sync(api.SYNC_DOWNLOAD_LAZY)
    try:
        for step in range(10):
            assert a.step(f_full, g, p)[0] == 0
            assert b.step(f_lazy, g, p)[0] == 0
        np.testing.assert_array_equal(f_lazy.u, f.u)  # the host copy was never touched
        with pytest.raises(Exception):
            b.set_host_sync(api.SYNC_FULL)  # refused while the device is ahead
        b.sync_host(f_lazy)
        b.set_host_sync(api.SYNC_FULL)
    finally:
        a.close()
        b.close()
    bitwise(f_lazy, f_full, name + " lazy vs full")
    np.testing.assert_array_equal(f_lazy.rho, f_full.rho)
    pc = abi.poisson_config("jacobi", max_iterations=40, on_nonconvergence=abi.CFD_B200_NONCONV_CONTINUE)
    for step in range(10):
        if name == "projection_gpu":
            assert orc.projection(fo, g, p, pc, 1)[0] == 0
        elif name == "rk4_gpu":
            assert orc.rk(fo, g, p, 4, 1) == 0
        else:
            assert orc.explicit_euler(fo, g, p, 1) == 0
    bitwise(f_full, fo, name + " full vs oracle")


def test_host_sync_is_refused_on_foreign_objects(clean):
    L = api.load()
    ps = L.create_cg_gpu_solver()
    import ctypes as C
    rc = L.cfd_b200_solver_set_host_sync(C.cast(ps, C.POINTER(abi.NSSolver)), 1)
    assert rc == abi.CFD_ERROR_INVALID
    C.CDLL(None).free(C.cast(ps, C.c_void_p))
    slv = api.GpuSolver("rk2_gpu")
    try:
        assert L.cfd_b200_solver_set_host_sync(slv._slv, 8) == abi.CFD_ERROR_INVALID
    finally:
        slv.close()


# ------------------------------------------------------------------------------------------------
# runner cache of the stateless symbols
# ------------------------------------------------------------------------------------------------
def test_stateless_symbols_cache_two_shapes_and_run_concurrently(clean, orc):
    api.set_poisson(abi.poisson_config("jacobi", max_iterations=30, on_nonconvergence=abi.CFD_B200_NONCONV_CONTINUE))
    pc = abi.poisson_config("jacobi", max_iterations=30, on_nonconvergence=abi.CFD_B200_NONCONV_CONTINUE)
    shapes = [(30, 22, 12), (44, 18, 1)]
    work = []
    for q, sh in enumerate(shapes):
        g, f = cases.random_field(*sh, seed=30 + q, scale=0.05)
        p = cases.params(dt=1e-3, amp=(0.1, 0.05))
        p.max_iter = 1
        work.append((g, f, f.copy(), p))
    errs = []

    def loop(g, f, fo, p):
        try:
            for _ in range(6):
                rc = api.solve_method_gpu("projection", f, g, p)
                if rc != 0:
                    errs.append(rc)
        except Exception as e:  # pragma: no cover
            errs.append(e)

    th = [threading.Thread(target=loop, args=w) for w in work]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert not errs
    for g, f, fo, p in work:
        for _ in range(6):
            assert orc.projection(fo, g, p, pc, 1)[0] == 0
        bitwise(f, fo, "cached stateless projection %dx%dx%d" % (g.nx, g.ny, g.nz))


# ------------------------------------------------------------------------------------------------
# the gang: G GPUs behind the drop-in symbols
# ------------------------------------------------------------------------------------------------
def _gang_or_skip(G, monkeypatch):
    import torch
    if torch.cuda.device_count() >= G:
        return
    if os.environ.get("CFD_B200_NO_VIRTUAL_RANKS"):
        pytest.skip("needs %d GPUs" % G)
    monkeypatch.setenv("CFD_B200_VIRTUAL_RANKS", "1")  # G slabs on one GPU, one stream each (peer transport)


GANG_CASES = [
    # (name, kind, method, shape, tol, poisson kwargs)
    ("jacobi z-slabs even", "projection", "jacobi", (34, 22, 26), 0.0, dict(max_iterations=30)),
    ("jacobi z-slabs uneven", "projection", "jacobi", (34, 22, 29), 0.0, dict(max_iterations=30)),
    ("rbsor z-slabs uneven", "projection", "rbsor", (34, 22, 31), 0.0, dict(max_iterations=25)),
    ("rbsor y-slabs uneven", "projection", "rbsor", (36, 41, 1), 0.0, dict(max_iterations=25)),
    # one-pass sweep on slabs (second halo layer): two own planes per rank at G = 4; 2 + 1 + 1 + 1 own planes (the
    # ranks agree on the half-sweeps instead); the reference's own stopping rule
    ("rbsor z-slabs two planes each", "projection", "rbsor", (34, 22, 10), 0.0, dict(max_iterations=25)),
    ("rbsor z-slabs too thin", "projection", "rbsor", (34, 22, 7), 0.0, dict(max_iterations=25)),
    ("rbsor z-slabs reference rule", "projection", "rbsor", (34, 22, 31), 0.0, dict(max_iterations=400)),
    ("multi-tile rbsor z-slabs", "projection", "rbsor", (130, 34, 83), 0.0, dict(max_iterations=12)),
    ("cg z-slabs uneven", "projection", "cg", (34, 22, 29), 1e-8, dict(max_iterations=60)),
    ("pcg y-slabs", "projection", "pcg", (36, 42, 1), 1e-8, dict(max_iterations=60)),
    ("bicgstab x12 z-slabs", "projection", "bicgstab", (34, 22, 27), 1e-8, dict(max_iterations=12)),
    ("multi-tile jacobi", "projection", "jacobi", (130, 34, 83), 0.0, dict(max_iterations=12)),
    ("multi-tile bicgstab x12", "projection", "bicgstab", (130, 34, 83), 1e-8, dict(max_iterations=12)),
    ("multi-tile rbsor y-slabs", "projection", "rbsor", (130, 71, 1), 0.0, dict(max_iterations=12)),
    ("rk4 z-slabs uneven", "rk4", None, (34, 22, 29), 0.0, {}),
    ("rk2 y-slabs uneven", "rk2", None, (36, 41, 1), 0.0, {}),
    ("euler z-slabs uneven", "explicit_euler", None, (34, 22, 29), 0.0, {}),
]


@pytest.mark.parametrize("G", [2, 3, 4])
@pytest.mark.parametrize("case", GANG_CASES, ids=[c[0].replace(" ", "_") for c in GANG_CASES])
def test_gang_behind_the_stateless_symbols_equals_the_undecomposed_oracle(clean, orc, monkeypatch, G, case):
    _gang_or_skip(G, monkeypatch)
    name, kind, method, shape, tol, pkw = case
    g, f = cases.random_field(*shape, seed=21, scale=0.05)
    f.rho[...] = 1.0 + 0.1 * np.random.default_rng(2).random(f.rho.shape)
    p = cases.params(dt=1e-3, amp=(0.1, 0.05))
    p.max_iter = 2
    fo = f.copy()
    pc = abi.poisson_config(method or "cg", on_nonconvergence=abi.CFD_B200_NONCONV_CONTINUE, **pkw)
    api.set_poisson(pc)
    api.set_device_count(G)
    assert api.get_device_count() == G
    rc = api.solve_method_gpu(kind, f, g, p)
    if kind == "projection":
        rco, _ = orc.projection(fo, g, p, pc, 2)
    elif kind == "explicit_euler":
        rco = orc.explicit_euler(fo, g, p, 2)
    else:
        rco = orc.rk(fo, g, p, int(kind[2]), 2)
    assert rc == rco == 0, api.last_error()
    if tol == 0.0:
        bitwise(f, fo, "%s G=%d" % (name, G))
        if kind != "projection":
            np.testing.assert_array_equal(f.rho, fo.rho)
    else:
        close(f, fo, tol, "%s G=%d" % (name, G))


@pytest.mark.parametrize("G", [2, 3])
@pytest.mark.parametrize("n,steps", [(64, 3)])
def test_gang_c3_cavity_rbsor_one_pass_and_half_sweeps(clean, orc, monkeypatch, G, n, steps):
    """C3 shape on slabs: lid-driven cavity, Red-Black SOR x 30, bitwise per step against the undecomposed oracle --
    with the one-pass sweep (two-layer exchange, poisson_tma3.inc) and with the half-sweeps it replaces."""
    _gang_or_skip(G, monkeypatch)
    pc = abi.poisson_config("rbsor", max_iterations=30, on_nonconvergence=abi.CFD_B200_NONCONV_CONTINUE)
    api.set_poisson(pc)
    api.set_device_count(G)
    for path in ("one_pass", "half_sweeps"):
        if path == "half_sweeps":
            monkeypatch.setenv("CFD_B200_NO_RB_FUSED", "1")
        else:
            monkeypatch.delenv("CFD_B200_NO_RB_FUSED", raising=False)
        g, f, p = cases.cavity(n, n, dt=1e-4)
        fo = f.copy()
        slv = api.GpuSolver("projection_gpu")
        try:
            for step in range(steps):
                cases.apply_cavity_bc(f)
                cases.apply_cavity_bc(fo)
                rc, _ = slv.step(f, g, p)
                rco, _ = orc.projection(fo, g, p, pc, 1)
                assert rc == rco == 0, api.last_error()
                bitwise(f, fo, "C3 %d^3 G=%d %s step %d" % (n, G, path, step))
        finally:
            slv.close()


@pytest.mark.parametrize("case", [c for c in GANG_CASES if c[0] in ("jacobi z-slabs uneven", "rbsor y-slabs uneven",
                                                                     "rbsor z-slabs uneven",
                                                                     "multi-tile bicgstab x12", "pcg y-slabs", "rk4 z-slabs uneven")],
                         ids=lambda c: c[0].replace(" ", "_"))
def test_gang_over_the_nccl_fallback(clean, orc, monkeypatch, case):
    """The library transport (ncclCommInitAll, ncclSend/Recv, ncclAllReduce) behind the same gang: what runs where
    devices cannot map each other's memory.  Needs two real GPUs (NCCL cannot put two ranks on one device)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    monkeypatch.setenv("CFD_B200_GANG_TRANSPORT", "nccl")
    name, kind, method, shape, tol, pkw = case
    g, f = cases.random_field(*shape, seed=21, scale=0.05)
    f.rho[...] = 1.0 + 0.1 * np.random.default_rng(2).random(f.rho.shape)
    p = cases.params(dt=1e-3, amp=(0.1, 0.05))
    p.max_iter = 2
    fo = f.copy()
    pc = abi.poisson_config(method or "cg", on_nonconvergence=abi.CFD_B200_NONCONV_CONTINUE, **pkw)
    api.set_poisson(pc)
    api.set_device_count(2)
    rc = api.solve_method_gpu(kind, f, g, p)
    rco = orc.projection(fo, g, p, pc, 2)[0] if kind == "projection" else orc.rk(fo, g, p, int(kind[2]), 2)
    assert rc == rco == 0, api.last_error()
    if tol == 0.0:
        bitwise(f, fo, "%s over NCCL" % name)
    else:
        close(f, fo, tol, "%s over NCCL" % name)


@pytest.mark.parametrize("G", [2, 4])
def test_gang_natural_convection_y_slabs(clean, orc, monkeypatch, G):
    """C4 shape: energy + buoyancy + thermal BCs, PCG, y-slabs, persistent solver object with stats."""
    _gang_or_skip(G, monkeypatch)
    g, f, p = cases.natural_convection(130, 1, dt=2e-4)
    pc = abi.poisson_config("pcg", max_iterations=60, on_nonconvergence=abi.CFD_B200_NONCONV_CONTINUE)
    api.set_poisson(pc)
    api.set_device_count(G)
    fo = f.copy()
    slv = api.GpuSolver("projection_gpu")
    try:
        for step in range(3):
            rc, st = slv.step(f, g, p)
            rco, _ = orc.projection(fo, g, p, pc, 1)
            assert rc == rco == 0, api.last_error()
            close(f, fo, 1e-8, "natconv gang G=%d step %d" % (G, step))
            mv, mp, mt = (np.zeros(1) for _ in range(3))
            orc.lib.orc_field_stats(fo.c, abi._ptr(mv), abi._ptr(mp), abi._ptr(mt))
            assert abs(st.max_velocity - mv[0]) <= 1e-8 * max(mv[0], 1e-30)
            assert abs(st.max_temperature - mt[0]) <= 1e-12 * abs(mt[0])
    finally:
        slv.close()


@pytest.mark.parametrize("G", [2, 4])
def test_gang_nan_in_one_slab_takes_every_rank_out(clean, orc, monkeypatch, G):
    """A NaN seeded in the last slab's pressure: every rank must leave at the same step with DIVERGED (the
    flag is all-reduced), not hang in the next halo exchange."""
    _gang_or_skip(G, monkeypatch)
    g, f = cases.random_field(30, 20, 34, seed=5, scale=0.05)
    f.rho[...] = 1.0
    f.p[30, 7, 9] = np.nan
    p = cases.params(dt=1e-3)
    p.max_iter = 3
    fo = f.copy()
    api.set_device_count(G)
    rc = api.solve_method_gpu("rk4", f, g, p)
    rco = orc.rk(fo, g, p, 4, 3)
    assert rc == rco == abi.CFD_ERROR_DIVERGED


@pytest.mark.parametrize("G", [2, 4])
@pytest.mark.parametrize("name,kind", [("projection_gpu", "projection"), ("rk4_gpu", "rk4")])
def test_gang_through_the_reference_registry(clean, orc, monkeypatch, G, name, kind):
    """run_simulation_step's path (simulation_api.c:185-202 -> solver_registry.c:1121-1139) with the
    UNMODIFIED reference core/api linked against libcfd_b200.so, CFD_B200_NGPU=G in the environment."""
    import oracle
    _gang_or_skip(G, monkeypatch)
    try:
        ref = oracle.Ref("dropin")
    except (FileNotFoundError, OSError) as e:
        pytest.skip(str(e))
    monkeypatch.setenv("CFD_B200_NGPU", str(G))
    api.set_device_count(0)  # the environment decides
    pc = abi.poisson_config("jacobi", max_iterations=30, on_nonconvergence=abi.CFD_B200_NONCONV_CONTINUE)
    api.set_poisson(pc)
    g, f = cases.random_field(26, 18, 23, seed=4, scale=0.05)
    p = cases.params(dt=1e-3, amp=(0.1, 0.05))
    fo = f.copy()
    rc, st = ref.run_solver(name, f, g, p, nsteps=2)
    assert rc == 0
    for _ in range(2):
        rco = orc.projection(fo, g, p, pc, 1)[0] if kind == "projection" else orc.rk(fo, g, p, 4, 1)
        assert rco == 0
    bitwise(f, fo, "registry %s with CFD_B200_NGPU=%d" % (name, G))
