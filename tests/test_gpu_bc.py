"""GPU boundary-condition API on device pointers (cfd/boundary/boundary_conditions_gpu.cuh)
against the reference's own CPU implementation, compiled unmodified into oracle/_ref/libcfd_ref.so
(bc_apply_*_cpu).  Pure data movement, so the bar is bitwise, edges and corners included -- which is
where the reference's own 3-D GPU kernels deviate from its CPU passes."""
import ctypes as C

import numpy as np
import pytest

from cfd_b200 import abi, lib

import oracle

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def ref():
    try:
        r = oracle.Ref()
    except (FileNotFoundError, OSError) as e:
        pytest.skip("compiled reference not available: %s" % e)
    L = r.lib
    sz, dp = C.c_size_t, abi.c_double_p
    DV = C.POINTER(abi.DirichletValues)
    for name, args in {"bc_apply_scalar_cpu": [dp, sz, sz, C.c_int], "bc_apply_velocity_cpu": [dp, dp, sz, sz, C.c_int],
                       "bc_apply_dirichlet_scalar_cpu": [dp, sz, sz, DV],
                       "bc_apply_dirichlet_velocity_cpu": [dp, dp, sz, sz, DV, DV],
                       "bc_apply_inlet_cpu": [dp, dp, sz, sz, C.POINTER(abi.BcInletConfig)],
                       "bc_apply_scalar_cpu_3d": [dp, sz, sz, sz, sz, C.c_int],
                       "bc_apply_velocity_cpu_3d": [dp, dp, dp, sz, sz, sz, sz, C.c_int]}.items():
        fn = getattr(L, name)
        fn.restype, fn.argtypes = C.c_int, args
    return L


def dev(a):
    return torch.from_numpy(a.copy()).cuda()


def stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def back(t):
    torch.cuda.synchronize()
    return t.cpu().numpy()


def field(shape, seed):
    return np.random.default_rng(seed).standard_normal(shape)


SHAPES2 = [(3, 3), (4, 3), (17, 9), (64, 130), (257, 33)]          # (ny, nx)
SHAPES3 = [(3, 3, 3), (4, 5, 3), (9, 12, 17), (33, 20, 34)]        # (nz, ny, nx)
TYPES = [abi.BC_TYPE_NEUMANN, abi.BC_TYPE_PERIODIC, abi.BC_TYPE_DIRICHLET, abi.BC_TYPE_NOSLIP, abi.BC_TYPE_OUTLET]


def ref_type(t):  # the GPU entry points treat everything but PERIODIC as zero-gradient (boundary_conditions_gpu.cu:190-207)
    return t if t == abi.BC_TYPE_PERIODIC else abi.BC_TYPE_NEUMANN


@pytest.mark.parametrize("bct", TYPES)
@pytest.mark.parametrize("shape", SHAPES2)
def test_scalar_and_velocity_2d(ref, shape, bct):
    L = lib.load()
    ny, nx = shape
    a, u, v = field(shape, 1), field(shape, 2), field(shape, 3)
    ea, eu, ev = a.copy(), u.copy(), v.copy()
    assert ref.bc_apply_scalar_cpu(abi._ptr(ea), nx, ny, ref_type(bct)) == 0
    assert ref.bc_apply_velocity_cpu(abi._ptr(eu), abi._ptr(ev), nx, ny, ref_type(bct)) == 0
    da, du, dv = dev(a), dev(u), dev(v)
    L.bc_apply_scalar_gpu(da.data_ptr(), nx, ny, bct, stream())
    L.bc_apply_velocity_gpu(du.data_ptr(), dv.data_ptr(), nx, ny, bct, stream())
    np.testing.assert_array_equal(back(da), ea)
    np.testing.assert_array_equal(back(du), eu)
    np.testing.assert_array_equal(back(dv), ev)
    if bct == abi.BC_TYPE_NEUMANN:
        dn = dev(a)
        L.bc_apply_neumann_gpu(dn.data_ptr(), nx, ny, stream())
        np.testing.assert_array_equal(back(dn), ea)


@pytest.mark.parametrize("shape", SHAPES2)
def test_dirichlet_2d(ref, shape):
    L = lib.load()
    ny, nx = shape
    a, u, v = field(shape, 4), field(shape, 5), field(shape, 6)
    va = abi.DirichletValues(1.0, 2.0, 3.0, 4.0, 5.0, 6.0)
    vu = abi.DirichletValues(-1.0, -2.0, -3.0, -4.0, -5.0, -6.0)
    vv = abi.DirichletValues(0.5, 0.25, 0.125, 8.0, 16.0, 32.0)
    ea, eu, ev = a.copy(), u.copy(), v.copy()
    assert ref.bc_apply_dirichlet_scalar_cpu(abi._ptr(ea), nx, ny, C.byref(va)) == 0
    assert ref.bc_apply_dirichlet_velocity_cpu(abi._ptr(eu), abi._ptr(ev), nx, ny, C.byref(vu), C.byref(vv)) == 0
    da, du, dv = dev(a), dev(u), dev(v)
    L.bc_apply_dirichlet_scalar_gpu(da.data_ptr(), nx, ny, C.byref(va), stream())
    L.bc_apply_dirichlet_velocity_gpu(du.data_ptr(), dv.data_ptr(), nx, ny, C.byref(vu), C.byref(vv), stream())
    np.testing.assert_array_equal(back(da), ea)
    np.testing.assert_array_equal(back(du), eu)
    np.testing.assert_array_equal(back(dv), ev)


@pytest.mark.parametrize("bct", TYPES[:3])
@pytest.mark.parametrize("shape", SHAPES3)
def test_scalar_and_velocity_3d(ref, shape, bct):
    L = lib.load()
    nz, ny, nx = shape
    a, u, v, w = (field(shape, s) for s in (7, 8, 9, 10))
    ea, eu, ev, ew = a.copy(), u.copy(), v.copy(), w.copy()
    assert ref.bc_apply_scalar_cpu_3d(abi._ptr(ea), nx, ny, nz, nx * ny, ref_type(bct)) == 0
    assert ref.bc_apply_velocity_cpu_3d(abi._ptr(eu), abi._ptr(ev), abi._ptr(ew), nx, ny, nz, nx * ny, ref_type(bct)) == 0
    da, du, dv, dw = dev(a), dev(u), dev(v), dev(w)
    L.bc_apply_scalar_3d_gpu(da.data_ptr(), nx, ny, nz, bct, stream())
    L.bc_apply_velocity_3d_gpu(du.data_ptr(), dv.data_ptr(), dw.data_ptr(), nx, ny, nz, bct, stream())
    for got, exp in ((da, ea), (du, eu), (dv, ev), (dw, ew)):
        np.testing.assert_array_equal(back(got), exp)


def test_3d_entry_points_delegate_and_reject_like_the_reference(ref):
    L = lib.load()
    a = field((1, 9, 12), 11)
    e = a.copy()
    assert ref.bc_apply_scalar_cpu(abi._ptr(e), 12, 9, abi.BC_TYPE_PERIODIC) == 0
    d = dev(a)
    L.bc_apply_scalar_3d_gpu(d.data_ptr(), 12, 9, 1, abi.BC_TYPE_PERIODIC, stream())   # nz == 1 -> the 2-D function
    np.testing.assert_array_equal(back(d), e)
    b = field((2, 9, 12), 12)
    d = dev(b)
    L.bc_apply_scalar_3d_gpu(d.data_ptr(), 12, 9, 2, abi.BC_TYPE_NEUMANN, stream())    # nz == 2 -> untouched
    L.bc_apply_scalar_gpu(d.data_ptr(), 2, 9, abi.BC_TYPE_NEUMANN, stream())           # nx < 3  -> untouched
    L.bc_apply_scalar_gpu(None, 12, 9, abi.BC_TYPE_NEUMANN, stream())                  # NULL    -> no-op
    np.testing.assert_array_equal(back(d), b)


def inlet(edge, profile, spec, *vals):
    c = abi.BcInletConfig()
    c.edge, c.profile, c.spec_type = edge, profile, spec
    for k, x in enumerate(vals):
        c.spec.mass_flow[k] = x  # the union members overlay each other
    return c


EDGES = [abi.BC_EDGE_LEFT, abi.BC_EDGE_RIGHT, abi.BC_EDGE_BOTTOM, abi.BC_EDGE_TOP]


@pytest.mark.parametrize("edge", EDGES)
@pytest.mark.parametrize("profile", [abi.BC_INLET_PROFILE_UNIFORM, abi.BC_INLET_PROFILE_PARABOLIC])
@pytest.mark.parametrize("spec,vals", [(abi.BC_INLET_SPEC_VELOCITY, (1.25, -0.5)),
                                       (abi.BC_INLET_SPEC_MAGNITUDE_DIR, (2.0, 0.7)),
                                       (abi.BC_INLET_SPEC_MASS_FLOW, (3.0, 1.2, 0.8))])
def test_inlet(ref, edge, profile, spec, vals):
    L = lib.load()
    ny, nx = 33, 20
    u, v = field((ny, nx), 13), field((ny, nx), 14)
    cfg = inlet(edge, profile, spec, *vals)
    eu, ev = u.copy(), v.copy()
    assert ref.bc_apply_inlet_cpu(abi._ptr(eu), abi._ptr(ev), nx, ny, C.byref(cfg)) == 0
    du, dv = dev(u), dev(v)
    assert L.bc_apply_inlet_gpu(du.data_ptr(), dv.data_ptr(), nx, ny, C.byref(cfg), stream()) == 0
    np.testing.assert_array_equal(back(du), eu)
    np.testing.assert_array_equal(back(dv), ev)


def test_inlet_custom_profile_and_errors(ref):
    L = lib.load()
    ny, nx = 17, 12
    u, v = field((ny, nx), 15), field((ny, nx), 16)

    @abi.INLET_PROFILE_FN
    def prof(pos, uo, vo, _):
        uo[0] = 3.0 * pos * pos
        vo[0] = -pos

    cfg = inlet(abi.BC_EDGE_LEFT, abi.BC_INLET_PROFILE_CUSTOM, abi.BC_INLET_SPEC_VELOCITY, 1.0, 0.0)
    cfg.custom_profile = prof
    eu, ev = u.copy(), v.copy()
    assert ref.bc_apply_inlet_cpu(abi._ptr(eu), abi._ptr(ev), nx, ny, C.byref(cfg)) == 0
    du, dv = dev(u), dev(v)
    assert L.bc_apply_inlet_gpu(du.data_ptr(), dv.data_ptr(), nx, ny, C.byref(cfg), stream()) == 0  # host-evaluated
    np.testing.assert_array_equal(back(du), eu)
    np.testing.assert_array_equal(back(dv), ev)
    bad = inlet(abi.BC_EDGE_FRONT, abi.BC_INLET_PROFILE_UNIFORM, abi.BC_INLET_SPEC_VELOCITY, 1.0, 0.0)
    assert L.bc_apply_inlet_gpu(du.data_ptr(), dv.data_ptr(), nx, ny, C.byref(bad), stream()) == abi.CFD_ERROR_INVALID
    bad = inlet(abi.BC_EDGE_LEFT, abi.BC_INLET_PROFILE_UNIFORM, abi.BC_INLET_SPEC_MASS_FLOW, 1.0, 0.0, 1.0)
    assert L.bc_apply_inlet_gpu(du.data_ptr(), dv.data_ptr(), nx, ny, C.byref(bad), stream()) == abi.CFD_ERROR_INVALID
    assert L.bc_apply_inlet_gpu(None, dv.data_ptr(), nx, ny, C.byref(cfg), stream()) == abi.CFD_ERROR_INVALID
