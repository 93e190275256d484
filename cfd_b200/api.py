"""Host-side mirror of the reference's solver interface for the projection hot path.

Mirrors (same names, argument meaning, status codes):
  solver_step / solver_solve on the "projection_gpu" / "rk2_gpu" / "rk4_gpu" names
      (lib/src/api/solver_registry.c:1121-1291)  ->  GpuSolver.step / GpuSolver.solve
  poisson_solver_create(method, POISSON_BACKEND_GPU) + init + solve
      (lib/src/solvers/linear/linear_solver.c:150-286, 487-509)  ->  PoissonGpuSolver
plus the device-resident Session used for multi-step and multi-GPU runs.
Everything computes through the C-ABI of libcfd_b200.so; there is no CPU path here.
"""
import ctypes as C

import numpy as np

from . import abi
from .lib import CfdB200Error, check, last_error, load

FM_U, FM_V, FM_W, FM_P, FM_RHO, FM_T, FM_ALL = 1, 2, 4, 8, 16, 32, 63
SYNC_FULL, SYNC_UPLOAD_BOUNDARY, SYNC_DOWNLOAD_LAZY = 0, 1, 2


def set_device_count(n):
    """GPUs behind every drop-in call of this process (cfd_b200_set_device_count; 0 = CFD_B200_NGPU or 1)."""
    check(load().cfd_b200_set_device_count(int(n)), "cfd_b200_set_device_count")


def set_arithmetic(fast):
    """Arithmetic policy of sessions / solvers created afterwards: 0 exact (default), 1 fast, -1 environment."""
    check(load().cfd_b200_set_arithmetic(int(fast)), "cfd_b200_set_arithmetic")


def get_device_count():
    return int(load().cfd_b200_get_device_count())


def gpu_is_available():
    return bool(load().gpu_is_available())


def set_poisson(method="cg", **kw):
    """Select the Poisson method of projection_gpu (the reference hard-codes CG)."""
    cfg = abi.poisson_config(method, **kw) if isinstance(method, str) else method
    check(load().cfd_b200_set_poisson_config(C.byref(cfg)), "cfd_b200_set_poisson_config")
    return cfg


class GpuSolver:
    """ns_solver_t for one of the GPU names, driven like the reference drives it:
    host flow_field in, host flow_field out (navier_stokes_solver.h:227-232)."""

    NAMES = {"projection_gpu": "cfd_b200_create_projection_gpu_solver",
             "explicit_euler_gpu": "cfd_b200_create_explicit_euler_gpu_solver",
             "rk2_gpu": "cfd_b200_create_rk2_gpu_solver",
             "rk4_gpu": "cfd_b200_create_rk4_gpu_solver"}

    def __init__(self, name):
        if name not in self.NAMES:
            raise KeyError("unknown GPU solver name %r" % name)
        L = load()
        self._slv = getattr(L, self.NAMES[name])()
        if not self._slv:
            raise CfdB200Error(abi.CFD_ERROR_UNSUPPORTED, "cfd_solver_create(%s)" % name, last_error())
        s = self._slv.contents
        FF, GR, NP, NS = (C.POINTER(abi.FlowField), C.POINTER(abi.Grid), C.POINTER(abi.NSSolverParams),
                          C.POINTER(abi.NSSolverStats))
        fn = C.CFUNCTYPE(C.c_int, C.POINTER(abi.NSSolver), FF, GR, NP, NS)
        self._step, self._solve = fn(s.step), fn(s.solve)
        self._init = C.CFUNCTYPE(C.c_int, C.POINTER(abi.NSSolver), GR, NP)(s.init)
        self._destroy = C.CFUNCTYPE(None, C.POINTER(abi.NSSolver))(s.destroy)
        self.name = s.name.decode()

    def init(self, grid, params):
        return self._init(self._slv, C.byref(grid.c), C.byref(params))

    def step(self, field, grid, params):
        st = abi.NSSolverStats()
        rc = self._step(self._slv, C.byref(field.c), C.byref(grid.c), C.byref(params), C.byref(st))
        return rc, st

    def solve(self, field, grid, params):
        st = abi.NSSolverStats()
        rc = self._solve(self._slv, C.byref(field.c), C.byref(grid.c), C.byref(params), C.byref(st))
        return rc, st

    def set_host_sync(self, flags):
        """SYNC_FULL | SYNC_UPLOAD_BOUNDARY | SYNC_DOWNLOAD_LAZY (cfd_b200_solver_set_host_sync)."""
        check(load().cfd_b200_solver_set_host_sync(self._slv, int(flags)), "cfd_b200_solver_set_host_sync")

    def sync_host(self, field):
        check(load().cfd_b200_solver_sync_host(self._slv, C.byref(field.c)), "cfd_b200_solver_sync_host")

    def close(self):
        if self._slv:
            self._destroy(self._slv)
            C.CDLL(None).free(C.cast(self._slv, C.c_void_p))
            self._slv = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def solve_method_gpu(kind, field, grid, params, config=None):
    """The stateless link-time entry points solve_{projection,rk2,rk4}_method_gpu
    (cfd/core/gpu_device.h:221-249).  Runs params.max_iter steps."""
    L = load()
    fn = {"projection": L.solve_projection_method_gpu, "rk2": L.solve_rk2_method_gpu,
          "rk4": L.solve_rk4_method_gpu, "explicit_euler": L.solve_explicit_euler_method_gpu}[kind]
    if config is None:
        config = L.gpu_config_default()
        config.min_grid_size, config.min_steps = 1, 1  # as the registry wrappers do (:1131-1133)
    return fn(C.byref(field.c), C.byref(grid.c), C.byref(params), C.byref(config))


class PoissonGpuSolver:
    """poisson_solver_t with POISSON_BACKEND_GPU (create / init / solve / destroy)."""

    FACTORIES = {abi.POISSON_METHOD_JACOBI: "create_jacobi_gpu_solver",
                 abi.POISSON_METHOD_REDBLACK_SOR: "create_redblack_gpu_solver",
                 abi.POISSON_METHOD_CG: "create_cg_gpu_solver",
                 abi.POISSON_METHOD_BICGSTAB: "create_bicgstab_gpu_solver",
                 abi.POISSON_METHOD_MULTIGRID: "cfd_b200_create_multigrid_gpu_solver"}

    def __init__(self, method, nx, ny, nz, dx, dy, dz, params=None):
        L = load()
        self._s = getattr(L, self.FACTORIES[method])()
        if not self._s:
            raise CfdB200Error(abi.CFD_ERROR_UNSUPPORTED, "poisson_solver_create", last_error())
        s = self._s.contents
        # poisson_solver_init (linear_solver.c:245-286): fill extents/params, then call init
        s.nx, s.ny, s.nz, s.dx, s.dy, s.dz = nx, ny, nz, dx, dy, dz
        if params is not None:
            s.params = params
        else:
            s.params = abi.poisson_params_default()
            if method == abi.POISSON_METHOD_JACOBI:
                s.params.max_iterations = 2000
        PSP = C.POINTER(abi.PoissonSolver)
        init = C.CFUNCTYPE(C.c_int, PSP, C.c_size_t, C.c_size_t, C.c_size_t, C.c_double, C.c_double,
                           C.c_double, C.POINTER(abi.PoissonParams))(s.init)
        self._solve = C.CFUNCTYPE(C.c_int, PSP, abi.c_double_p, abi.c_double_p, abi.c_double_p,
                                  C.POINTER(abi.PoissonStats))(s.solve)
        self._destroy = C.CFUNCTYPE(None, PSP)(s.destroy)
        check(init(self._s, nx, ny, nz, dx, dy, dz, C.byref(s.params)), "poisson_solver_init")

    def solve(self, x, rhs, x_temp=None):
        st = abi.PoissonStats()
        rc = self._solve(self._s, abi._ptr(x), abi._ptr(x_temp) if x_temp is not None else None,
                         abi._ptr(rhs), C.byref(st))
        return rc, st

    def close(self):
        if self._s:
            self._destroy(self._s)
            C.CDLL(None).free(C.cast(self._s, C.c_void_p))
            self._s = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Session:
    """Device-resident slab (cfd_b200_session_t)."""

    def __init__(self, nx, ny, nz, dx, dy, dz=0.0, xmin=0.0, ymin=0.0, zmin=0.0, rank=0, nranks=1,
                 device=-1):
        self.L = load()
        d = abi.SessionDesc(nx, ny, nz, dx, dy, dz, xmin, ymin, zmin, rank, nranks, device)
        self.desc = d
        self.h = self.L.cfd_b200_session_create(C.byref(d))
        if not self.h:
            raise CfdB200Error(abi.CFD_ERROR, "cfd_b200_session_create", last_error())
        self.ncell = nx * ny * nz

    @classmethod
    def for_grid(cls, grid, **kw):
        dx, dy, dz = grid.spacing
        return cls(grid.nx, grid.ny, grid.nz, dx, dy, dz, float(grid.x[0]), float(grid.y[0]),
                   float(grid.z[0]) if grid.nz > 1 else 0.0, **kw)

    def upload(self, field, mask=FM_ALL):
        check(self.L.cfd_b200_session_upload_mask(self.h, C.byref(field.c), mask), "session_upload")

    def download(self, field, mask=FM_ALL):
        check(self.L.cfd_b200_session_download_mask(self.h, C.byref(field.c), mask), "session_download")

    def set_slab(self, global_layers, global_lo):
        check(self.L.cfd_b200_session_set_slab(self.h, global_layers, global_lo), "session_set_slab")

    def set_coords(self, x=None, y=None):
        """Exact cell coordinates of this slab (slices of the global grid arrays)."""
        self._xy = (np.ascontiguousarray(x, dtype=np.float64) if x is not None else None,
                    np.ascontiguousarray(y, dtype=np.float64) if y is not None else None)
        check(self.L.cfd_b200_session_set_coords(self.h, abi._ptr(self._xy[0]) if x is not None else None,
                                                 abi._ptr(self._xy[1]) if y is not None else None), "set_coords")

    def init_synthetic(self, seed=12345, global_k0=0, global_nz=None):
        check(self.L.cfd_b200_session_init_synthetic(self.h, seed, global_k0,
                                                     global_nz or self.desc.nz), "init_synthetic")

    def projection(self, params, pcfg=None, nsteps=1):
        st = abi.NSSolverStats()
        rc = self.L.cfd_b200_session_projection(self.h, C.byref(params),
                                                C.byref(pcfg) if pcfg is not None else None, nsteps,
                                                C.byref(st))
        return rc, st

    def rk(self, order, params, nsteps=1):
        st = abi.NSSolverStats()
        rc = self.L.cfd_b200_session_rk(self.h, order, C.byref(params), nsteps, C.byref(st))
        return rc, st

    def poisson(self, pcfg, x=None, rhs=None):
        if x is not None or rhs is not None:
            check(self.L.cfd_b200_session_upload_poisson(
                self.h, abi._ptr(x) if x is not None else None,
                abi._ptr(rhs) if rhs is not None else None), "upload_poisson")
        st = abi.PoissonStats()
        rc = self.L.cfd_b200_session_poisson(self.h, C.byref(pcfg), C.byref(st))
        return rc, st

    def download_x(self, x):
        check(self.L.cfd_b200_session_download_poisson(self.h, abi._ptr(x)), "download_poisson")

    def last_poisson_stats(self):
        return self.L.cfd_b200_session_last_poisson_stats(self.h)

    def peer_export(self):
        out = C.create_string_buffer(64)
        check(self.L.cfd_b200_session_peer_export(self.h, out), "session_peer_export")
        return out.raw

    def peer_attach(self, handles):
        """handles: the 64-byte exports of ALL ranks, concatenated in rank order."""
        return self.L.cfd_b200_session_peer_attach(self.h, handles)

    def transport(self):
        return {0: "none", 1: "nccl", 2: "peer"}[self.L.cfd_b200_session_transport(self.h)]

    def init_comm(self, nccl_lib, unique_id):
        check(self.L.cfd_b200_session_init_comm(self.h, nccl_lib.encode() if nccl_lib else None,
                                                unique_id), "session_init_comm")

    def profile(self, on=True):
        self.L.cfd_b200_session_profile(self.h, int(on))

    def reset_kernel_stats(self):
        self.L.cfd_b200_session_reset_kernel_stats(self.h)

    def kernel_stats(self):
        arr = (abi.KernelStat * 48)()
        n = self.L.cfd_b200_session_kernel_stats(self.h, arr, 48)
        return [dict(name=arr[i].name.decode(), launches=int(arr[i].launches), total_ms=arr[i].total_ms,
                     bytes_per_cell=arr[i].bytes_per_cell) for i in range(n)]

    def stream(self):
        return self.L.cfd_b200_session_stream(self.h)

    def sync(self):
        check(self.L.cfd_b200_session_sync(self.h), "session_sync")

    def close(self):
        if self.h:
            self.L.cfd_b200_session_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def pinned_field(nx, ny, nz):
    """HostField whose arrays live in CUDA pinned memory (cfd_b200_host_alloc)."""
    L = load()
    keep = []

    def alloc(n):
        p = L.cfd_b200_host_alloc(n * 8)
        if not p:
            raise MemoryError("cfd_b200_host_alloc(%d bytes) failed" % (n * 8))
        keep.append(p)
        buf = (C.c_double * n).from_address(p)
        return np.frombuffer(buf, dtype=np.float64)

    f = abi.HostField(nx, ny, nz, alloc=alloc)
    f._pinned = keep

    def release():
        for p in keep:
            L.cfd_b200_host_free(p)
        keep.clear()

    f.release = release
    return f
