"""ctypes mirror of include/cfd_b200.h (Part 1: the reference's boundary types).

Each Structure cites the reference header it mirrors; tests/test_abi.py proves the layouts
agree with both the product library and the compiled reference.
"""
import ctypes as C

import numpy as np

c_double_p = C.POINTER(C.c_double)

# cfd/core/cfd_status.h:13-24
CFD_SUCCESS = 0
CFD_ERROR = -1
CFD_ERROR_NOMEM = -2
CFD_ERROR_INVALID = -3
CFD_ERROR_IO = -4
CFD_ERROR_UNSUPPORTED = -5
CFD_ERROR_DIVERGED = -6
CFD_ERROR_MAX_ITER = -7
CFD_ERROR_LIMIT_EXCEEDED = -8
CFD_ERROR_NOT_FOUND = -9

# cfd/boundary/boundary_conditions.h:19-27
BC_TYPE_PERIODIC, BC_TYPE_NEUMANN, BC_TYPE_DIRICHLET, BC_TYPE_NOSLIP = 0, 1, 2, 3
BC_TYPE_INLET, BC_TYPE_OUTLET, BC_TYPE_SYMMETRY = 4, 5, 6

# cfd/solvers/poisson_solver.h:53-95
POISSON_METHOD_JACOBI, POISSON_METHOD_GAUSS_SEIDEL, POISSON_METHOD_SOR = 0, 1, 2
POISSON_METHOD_REDBLACK_SOR, POISSON_METHOD_CG, POISSON_METHOD_BICGSTAB = 3, 4, 5
POISSON_METHOD_MULTIGRID = 6
POISSON_BACKEND_AUTO, POISSON_BACKEND_SCALAR, POISSON_BACKEND_OMP = 0, 1, 2
POISSON_BACKEND_SIMD, POISSON_BACKEND_GPU = 3, 4
POISSON_CONVERGED, POISSON_MAX_ITER, POISSON_DIVERGED, POISSON_STAGNATED, POISSON_ERROR = 0, 1, 2, 3, -1
POISSON_PRECOND_NONE, POISSON_PRECOND_JACOBI = 0, 1
CFD_B200_PRECOND_MULTIGRID = 100  # extension: CG preconditioned by a multigrid cycle (csrc/poisson_mgpc.cu)

CFD_B200_NONCONV_ERROR, CFD_B200_NONCONV_CONTINUE = 0, 1


class Grid(C.Structure):
    """cfd/core/grid.h:18-40"""
    _fields_ = [
        ("x", c_double_p), ("y", c_double_p), ("dx", c_double_p), ("dy", c_double_p),
        ("nx", C.c_size_t), ("ny", C.c_size_t),
        ("xmin", C.c_double), ("xmax", C.c_double), ("ymin", C.c_double), ("ymax", C.c_double),
        ("z", c_double_p), ("dz", c_double_p), ("nz", C.c_size_t),
        ("zmin", C.c_double), ("zmax", C.c_double),
        ("stride_z", C.c_size_t), ("inv_dz2", C.c_double),
        ("k_start", C.c_size_t), ("k_end", C.c_size_t),
    ]


class FlowField(C.Structure):
    """cfd/solvers/navier_stokes_solver.h:54-64"""
    _fields_ = [
        ("u", c_double_p), ("v", c_double_p), ("w", c_double_p), ("p", c_double_p),
        ("rho", c_double_p), ("T", c_double_p),
        ("nx", C.c_size_t), ("ny", C.c_size_t), ("nz", C.c_size_t),
    ]


class DirichletValues(C.Structure):
    """cfd/boundary/boundary_conditions.h:50-57"""
    _fields_ = [(n, C.c_double) for n in ("left", "right", "top", "bottom", "front", "back")]


# boundary_conditions.h:64-68, 85-89, 96-103
BC_INLET_PROFILE_UNIFORM, BC_INLET_PROFILE_PARABOLIC, BC_INLET_PROFILE_CUSTOM = 0, 1, 2
BC_INLET_SPEC_VELOCITY, BC_INLET_SPEC_MAGNITUDE_DIR, BC_INLET_SPEC_MASS_FLOW = 0, 1, 2
BC_EDGE_LEFT, BC_EDGE_RIGHT, BC_EDGE_BOTTOM, BC_EDGE_TOP, BC_EDGE_FRONT, BC_EDGE_BACK = 1, 2, 4, 8, 16, 32

INLET_PROFILE_FN = C.CFUNCTYPE(None, C.c_double, C.POINTER(C.c_double), C.POINTER(C.c_double), C.c_void_p)


class _TimeParams(C.Union):
    _fields_ = [("sinusoidal", C.c_double * 4), ("ramp", C.c_double * 4), ("step", C.c_double * 3)]


class BcTimeConfig(C.Structure):
    """boundary_conditions.h:183-198"""
    _fields_ = [("profile", C.c_int), ("params", _TimeParams), ("custom_fn", C.c_void_p),
                ("custom_user_data", C.c_void_p)]


class _InletSpec(C.Union):
    _fields_ = [("velocity", C.c_double * 2), ("magnitude_dir", C.c_double * 2), ("mass_flow", C.c_double * 3)]


class BcInletConfig(C.Structure):
    """boundary_conditions.h:233-267"""
    _fields_ = [("edge", C.c_int), ("profile", C.c_int), ("spec_type", C.c_int), ("spec", _InletSpec),
                ("custom_profile", INLET_PROFILE_FN), ("custom_profile_user_data", C.c_void_p),
                ("time_config", BcTimeConfig), ("custom_profile_time", C.c_void_p),
                ("custom_profile_time_user_data", C.c_void_p)]


class ThermalBC(C.Structure):
    """navier_stokes_solver.h:108-116"""
    _fields_ = [(n, C.c_int) for n in ("left", "right", "bottom", "top", "front", "back")] + [
        ("dirichlet_values", DirichletValues)]


class NSSolverParams(C.Structure):
    """navier_stokes_solver.h:121-158"""
    _fields_ = [
        ("dt", C.c_double), ("cfl", C.c_double), ("gamma", C.c_double), ("mu", C.c_double),
        ("k", C.c_double), ("max_iter", C.c_int), ("tolerance", C.c_double),
        ("source_amplitude_u", C.c_double), ("source_amplitude_v", C.c_double),
        ("source_decay_rate", C.c_double), ("pressure_coupling", C.c_double),
        ("source_func", C.c_void_p), ("source_context", C.c_void_p),
        ("alpha", C.c_double), ("beta", C.c_double), ("T_ref", C.c_double),
        ("gravity", C.c_double * 3),
        ("heat_source_func", C.c_void_p), ("heat_source_context", C.c_void_p),
        ("thermal_bc", ThermalBC),
    ]


class NSSolverStats(C.Structure):
    """navier_stokes_solver.h:198-207"""
    _fields_ = [
        ("iterations", C.c_int), ("residual", C.c_double), ("max_velocity", C.c_double),
        ("max_pressure", C.c_double), ("max_temperature", C.c_double), ("cfl_number", C.c_double),
        ("elapsed_time_ms", C.c_double), ("status", C.c_int),
    ]


class GpuConfig(C.Structure):
    """cfd/core/gpu_device.h:32-53"""
    _fields_ = [
        ("enable_gpu", C.c_int), ("min_grid_size", C.c_size_t), ("min_steps", C.c_int),
        ("block_size_x", C.c_int), ("block_size_y", C.c_int), ("poisson_max_iter", C.c_int),
        ("poisson_tolerance", C.c_double), ("persistent_memory", C.c_int),
        ("async_transfers", C.c_int), ("sync_after_kernel", C.c_int), ("verbose", C.c_int),
    ]


class GpuSolverStats(C.Structure):
    """gpu_device.h:74-82"""
    _fields_ = [
        ("kernel_time_ms", C.c_double), ("transfer_time_ms", C.c_double),
        ("poisson_time_ms", C.c_double), ("poisson_iterations", C.c_int),
        ("poisson_residual", C.c_double), ("memory_allocated", C.c_size_t),
        ("kernels_launched", C.c_int),
    ]


class PoissonParams(C.Structure):
    """poisson_solver.h:100-108"""
    _fields_ = [
        ("tolerance", C.c_double), ("absolute_tolerance", C.c_double), ("max_iterations", C.c_int),
        ("omega", C.c_double), ("check_interval", C.c_int), ("verbose", C.c_bool),
        ("preconditioner", C.c_int),
    ]


class PoissonStats(C.Structure):
    """poisson_solver.h:113-119"""
    _fields_ = [
        ("status", C.c_int), ("iterations", C.c_int), ("initial_residual", C.c_double),
        ("final_residual", C.c_double), ("elapsed_time_ms", C.c_double),
    ]


class PoissonSolver(C.Structure):
    """poisson_solver.h:206-233 (function pointers kept opaque)"""
    _fields_ = [
        ("name", C.c_char_p), ("description", C.c_char_p), ("method", C.c_int), ("backend", C.c_int),
        ("nx", C.c_size_t), ("ny", C.c_size_t), ("nz", C.c_size_t),
        ("dx", C.c_double), ("dy", C.c_double), ("dz", C.c_double),
        ("params", PoissonParams), ("context", C.c_void_p),
        ("init", C.c_void_p), ("destroy", C.c_void_p), ("solve", C.c_void_p),
        ("iterate", C.c_void_p), ("apply_bc", C.c_void_p),
    ]


class NSSolver(C.Structure):
    """navier_stokes_solver.h:254-277 (function pointers kept opaque)"""
    _fields_ = [
        ("name", C.c_char_p), ("description", C.c_char_p), ("version", C.c_char_p),
        ("capabilities", C.c_int), ("backend", C.c_int), ("context", C.c_void_p),
        ("init", C.c_void_p), ("destroy", C.c_void_p), ("step", C.c_void_p), ("solve", C.c_void_p),
        ("apply_boundary", C.c_void_p), ("compute_dt", C.c_void_p), ("get_name", C.c_void_p),
        ("get_description", C.c_void_p), ("get_capabilities", C.c_void_p),
    ]


class PoissonConfig(C.Structure):
    """include/cfd_b200.h: cfd_b200_poisson_config_t"""
    _fields_ = [("method", C.c_int), ("params", PoissonParams), ("on_nonconvergence", C.c_int)]


class SessionDesc(C.Structure):
    """include/cfd_b200.h: cfd_b200_session_desc_t"""
    _fields_ = [
        ("nx", C.c_size_t), ("ny", C.c_size_t), ("nz", C.c_size_t),
        ("dx", C.c_double), ("dy", C.c_double), ("dz", C.c_double),
        ("xmin", C.c_double), ("ymin", C.c_double), ("zmin", C.c_double),
        ("rank", C.c_int), ("nranks", C.c_int), ("device", C.c_int),
    ]


class KernelStat(C.Structure):
    """include/cfd_b200.h: cfd_b200_kernel_stat_t"""
    _fields_ = [("name", C.c_char_p), ("launches", C.c_ulonglong), ("total_ms", C.c_double),
                ("bytes_per_cell", C.c_double)]


def _ptr(a):
    return a.ctypes.data_as(c_double_p)


def aligned_zeros(n, align=64):
    """numpy float64 array whose data pointer is `align`-byte aligned (the reference uses
    32-byte aligned calloc, cpu/solver_explicit_euler.c:79-110)."""
    raw = np.zeros(n * 8 + align, dtype=np.uint8)
    off = (-raw.ctypes.data) % align
    return raw[off:off + n * 8].view(np.float64)


class HostGrid:
    """Uniform grid with the semantics of grid_create + grid_initialize_uniform
    (lib/src/core/grid.c:9-120): dx = (xmax-xmin)/(nx-1), x[i] = xmin + i*dx."""

    def __init__(self, nx, ny, nz=1, xmin=0.0, xmax=1.0, ymin=0.0, ymax=1.0, zmin=0.0, zmax=0.0):
        if nz > 1 and zmax <= zmin:
            raise ValueError("zmax must exceed zmin when nz > 1")
        self.nx, self.ny, self.nz = nx, ny, nz
        dx = (xmax - xmin) / (nx - 1)
        dy = (ymax - ymin) / (ny - 1)
        self.x = xmin + np.arange(nx, dtype=np.float64) * dx
        self.y = ymin + np.arange(ny, dtype=np.float64) * dy
        self.dx = np.full(nx - 1, dx)
        self.dy = np.full(ny - 1, dy)
        g = Grid()
        g.x, g.y, g.dx, g.dy = _ptr(self.x), _ptr(self.y), _ptr(self.dx), _ptr(self.dy)
        g.nx, g.ny, g.nz = nx, ny, nz
        g.xmin, g.xmax, g.ymin, g.ymax = xmin, xmax, ymin, ymax
        if nz > 1:
            dz = (zmax - zmin) / (nz - 1)
            self.z = zmin + np.arange(nz, dtype=np.float64) * dz
            self.dz = np.full(nz - 1, dz)
            g.z, g.dz = _ptr(self.z), _ptr(self.dz)
            g.zmin, g.zmax = zmin, zmax
            g.stride_z, g.inv_dz2, g.k_start, g.k_end = nx * ny, 1.0 / (dz * dz), 1, nz - 1
        else:
            self.z = self.dz = None
            g.stride_z, g.inv_dz2, g.k_start, g.k_end = 0, 0.0, 0, 1
        self.c = g

    @property
    def spacing(self):
        return float(self.dx[0]), float(self.dy[0]), float(self.dz[0]) if self.nz > 1 else 0.0


class HostField:
    """Host flow_field: six nx*ny*nz float64 arrays, x fastest (cfd/core/indexing.h:17-20).
    Arrays are exposed shaped (nz, ny, nx)."""

    NAMES = ("u", "v", "w", "p", "rho", "T")

    def __init__(self, nx, ny, nz=1, alloc=None):
        self.nx, self.ny, self.nz = nx, ny, nz
        n = nx * ny * nz
        alloc = alloc or aligned_zeros
        self.c = FlowField()
        self.c.nx, self.c.ny, self.c.nz = nx, ny, nz
        for name in self.NAMES:
            a = alloc(n)
            setattr(self, name, a.reshape(nz, ny, nx))
            setattr(self.c, name, _ptr(a))
        self.rho[...] = 1.0

    def copy(self):
        o = HostField(self.nx, self.ny, self.nz)
        for name in self.NAMES:
            getattr(o, name)[...] = getattr(self, name)
        return o

    def arrays(self):
        return {n: getattr(self, n) for n in self.NAMES}


def ns_params_default():
    """ns_solver_params_default() -- cpu/solver_explicit_euler.c:58-78."""
    p = NSSolverParams()
    p.dt, p.cfl, p.gamma, p.mu, p.k = 0.001, 0.2, 1.4, 0.01, 0.0242
    p.max_iter, p.tolerance = 100, 1e-6
    p.source_amplitude_u, p.source_amplitude_v = 0.1, 0.05
    p.source_decay_rate, p.pressure_coupling = 0.1, 0.1
    return p


def poisson_params_default():
    """poisson_solver_params_default() -- linear_solver.c:37-47."""
    p = PoissonParams()
    p.tolerance, p.absolute_tolerance, p.max_iterations = 1e-6, 1e-10, 5000
    p.omega, p.check_interval, p.verbose, p.preconditioner = 0.0, 1, False, POISSON_PRECOND_NONE
    return p


def poisson_config(method="cg", **kw):
    """Build a cfd_b200_poisson_config_t. method in {jacobi, rbsor, cg, pcg, mgpcg, bicgstab, multigrid}."""
    names = {"jacobi": POISSON_METHOD_JACOBI, "rbsor": POISSON_METHOD_REDBLACK_SOR,
             "cg": POISSON_METHOD_CG, "pcg": POISSON_METHOD_CG, "mgpcg": POISSON_METHOD_CG,
             "bicgstab": POISSON_METHOD_BICGSTAB, "multigrid": POISSON_METHOD_MULTIGRID}
    cfg = PoissonConfig()
    cfg.method = names[method]
    cfg.params = poisson_params_default()
    if method == "jacobi":
        cfg.params.max_iterations = 2000  # linear_solver.c:275-278
    if method == "multigrid":
        cfg.params.max_iterations = 100  # V-cycles
    if method == "pcg":
        cfg.params.preconditioner = POISSON_PRECOND_JACOBI
    if method == "mgpcg":
        cfg.params.preconditioner = CFD_B200_PRECOND_MULTIGRID
    cfg.on_nonconvergence = CFD_B200_NONCONV_ERROR
    for k, v in kw.items():
        if k == "on_nonconvergence":
            cfg.on_nonconvergence = v
        else:
            setattr(cfg.params, k, v)
    return cfg
