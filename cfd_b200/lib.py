"""ctypes binding of libcfd_b200.so (include/cfd_b200.h).  Fails loudly: there is no fallback."""
import ctypes as C
import os

from . import abi

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libcfd_b200.so")

_lib = None


class CfdB200Error(RuntimeError):
    def __init__(self, status, where, detail=""):
        self.status = status
        super().__init__("%s failed with cfd_status_t %d%s" % (where, status, (": " + detail) if detail else ""))


def load():
    """Load libcfd_b200.so and declare every prototype of include/cfd_b200.h."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError("libcfd_b200.so is missing (%s). Build it with `python -m cfd_b200.build`; "
                          "cfd_b200 has no CPU fallback." % LIB_PATH)
    L = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    vp, i32, dbl = C.c_void_p, C.c_int, C.c_double
    FF, GR, NP, NS = (C.POINTER(abi.FlowField), C.POINTER(abi.Grid), C.POINTER(abi.NSSolverParams),
                      C.POINTER(abi.NSSolverStats))
    PC, PS, GC = C.POINTER(abi.PoissonConfig), C.POINTER(abi.PoissonStats), C.POINTER(abi.GpuConfig)

    def proto(name, res, args):
        f = getattr(L, name)
        f.restype, f.argtypes = res, args

    # Part 2: drop-in entry points
    proto("gpu_is_available", i32, [])
    proto("gpu_config_default", abi.GpuConfig, [])
    for n in ("solve_projection_method_gpu", "solve_rk2_method_gpu", "solve_rk4_method_gpu",
              "solve_explicit_euler_method_gpu", "solve_navier_stokes_gpu"):
        proto(n, i32, [FF, GR, NP, GC])
    for n in ("create_jacobi_gpu_solver", "create_redblack_gpu_solver", "create_cg_gpu_solver",
              "create_bicgstab_gpu_solver", "cfd_b200_create_multigrid_gpu_solver"):
        proto(n, C.POINTER(abi.PoissonSolver), [])
    proto("gpu_should_use", i32, [GC, C.c_size_t, C.c_size_t, C.c_size_t, i32])
    proto("gpu_select_device", i32, [i32])
    proto("gpu_solver_create", vp, [C.c_size_t, C.c_size_t, C.c_size_t, GC])
    proto("gpu_solver_destroy", None, [vp])
    proto("gpu_solver_upload", i32, [vp, FF])
    proto("gpu_solver_download", i32, [vp, FF])
    proto("gpu_solver_step", i32, [vp, GR, NP, C.POINTER(abi.GpuSolverStats)])
    proto("gpu_solver_get_stats", abi.GpuSolverStats, [vp])
    proto("gpu_solver_reset_stats", None, [vp])
    for n in ("cfd_b200_create_projection_gpu_solver", "cfd_b200_create_explicit_euler_gpu_solver",
              "cfd_b200_create_rk2_gpu_solver",
              "cfd_b200_create_rk4_gpu_solver"):
        proto(n, C.POINTER(abi.NSSolver), [])
    # GPU boundary-condition API on device pointers (boundary_conditions_gpu.cuh)
    sz, dp, DV = C.c_size_t, C.c_void_p, C.POINTER(abi.DirichletValues)
    proto("bc_apply_neumann_gpu", None, [dp, sz, sz, vp])
    proto("bc_apply_scalar_gpu", None, [dp, sz, sz, i32, vp])
    proto("bc_apply_velocity_gpu", None, [dp, dp, sz, sz, i32, vp])
    proto("bc_apply_dirichlet_scalar_gpu", None, [dp, sz, sz, DV, vp])
    proto("bc_apply_dirichlet_velocity_gpu", None, [dp, dp, sz, sz, DV, DV, vp])
    proto("bc_apply_inlet_gpu", i32, [dp, dp, sz, sz, C.POINTER(abi.BcInletConfig), vp])
    proto("bc_apply_scalar_3d_gpu", None, [dp, sz, sz, sz, i32, vp])
    proto("bc_apply_velocity_3d_gpu", None, [dp, dp, dp, sz, sz, sz, i32, vp])
    # Part 3: extensions
    proto("cfd_b200_poisson_config_default", abi.PoissonConfig, [])
    proto("cfd_b200_set_poisson_config", i32, [PC])
    proto("cfd_b200_get_poisson_config", abi.PoissonConfig, [])
    proto("cfd_b200_last_error", C.c_char_p, [])
    proto("cfd_b200_version", C.c_char_p, [])
    proto("cfd_b200_kernel_launch_count", C.c_ulonglong, [])
    proto("cfd_b200_host_alloc", vp, [C.c_size_t])
    proto("cfd_b200_host_free", None, [vp])
    proto("cfd_b200_session_create", vp, [C.POINTER(abi.SessionDesc)])
    proto("cfd_b200_session_destroy", None, [vp])
    proto("cfd_b200_nccl_unique_id", i32, [C.c_char_p, C.c_char_p])
    proto("cfd_b200_session_init_comm", i32, [vp, C.c_char_p, C.c_char_p])
    proto("cfd_b200_session_upload", i32, [vp, FF])
    proto("cfd_b200_session_download", i32, [vp, FF])
    proto("cfd_b200_session_upload_mask", i32, [vp, FF, i32])
    proto("cfd_b200_session_download_mask", i32, [vp, FF, i32])
    proto("cfd_b200_session_init_synthetic", i32, [vp, C.c_uint64, C.c_size_t, C.c_size_t])
    proto("cfd_b200_session_set_coords", i32, [vp, C.POINTER(C.c_double), C.POINTER(C.c_double)])
    proto("cfd_b200_session_projection", i32, [vp, NP, PC, i32, NS])
    proto("cfd_b200_session_rk", i32, [vp, i32, NP, i32, NS])
    proto("cfd_b200_session_poisson", i32, [vp, PC, PS])
    proto("cfd_b200_session_upload_poisson", i32, [vp, abi.c_double_p, abi.c_double_p])
    proto("cfd_b200_session_download_poisson", i32, [vp, abi.c_double_p])
    proto("cfd_b200_session_last_poisson_stats", abi.PoissonStats, [vp])
    proto("cfd_b200_session_profile", None, [vp, i32])
    proto("cfd_b200_session_kernel_stats", i32, [vp, C.POINTER(abi.KernelStat), i32])
    proto("cfd_b200_session_reset_kernel_stats", None, [vp])
    proto("cfd_b200_session_stream", vp, [vp])
    proto("cfd_b200_session_sync", i32, [vp])
    proto("cfd_b200_abi_table", i32, [C.POINTER(C.c_size_t), i32])
    proto("cfd_b200_session_peer_export", i32, [vp, C.c_char_p])
    proto("cfd_b200_session_peer_attach", i32, [vp, C.c_char_p])
    proto("cfd_b200_session_transport", i32, [vp])
    proto("cfd_b200_noalloc_max_cells", C.c_size_t, [])
    proto("cfd_b200_poisson_residual_linf_noalloc", i32, [vp, vp, C.c_size_t, C.c_size_t, C.c_size_t, dbl, dbl, dbl,
                                                          vp, vp])
    proto("cfd_b200_set_arithmetic", i32, [i32])
    proto("cfd_b200_session_arithmetic", i32, [vp])
    proto("cfd_b200_set_device_count", i32, [i32])
    proto("cfd_b200_get_device_count", i32, [])
    proto("cfd_b200_session_set_slab", i32, [vp, C.c_size_t, C.c_size_t])
    proto("cfd_b200_release_cached_sessions", None, [])
    proto("cfd_b200_solver_set_host_sync", i32, [C.POINTER(abi.NSSolver), i32])
    proto("cfd_b200_solver_sync_host", i32, [C.POINTER(abi.NSSolver), FF])
    _lib = L
    return L


# Every symbol include/cfd_b200.h declares (checked by tests/test_abi.py against the .so).
DECLARED_SYMBOLS = [
    "gpu_is_available", "gpu_config_default", "solve_projection_method_gpu", "solve_rk2_method_gpu",
    "solve_rk4_method_gpu", "solve_explicit_euler_method_gpu", "create_jacobi_gpu_solver",
    "create_redblack_gpu_solver", "create_cg_gpu_solver", "create_bicgstab_gpu_solver",
    "gpu_get_device_info", "gpu_select_device", "gpu_should_use", "solve_navier_stokes_gpu",
    "gpu_solver_create", "gpu_solver_destroy", "gpu_solver_upload", "gpu_solver_download",
    "gpu_solver_step", "gpu_solver_get_stats", "gpu_solver_reset_stats",
    "bc_apply_neumann_gpu", "bc_apply_scalar_gpu", "bc_apply_velocity_gpu", "bc_apply_dirichlet_scalar_gpu",
    "bc_apply_dirichlet_velocity_gpu", "bc_apply_inlet_gpu", "bc_apply_scalar_3d_gpu", "bc_apply_velocity_3d_gpu",
    "cfd_b200_create_projection_gpu_solver", "cfd_b200_create_explicit_euler_gpu_solver",
    "cfd_b200_create_rk2_gpu_solver",
    "cfd_b200_create_rk4_gpu_solver", "cfd_b200_create_multigrid_gpu_solver", "cfd_b200_poisson_config_default", "cfd_b200_set_poisson_config",
    "cfd_b200_get_poisson_config", "cfd_b200_last_error", "cfd_b200_version",
    "cfd_b200_kernel_launch_count", "cfd_b200_host_alloc", "cfd_b200_host_free",
    "cfd_b200_session_create", "cfd_b200_session_destroy", "cfd_b200_nccl_unique_id",
    "cfd_b200_session_init_comm", "cfd_b200_session_upload", "cfd_b200_session_download",
    "cfd_b200_session_upload_mask", "cfd_b200_session_download_mask",
    "cfd_b200_session_init_synthetic", "cfd_b200_session_set_coords", "cfd_b200_session_projection", "cfd_b200_session_rk",
    "cfd_b200_session_poisson", "cfd_b200_session_upload_poisson", "cfd_b200_session_download_poisson",
    "cfd_b200_session_last_poisson_stats", "cfd_b200_session_profile", "cfd_b200_session_kernel_stats",
    "cfd_b200_session_reset_kernel_stats", "cfd_b200_session_stream", "cfd_b200_session_sync",
    "cfd_b200_set_device_count", "cfd_b200_get_device_count", "cfd_b200_session_set_slab",
    "cfd_b200_release_cached_sessions", "cfd_b200_solver_set_host_sync", "cfd_b200_solver_sync_host",
    "cfd_b200_noalloc_max_cells", "cfd_b200_poisson_residual_linf_noalloc",
    "cfd_b200_set_arithmetic", "cfd_b200_session_arithmetic", "cfd_b200_session_peer_export", "cfd_b200_session_peer_attach", "cfd_b200_session_transport",
]


def last_error():
    return load().cfd_b200_last_error().decode(errors="replace")


def check(rc, where):
    if rc != abi.CFD_SUCCESS:
        raise CfdB200Error(rc, where, last_error())
