"""oracle -- parity checkers for the projection-method hot path.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this package.  The product (cfd_b200/) never does.

  Oracle : ctypes binding of liboracle.so   (oracle.c, our CPU restatement)
  Ref    : ctypes binding of _ref/libcfd_ref*.so (the unmodified reference, compiled by
           oracle/Makefile from /root/reference; prebuilt copy travels to the GPU box)
"""
import ctypes as C
import os
import subprocess

from cfd_b200 import abi

HERE = os.path.dirname(os.path.abspath(__file__))


def build(quiet=True):
    """Compile liboracle.so and, when /root/reference is present, oracle/_ref/*.so."""
    out = subprocess.run(["make", "-C", HERE, "all"], capture_output=True, text=True)
    if out.returncode != 0:
        raise RuntimeError("oracle build failed:\n" + out.stdout[-2000:] + out.stderr[-2000:])
    if not quiet:
        print(out.stdout[-500:])


class Oracle:
    """CPU restatement (oracle.c)."""

    def __init__(self, variant=""):
        """variant "ld": Krylov dot products accumulated in long double (sensitivity probe)."""
        if variant == "ref":
            # timed CPU baseline: OpenMP stencil loops + the reference's own Poisson solver objects
            path = os.path.join(HERE, "_ref", "liboracle_ref.so")
            if not os.path.exists(path):
                raise FileNotFoundError(path + " (run `make -C oracle ref` where /root/reference exists)")
        else:
            path = os.path.join(HERE, "liboracle%s.so" % (("_" + variant) if variant else ""))
            if not os.path.exists(path):
                build()
        L = self.lib = C.CDLL(path)
        if variant == "ref":
            L.orc_set_ref_backend.argtypes = [C.c_int]
            L.orc_ref_backend_available.argtypes = [C.c_int, C.c_int]
            L.orc_ref_backend_available.restype = C.c_int
        L.orc_projection.restype = C.c_int
        L.orc_projection.argtypes = [C.POINTER(abi.FlowField), C.POINTER(abi.Grid),
                                     C.POINTER(abi.NSSolverParams), C.POINTER(abi.PoissonConfig),
                                     C.c_int, C.POINTER(abi.PoissonStats)]
        L.orc_rk.restype = C.c_int
        L.orc_rk.argtypes = [C.POINTER(abi.FlowField), C.POINTER(abi.Grid),
                             C.POINTER(abi.NSSolverParams), C.c_int, C.c_int]
        L.orc_explicit_euler.restype = C.c_int
        L.orc_explicit_euler.argtypes = [C.POINTER(abi.FlowField), C.POINTER(abi.Grid),
                                         C.POINTER(abi.NSSolverParams), C.c_int]
        L.orc_poisson_solve.restype = C.c_int
        L.orc_poisson_solve.argtypes = [C.c_int, C.c_size_t, C.c_size_t, C.c_size_t, C.c_double,
                                        C.c_double, C.c_double, C.POINTER(abi.PoissonParams),
                                        abi.c_double_p, abi.c_double_p, abi.c_double_p,
                                        C.POINTER(abi.PoissonStats)]
        L.orc_optimal_omega.restype = C.c_double
        L.orc_optimal_omega.argtypes = [C.c_size_t] * 3 + [C.c_double] * 3
        L.orc_thermal_bcs.restype = C.c_int
        L.orc_thermal_bcs.argtypes = [C.POINTER(abi.FlowField), C.POINTER(abi.NSSolverParams)]
        L.orc_apply_periodic.restype = None
        L.orc_apply_periodic.argtypes = [C.POINTER(abi.FlowField)]
        L.orc_neumann_scalar.restype = None
        L.orc_neumann_scalar.argtypes = [abi.c_double_p, C.c_size_t, C.c_size_t, C.c_size_t]
        L.orc_mg_levels.restype = C.c_int
        L.orc_mg_levels.argtypes = [C.c_size_t] * 3 + [C.POINTER(C.c_int)]
        L.orc_field_stats.restype = None
        L.orc_field_stats.argtypes = [C.POINTER(abi.FlowField)] + [abi.c_double_p] * 3

    def projection(self, field, grid, params, pcfg, nsteps=1):
        st = abi.PoissonStats()
        rc = self.lib.orc_projection(C.byref(field.c), C.byref(grid.c), C.byref(params),
                                     C.byref(pcfg), nsteps, C.byref(st))
        return rc, st

    def rk(self, field, grid, params, order, nsteps=1):
        return self.lib.orc_rk(C.byref(field.c), C.byref(grid.c), C.byref(params), order, nsteps)

    def explicit_euler(self, field, grid, params, nsteps=1):
        return self.lib.orc_explicit_euler(C.byref(field.c), C.byref(grid.c), C.byref(params), nsteps)

    def neumann(self, x):
        """Boundary cells = nearest interior cell, in place (orc_neumann_scalar); x is [nz, ny, nx]."""
        nz, ny, nx = x.shape
        self.lib.orc_neumann_scalar(abi._ptr(x), nx, ny, nz)

    def mg_levels(self, nx, ny, nz):
        """Extents (nx, ny, nz) of every level of the multigrid hierarchy (oracle_mg.inc)."""
        ext = (C.c_int * (3 * 32))()
        n = self.lib.orc_mg_levels(nx, ny, nz, ext)
        return [tuple(ext[3 * l:3 * l + 3]) for l in range(n)]

    def poisson(self, method, nx, ny, nz, dx, dy, dz, pparams, x, rhs, x_temp=None):
        st = abi.PoissonStats()
        xt = abi._ptr(x_temp) if x_temp is not None else None
        rc = self.lib.orc_poisson_solve(method, nx, ny, nz, dx, dy, dz, C.byref(pparams),
                                        abi._ptr(x), xt, abi._ptr(rhs), C.byref(st))
        return rc, st


class Ref:
    """The compiled reference library (oracle/_ref/libcfd_ref.so or _fast / _dropin)."""

    def __init__(self, variant=""):
        name = "libcfd_ref%s.so" % (("_" + variant) if variant else "")
        path = os.path.join(HERE, "_ref", name)
        if not os.path.exists(path):
            raise FileNotFoundError(path + " (run `make -C oracle ref` where /root/reference exists)")
        L = self.lib = C.CDLL(path, mode=C.RTLD_GLOBAL)
        vp = C.c_void_p
        L.cfd_registry_create.restype = vp
        L.cfd_registry_destroy.argtypes = [vp]
        L.cfd_registry_register_defaults.argtypes = [vp]
        L.cfd_solver_create.restype = vp
        L.cfd_solver_create.argtypes = [vp, C.c_char_p]
        L.solver_destroy.argtypes = [vp]
        L.solver_init.restype = C.c_int
        L.solver_init.argtypes = [vp, C.POINTER(abi.Grid), C.POINTER(abi.NSSolverParams)]
        for fn in (L.solver_step, L.solver_solve):
            fn.restype = C.c_int
            fn.argtypes = [vp, C.POINTER(abi.FlowField), C.POINTER(abi.Grid),
                           C.POINTER(abi.NSSolverParams), C.POINTER(abi.NSSolverStats)]
        L.poisson_solver_create.restype = vp
        L.poisson_solver_create.argtypes = [C.c_int, C.c_int]
        L.poisson_solver_init.restype = C.c_int
        L.poisson_solver_init.argtypes = [vp, C.c_size_t, C.c_size_t, C.c_size_t, C.c_double,
                                          C.c_double, C.c_double, C.POINTER(abi.PoissonParams)]
        L.poisson_solver_solve.restype = C.c_int
        L.poisson_solver_solve.argtypes = [vp, abi.c_double_p, abi.c_double_p, abi.c_double_p,
                                           C.POINTER(abi.PoissonStats)]
        L.poisson_solver_destroy.argtypes = [vp]
        L.ref_abi_table.restype = C.c_int
        L.ref_abi_table.argtypes = [C.POINTER(C.c_size_t), C.c_int]
        L.bc_apply_scalar_3d.restype = C.c_int
        L.energy_apply_thermal_bcs.restype = C.c_int
        L.energy_apply_thermal_bcs.argtypes = [C.POINTER(abi.FlowField), C.POINTER(abi.NSSolverParams)]
        L.apply_boundary_conditions.argtypes = [C.POINTER(abi.FlowField), C.POINTER(abi.Grid)]

    def run_solver(self, name, field, grid, params, nsteps=1, use_solve=False):
        """cfd_solver_create(name) then nsteps x solver_step (or one solver_solve(max_iter=nsteps))."""
        L = self.lib
        reg = L.cfd_registry_create()
        L.cfd_registry_register_defaults(reg)
        slv = L.cfd_solver_create(reg, name.encode())
        if not slv:
            L.cfd_registry_destroy(reg)
            return abi.CFD_ERROR_NOT_FOUND, None
        rc = L.solver_init(slv, C.byref(grid.c), C.byref(params))
        st = abi.NSSolverStats()
        if rc == 0:
            if use_solve:
                p2 = abi.NSSolverParams.from_buffer_copy(params)
                p2.max_iter = nsteps
                rc = L.solver_solve(slv, C.byref(field.c), C.byref(grid.c), C.byref(p2), C.byref(st))
            else:
                for _ in range(nsteps):
                    rc = L.solver_step(slv, C.byref(field.c), C.byref(grid.c), C.byref(params),
                                       C.byref(st))
                    if rc != 0:
                        break
        L.solver_destroy(slv)
        L.cfd_registry_destroy(reg)
        return rc, st

    def poisson(self, method, backend, nx, ny, nz, dx, dy, dz, pparams, x, rhs, x_temp=None):
        L = self.lib
        s = L.poisson_solver_create(method, backend)
        if not s:
            return abi.CFD_ERROR_UNSUPPORTED, None
        rc = L.poisson_solver_init(s, nx, ny, nz, dx, dy, dz,
                                   C.byref(pparams) if pparams is not None else None)
        st = abi.PoissonStats()
        if rc == 0:
            xt = abi._ptr(x_temp) if x_temp is not None else None
            rc = L.poisson_solver_solve(s, abi._ptr(x), xt, abi._ptr(rhs), C.byref(st))
        L.poisson_solver_destroy(s)
        return rc, st
