"""Generate tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref/libcfd_ref.so, built
from /root/reference by oracle/Makefile).  Run in the build container only:

    python tests/golden/make_golden.py

Each file holds the inputs and the reference outputs of one small case so that the oracle and
the CUDA path can be checked on machines where /root/reference does not exist.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import cases  # noqa: E402
import oracle  # noqa: E402
from cfd_b200 import abi  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
R = oracle.Ref()


def save_step(tag, solver, g, f, p, nsteps):
    inp = {("in_" + k): v.copy() for k, v in f.arrays().items()}
    rc, _ = R.run_solver(solver, f, g, p, nsteps)
    assert rc == 0, (tag, rc)
    out = {("out_" + k): v.copy() for k, v in f.arrays().items()}
    np.savez_compressed(os.path.join(OUT, tag + ".npz"), nsteps=nsteps, solver=solver, **inp, **out)
    print("wrote", tag)


def main():
    for solver in ("projection", "rk2", "rk4", "explicit_euler"):
        g, f, p = cases.golden16()
        save_step("golden16_" + solver, solver, g, f, p, 1)
    g, f, p = cases.taylor_green(12, 10, dt=1e-3)
    save_step("tg3d_rk4", "rk4", g, f, p, 3)
    g, f, p = cases.taylor_green(12, 10, dt=1e-3)
    save_step("tg3d_explicit_euler", "explicit_euler", g, f, p, 3)
    g, f, p = cases.natural_convection(17, 1, dt=5e-5)
    save_step("natconv2d_explicit_euler", "explicit_euler", g, f, p, 3)
    g, f, p = cases.taylor_green(12, 10, dt=1e-3)
    save_step("tg3d_projection", "projection", g, f, p, 2)
    g, f, p = cases.natural_convection(17, 1, dt=2e-3)
    save_step("natconv2d_projection", "projection", g, f, p, 3)
    g, f, p = cases.golden16()
    p.source_amplitude_u, p.source_amplitude_v = 0.1, 0.05
    save_step("source16_projection", "projection", g, f, p, 2)
    # stand-alone Poisson solves, scalar backend, every method
    meth = {"jacobi": abi.POISSON_METHOD_JACOBI, "rbsor": abi.POISSON_METHOD_REDBLACK_SOR,
            "cg": abi.POISSON_METHOD_CG, "pcg": abi.POISSON_METHOD_CG, "bicgstab": abi.POISSON_METHOD_BICGSTAB}
    for shape in ((17, 15, 1), (12, 10, 9)):
        for name, m in meth.items():
            g, x, rhs = cases.poisson_problem(*shape, seed=3)
            pp = abi.poisson_params_default()
            if name == "jacobi":
                pp.max_iterations = 2000
            if name == "pcg":
                pp.preconditioner = abi.POISSON_PRECOND_JACOBI
            x0 = x.copy()
            xt = np.zeros_like(x)
            dx, dy, dz = g.spacing
            rc, st = R.poisson(m, abi.POISSON_BACKEND_SCALAR, *shape, dx, dy, dz, pp, x, rhs, xt)
            tag = "poisson_%s_%dx%dx%d" % ((name,) + shape)
            np.savez_compressed(os.path.join(OUT, tag + ".npz"), x0=x0, rhs=rhs, x=x, rc=rc,
                                iterations=st.iterations, status=st.status,
                                initial_residual=st.initial_residual, final_residual=st.final_residual)
            print("wrote", tag, rc, st.iterations)


if __name__ == "__main__":
    main()
