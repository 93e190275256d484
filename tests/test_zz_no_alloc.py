"""The entry point that allocates nothing (csrc/no_alloc.cu):
cfd_b200_poisson_residual_linf_noalloc -- max |lap x - rhs| over the interior, reduced through a file-scope static
device array instead of a session's cudaMalloc'ed buffers -- against the EXISTING entry point for the same quantity:
the initial residual the Jacobi solver object reports (create_jacobi_gpu_solver -> solve -> stats.initial_residual),
itself bit-identical to the reference (tests/test_gpu_poisson.py).  Bitwise, as everywhere for the L-inf residual.
Consecutive calls share the scratch: a small grid right after a large one must not see the large one's partials.

Collected last on purpose (a failure here must not hide other tests); passes alone in a
fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest

import cases
from cfd_b200 import abi, api

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def reference_residual(g, x, rhs):
    """The existing entry point: one capped Jacobi solve; initial_residual = max |lap x0 - rhs|."""
    dx, dy, dz = g.spacing
    pp = abi.poisson_config("jacobi", max_iterations=1).params
    slv = api.PoissonGpuSolver(abi.POISSON_METHOD_JACOBI, g.nx, g.ny, g.nz, dx, dy, dz, pp)
    try:
        _, st = slv.solve(x.copy(), rhs, None)
    finally:
        slv.close()
    return st.initial_residual


def noalloc_residual(L, g, dx_, drhs, dout):
    dx, dy, dz = g.spacing
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = L.cfd_b200_poisson_residual_linf_noalloc(dx_.data_ptr(), drhs.data_ptr(), g.nx, g.ny, g.nz, dx, dy, dz,
                                                  dout.data_ptr(), st)
    assert rc == 0, api.last_error()


def test_noalloc_residual_equals_the_solver_objects_residual_on_consecutive_calls():
    L = api.load()
    shapes = [(258, 130, 66), (34, 30, 1), (66, 50, 1), (258, 130, 66), (40, 30, 20), (33, 33, 1), (21, 17, 9)]
    dout = torch.full((1,), -1.0, dtype=torch.float64, device="cuda")
    for q, shape in enumerate(shapes):
        g, x, rhs = cases.poisson_problem(*shape, seed=40 + q)
        x *= 10.0 ** (3 - q)  # residual magnitudes fall from call to call: stale partials would win the maximum
        want = reference_residual(g, x, rhs)
        dx_, drhs = torch.from_numpy(x.copy()).cuda(), torch.from_numpy(rhs.copy()).cuda()
        free0 = None
        for rep in range(3):  # calls in a row on the same input, then straight on to the next shape
            dout.fill_(-1.0)
            noalloc_residual(L, g, dx_, drhs, dout)
            torch.cuda.synchronize()
            got = float(dout.item())
            assert got == want, (shape, rep, got, want)
            if rep == 0:  # (the very first launch of a process may load modules: measured from the second call on)
                free0 = torch.cuda.mem_get_info()[0]
        assert torch.cuda.mem_get_info()[0] == free0, "the entry point must not allocate device memory"
        assert want > 0.0


def test_noalloc_residual_back_to_back_without_host_sync():
    """Stream order is all the ordering consecutive calls need: ten calls queued without a host synchronisation,
    alternating two problems and two output slots."""
    L = api.load()
    probs = []
    for q, shape in enumerate([(130, 66, 34), (50, 34, 1)]):
        g, x, rhs = cases.poisson_problem(*shape, seed=60 + q)
        probs.append((g, torch.from_numpy(x.copy()).cuda(), torch.from_numpy(rhs.copy()).cuda(),
                      reference_residual(g, x, rhs)))
    outs = torch.zeros(10, dtype=torch.float64, device="cuda")
    for q in range(10):
        g, dx_, drhs, _ = probs[q & 1]
        noalloc_residual(L, g, dx_, drhs, outs[q:q + 1])
    torch.cuda.synchronize()
    got = outs.cpu().numpy()
    for q in range(10):
        assert got[q] == probs[q & 1][3], (q, got[q], probs[q & 1][3])


def test_noalloc_residual_rejects_what_it_cannot_serve():
    L = api.load()
    assert L.cfd_b200_noalloc_max_cells() == 1 << 24
    d = torch.zeros(64, dtype=torch.float64, device="cuda")
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    f = L.cfd_b200_poisson_residual_linf_noalloc
    # beyond the static scratch: rejected BEFORE anything is touched (the small buffer is never read)
    assert f(d.data_ptr(), d.data_ptr(), 4097, 4096, 1, 0.1, 0.1, 0.0, d.data_ptr(), st) == abi.CFD_ERROR_LIMIT_EXCEEDED
    assert "maximum" in api.last_error()
    assert f(d.data_ptr(), d.data_ptr(), 1 << 40, 1 << 40, 1 << 40, 0.1, 0.1, 0.1, d.data_ptr(), st) == abi.CFD_ERROR_LIMIT_EXCEEDED
    assert f(None, d.data_ptr(), 8, 8, 1, 0.1, 0.1, 0.0, d.data_ptr(), st) == abi.CFD_ERROR_INVALID
    assert f(d.data_ptr(), d.data_ptr(), 8, 8, 2, 0.1, 0.1, 0.1, d.data_ptr(), st) == abi.CFD_ERROR_INVALID
    assert f(d.data_ptr(), d.data_ptr(), 2, 8, 1, 0.1, 0.1, 0.0, d.data_ptr(), st) == abi.CFD_ERROR_INVALID
    assert f(d.data_ptr(), d.data_ptr(), 8, 8, 1, 0.0, 0.1, 0.0, d.data_ptr(), st) == abi.CFD_ERROR_INVALID
    # exactly at the limit is served (4096 x 4096 x 1 = 2^24 cells)
    n = 4096
    big = torch.zeros(n * n, dtype=torch.float64, device="cuda")
    out = torch.full((1,), -1.0, dtype=torch.float64, device="cuda")
    assert f(big.data_ptr(), big.data_ptr(), n, n, 1, 1.0 / (n - 1), 1.0 / (n - 1), 0.0, out.data_ptr(), st) == 0
    torch.cuda.synchronize()
    assert float(out.item()) == 0.0
