"""The C-ABI library loads, exports every symbol include/cfd_b200.h declares, and its view of
the boundary types agrees with the compiled reference and with the ctypes mirror.  No compute."""
import ctypes as C
import os
import re

import pytest

from cfd_b200 import abi, lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _table(fn):
    buf = (C.c_size_t * 256)()
    n = fn(buf, 256)
    return list(buf[:n])


def test_library_loads_and_exports_every_declared_symbol():
    L = lib.load()
    for name in lib.DECLARED_SYMBOLS:
        assert hasattr(L, name), "libcfd_b200.so does not export %s" % name
    assert b"sm_100a" in L.cfd_b200_version()


def test_header_and_binding_list_agree():
    text = open(os.path.join(ROOT, "include", "cfd_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)  # drop comments
    text = text[text.index("Part 2") if "Part 2" in text else 0:]
    declared = set(re.findall(r"CFD_B200_API[^;(]*?\b([A-Za-z_][A-Za-z0-9_]*)\s*\(", text))
    declared.discard("__attribute__")  # the macro definition itself
    assert declared == set(lib.DECLARED_SYMBOLS), declared ^ set(lib.DECLARED_SYMBOLS)


def test_abi_matches_compiled_reference(ref):
    ours = _table(lib.load().cfd_b200_abi_table)
    theirs = _table(ref.lib.ref_abi_table)
    assert ours == theirs


def test_ctypes_mirror_sizes():
    t = _table(lib.load().cfd_b200_abi_table)
    assert t[0] == C.sizeof(abi.FlowField)
    sizes = {C.sizeof(abi.Grid), C.sizeof(abi.NSSolverParams), C.sizeof(abi.ThermalBC),
             C.sizeof(abi.DirichletValues), C.sizeof(abi.NSSolverStats), C.sizeof(abi.NSSolver),
             C.sizeof(abi.GpuConfig), C.sizeof(abi.GpuSolverStats), C.sizeof(abi.PoissonParams),
             C.sizeof(abi.PoissonStats), C.sizeof(abi.PoissonSolver), C.sizeof(abi.BcInletConfig),
             C.sizeof(abi.BcTimeConfig)}
    assert sizes <= set(t)
    assert abi.NSSolverParams.thermal_bc.offset in t and abi.NSSolverParams.alpha.offset in t
    assert abi.PoissonSolver.params.offset in t and abi.PoissonSolver.solve.offset in t
    assert abi.BcInletConfig.time_config.offset in t and abi.BcInletConfig.custom_profile.offset in t


def test_poisson_config_roundtrip_without_gpu():
    L = lib.load()
    d = L.cfd_b200_poisson_config_default()
    assert d.method == abi.POISSON_METHOD_CG and d.params.max_iterations == 5000
    assert d.params.tolerance == 1e-6 and d.params.absolute_tolerance == 1e-10
    cfg = abi.poisson_config("bicgstab", max_iterations=50)
    assert L.cfd_b200_set_poisson_config(C.byref(cfg)) == 0
    got = L.cfd_b200_get_poisson_config()
    assert got.method == abi.POISSON_METHOD_BICGSTAB and got.params.max_iterations == 50
    mg = abi.poisson_config("multigrid")  # the reference's placeholder method: implemented here
    assert mg.method == abi.POISSON_METHOD_MULTIGRID and mg.params.max_iterations == 100
    assert L.cfd_b200_set_poisson_config(C.byref(mg)) == 0
    assert L.cfd_b200_get_poisson_config().method == abi.POISSON_METHOD_MULTIGRID
    bad = abi.poisson_config("cg")
    for method in (abi.POISSON_METHOD_GAUSS_SEIDEL, abi.POISSON_METHOD_SOR):  # no GPU backend in the reference either
        bad.method = method
        assert L.cfd_b200_set_poisson_config(C.byref(bad)) == abi.CFD_ERROR_UNSUPPORTED
    L.cfd_b200_set_poisson_config(C.byref(d))


def test_fails_loudly_without_cuda():
    """No CPU fallback: on a box without a GPU every compute entry point refuses."""
    L = lib.load()
    if L.gpu_is_available():
        pytest.skip("a GPU is present")
    from cfd_b200 import api
    import cases
    g, f, p = cases.golden16()
    p.max_iter = 1
    assert api.solve_method_gpu("projection", f, g, p) == abi.CFD_ERROR
    with pytest.raises(lib.CfdB200Error):
        api.Session(16, 16, 1, 0.1, 0.1)
    with pytest.raises(lib.CfdB200Error):
        api.GpuSolver("projection_gpu")


def test_dropin_library_registers_the_gpu_names(ref):
    """oracle/_ref/libcfd_ref_dropin.so = unmodified reference core/api + libcfd_b200.so in place of
    libcfd_cuda.a.  Without a GPU the reference's factory must get NULL from our side
    (solver_registry.c:1157-1160) -- never a silent CPU path."""
    import ctypes as C
    import oracle
    try:
        d = oracle.Ref("dropin")
    except (FileNotFoundError, OSError) as e:
        pytest.skip(str(e))
    L = d.lib
    reg = L.cfd_registry_create()
    L.cfd_registry_register_defaults(reg)
    L.cfd_registry_has.restype = C.c_int
    L.cfd_registry_has.argtypes = [C.c_void_p, C.c_char_p]
    for name in (b"projection_gpu", b"rk2_gpu", b"rk4_gpu"):
        assert L.cfd_registry_has(reg, name) == 1
    if not lib.load().gpu_is_available():
        assert not L.cfd_solver_create(reg, b"projection_gpu")
    L.cfd_registry_destroy(reg)


def test_every_kernel_kind_has_a_name_and_a_byte_figure():
    """kernels.h's KernelKind, session.cu's name and bytes-per-cell tables must stay the same length: a missing
    initialiser would be a NULL name handed to the caller of cfd_b200_session_kernel_stats."""
    import os
    import re
    root = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "cfd_b200", "csrc")
    hdr = open(os.path.join(root, "kernels.h")).read()
    body = re.search(r"enum KernelKind \{(.*?)\};", hdr, re.S).group(1)
    kinds = [k.strip().split("=")[0].strip() for k in body.replace("\n", " ").split(",") if k.strip()]
    assert kinds[-1] == "KK_COUNT"
    src = open(os.path.join(root, "session.cu")).read()
    names = re.search(r"kKernelNames\[KK_COUNT\] = \{(.*?)\};", src, re.S).group(1)
    n_names = len(re.findall(r'"[^"]+"', names))
    table = re.search(r"kKernelBytesPerCell\[KK_COUNT\] = \{(.*?)\};", src, re.S).group(1)
    n_bytes = len(re.findall(r"\d+(?:\.\d+)?", re.sub(r"//.*", "", table)))
    assert n_names == n_bytes == len(kinds) - 1, (n_names, n_bytes, len(kinds) - 1)
