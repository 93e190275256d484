#!/usr/bin/env python
"""bench.py -- projection-step MLUPS (FP64) and Poisson-sweep HBM GB/s against the roofline.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--box 1024] [--iters 50]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

Workload (BASELINE.json configs[4], the configuration the metric is quoted on): synthetic 3-D box,
n^3 cells PER GPU (n = 1024 when it fits), slab-decomposed along z, one projection step =
predictor + divergence + pressure Poisson (BiCGSTAB, exactly `--iters` iterations: cap reached =>
continue, SURVEY 8d C5 "fixed-50-iteration timing mode") + corrector.  A fixed iteration count
makes MLUPS size-normalised, so the CPU arms can run the same step on a smaller box.

One JSON line on stdout (rank 0).  `value` is device-resident throughput (CUDA events on the
session stream, max over ranks); `e2e` goes through the drop-in C-ABI entry point
solve_projection_method_gpu() with pinned HOST flow_field arrays (H2D + D2H inside the timed
region); `roofline` is the dominant kernel, timed with per-launch CUDA events inside the same
timed region; `cpu_baseline` is the reference CPU code on this host (see oracle/Makefile).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "projection_step_mlups_fp64"
UNIT = "MLUPS"
DT = 1e-4
NU = 0.01


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--box", dest="n", type=int, default=0, help="cells per axis per GPU (0 = 1024 if it fits, else smaller)")
    ap.add_argument("--iters", type=int, default=50, help="BiCGSTAB iterations per step (fixed)")
    ap.add_argument("--method", default="bicgstab", choices=["jacobi", "rbsor", "cg", "pcg", "bicgstab"])
    ap.add_argument("--cpu-n", type=int, default=0, help="box edge of the CPU sample (0 = auto)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--budget", type=float, default=150.0, help="seconds of CPU work allowed to --impl reference")
    ap.add_argument("--config", default="c5", choices=["c5", "c4"],
                    help="c5 (default, the metric's configuration): 1024^3 box per GPU + BiCGSTAB, weak scaling; "
                         "c4: BASELINE.json configs[3], 4096^2 natural convection + PCG x iters on y-slabs, STRONG scaling")
    return ap.parse_args()


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


# ------------------------------------------------------------------------------------------
# clocks sampler (B200_PROFILING.md: sample nvidia-smi DURING the timed region)
# ------------------------------------------------------------------------------------------
class Clocks:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows, self.stop, self.index = [], False, index
        self.t = threading.Thread(target=self.run, daemon=True)

    def run(self):
        while not self.stop:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5)
                if out.returncode == 0 and out.stdout.strip():
                    self.rows.append([c.strip() for c in out.stdout.strip().split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def __enter__(self):
        self.t.start()
        return self

    def __exit__(self, *a):
        self.stop = True
        self.t.join(timeout=6)

    def summary(self):
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for q, n in enumerate(names) if any(len(r) > 2 + q and r[2 + q].startswith("Active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.rows)}


# ------------------------------------------------------------------------------------------
# workload helpers shared by both arms
# ------------------------------------------------------------------------------------------
def make_params():
    from cfd_b200 import abi
    p = abi.ns_params_default()
    p.dt, p.mu = DT, NU
    p.source_amplitude_u = p.source_amplitude_v = 0.0  # amplitudes 0
    p.max_iter = 1
    return p


def poisson_cfg(method, iters):
    from cfd_b200 import abi
    # tolerances 0: the stopping rule can never fire, so every step does EXACTLY `iters` iterations
    # whatever the box (an 8-GPU box converged after ~41 with the default 1e-6)
    return abi.poisson_config(method, max_iterations=iters, tolerance=0.0, absolute_tolerance=0.0,
                              on_nonconvergence=abi.CFD_B200_NONCONV_CONTINUE)


def synthetic_host_field(n, alloc=None):
    """The C5 state on the host (numpy): u = 0.1 sin2pix cos2piy cos2piz, v = -0.1 cos2pix sin2piy cos2piz,
    w = 0, each + 1e-3 uniform noise (seed 12345); p = 0, rho = 1, T = 300."""
    import numpy as np
    from cfd_b200 import abi
    g = abi.HostGrid(n, n, n, 0.0, 1.0, 0.0, 1.0, 0.0, 1.0)
    f = abi.HostField(n, n, n, alloc=alloc) if alloc else abi.HostField(n, n, n)
    x = np.linspace(0.0, 1.0, n)
    sx, cx = np.sin(2 * np.pi * x), np.cos(2 * np.pi * x)
    rng = np.random.default_rng(12345)
    for k in range(n):  # plane by plane: keeps the temporary small at 1024^3
        cz = cx[k]
        f.u[k] = 0.1 * cz * np.outer(cx, sx) + 1e-3 * (2.0 * rng.random((n, n)) - 1.0)
        f.v[k] = -0.1 * cz * np.outer(sx, cx) + 1e-3 * (2.0 * rng.random((n, n)) - 1.0)
        f.w[k] = 1e-3 * (2.0 * rng.random((n, n)) - 1.0)
    f.p[...] = 0.0
    f.rho[...] = 1.0
    f.T[...] = 300.0
    return g, f


# ------------------------------------------------------------------------------------------
# C4 (BASELINE.json configs[3]): natural convection 2-D 4096 x 4096 (Boussinesq energy + buoyancy + thermal BCs),
# projection + PCG, 1/2/4 GPUs on y-slabs.  The grid is FIXED, so this is strong scaling; SURVEY 8e asks for it to
# be reported honestly: a PCG iteration on a quarter of 16.7 M cells is ~25 us of kernels against two all-reduces
# and a halo row per iteration.
# ------------------------------------------------------------------------------------------
def c4_params():
    """de Vahl Davis Ra = 1e3 set-up of tests/validation/test_natural_convection.c:49-58,139-200; dt = 2.5e-7
    (< dx^2 / (4 alpha) at 4096^2)."""
    import math
    from cfd_b200 import abi
    p = abi.ns_params_default()
    Pr, Ra, beta, gmag, dT = 0.71, 1e3, 0.003333, 9.81, 20.0
    nu = math.sqrt(gmag * beta * dT * Pr / Ra)
    p.dt, p.mu, p.alpha, p.beta, p.T_ref = 2.5e-7, nu, nu / Pr, beta, 300.0
    p.gravity[0], p.gravity[1], p.gravity[2] = 0.0, -gmag, 0.0
    p.source_amplitude_u = p.source_amplitude_v = 0.0
    tb = p.thermal_bc
    tb.left = tb.right = abi.BC_TYPE_DIRICHLET
    tb.bottom = tb.top = tb.front = tb.back = abi.BC_TYPE_NEUMANN
    tb.dirichlet_values.left, tb.dirichlet_values.right = 310.0, 290.0
    p.max_iter = 1
    return p


def c4_host_field(n, rows=None, lo=0, alloc=None):
    """Rows [lo, lo + rows) of the global n x n state: quiescent, T = 310 - 20 x, p = 0, rho = 1."""
    import numpy as np
    from cfd_b200 import abi
    rows = rows or n
    f = abi.HostField(n, rows, 1, alloc=alloc) if alloc else abi.HostField(n, rows, 1)
    x = np.linspace(0.0, 1.0, n)
    for nm in ("u", "v", "w", "p"):
        getattr(f, nm)[...] = 0.0
    f.T[...] = 310.0 - 20.0 * x[None, None, :]
    f.rho[...] = 1.0
    return f


def slab_partition(layers, world):
    """(lo, nloc) per rank: global first layer and local layer count incl. halos -- the gang's rule (csrc/gang.cu)."""
    interior, out, off = layers - 2, [], 0
    for q in range(world):
        own = interior // world + (1 if q < interior % world else 0)
        out.append((off, own + 2))
        off += own
    return out


def c4_main(a, real_out):
    import numpy as np
    import torch
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    td = None
    if world > 1:
        import torch.distributed as td
        td.init_process_group("nccl", device_id=torch.device("cuda", local))
    from cfd_b200 import abi, api, dist, lib
    L = lib.load()
    n = a.n or 4096
    method = "pcg" if a.method == "bicgstab" else a.method
    parity = None
    if td is not None:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import multigpu_parity
        cx = multigpu_parity.Ctx(td, rank, world, local)
        parity = multigpu_parity.run_case(cx, "C4 natconv pcg 258x(64w+2) y-slabs", "natconv_y", (258, 64 * world + 2, 1),
                                          "projection", "pcg", tol=1e-8, nsteps=3, max_iterations=60)
    h = 1.0 / (n - 1)
    lo, nloc = slab_partition(n, world)[rank]
    s = api.Session(n, nloc, 1, h, h, 0.0, 0.0, lo * h, 0.0, rank=rank, nranks=world, device=local)
    if world > 1:
        s.set_slab(n, lo)
    dist.attach_comm(s, rank, world)
    transport = s.transport()
    coords = np.linspace(0.0, 1.0, n)
    s.set_coords(coords, coords[lo:lo + nloc])
    f = c4_host_field(n, nloc, lo)
    s.upload(f)
    prm, pc = c4_params(), poisson_cfg(method, a.iters)
    stream = torch.cuda.ExternalStream(s.stream())

    def barrier():
        torch.cuda.synchronize()
        if td is not None:
            td.barrier()
        torch.cuda.synchronize()

    def run_steps(k):
        rc, st = s.projection(prm, pc, k)
        if rc != 0:
            raise RuntimeError("projection failed rc=%d: %s" % (rc, lib.last_error()))

    run_steps(a.warmup)
    s.reset_kernel_stats()
    s.profile(True)
    l0 = L.cfd_b200_kernel_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    with Clocks(local) as clk:
        e0.record(stream)
        run_steps(a.steps)
        e1.record(stream)
        barrier()
    ms = e0.elapsed_time(e1)
    launches = int(L.cfd_b200_kernel_launch_count() - l0)
    s.profile(False)
    if td is not None:
        t = torch.tensor([ms], device="cuda", dtype=torch.float64)
        td.all_reduce(t, op=td.ReduceOp.MAX)
        ms = float(t.item())
    value = n * n * a.steps / (ms * 1e-3) / 1e6
    interior = (n - 2) * (nloc - 2)
    peak, peak_src = peaks()
    ks = [k for k in s.kernel_stats() if k["total_ms"] > 0]
    ks.sort(key=lambda k: -k["total_ms"])
    ksum = sum(k["total_ms"] for k in ks) or 1.0
    table = []
    for k in ks:
        avg = k["total_ms"] / k["launches"]
        gbs = k["bytes_per_cell"] * interior / (avg * 1e-3) / 1e9
        table.append({"kernel": k["name"], "launches": k["launches"], "avg_us": round(avg * 1e3, 2),
                      "share": round(k["total_ms"] / ksum, 4), "bytes_per_cell": k["bytes_per_cell"],
                      "GBps": round(gbs, 1), "frac": round(gbs / peak, 4)})
    top = next((t for t in table if t["bytes_per_cell"] > 0), None)
    roofline = None
    if top:
        roofline = {"bound": "hbm", "kernel": top["kernel"], "achieved": top["GBps"], "peak": peak, "unit": "GB/s",
                    "frac": top["frac"], "traffic": None, "peak_source": peak_src,
                    "algorithmic_bytes_per_launch": top["bytes_per_cell"] * interior, "kernels": table,
                    "kernel_ms_per_step": round(ksum / a.steps, 3),
                    "note": "slab of %d rows per rank: sum of kernel times vs ms_per_step shows what launch gaps and "
                            "collectives cost at this size" % (nloc - 2)}
    s.close()
    # ---- e2e: the drop-in symbol with a pinned host flow_field (u, v, w, p, T both ways), rank 0 drives all GPUs
    e2e = None
    if not a.no_e2e:
        store = None
        if td is not None:
            td.barrier()
            store = td.distributed_c10d._get_default_store()
        if rank == 0:
            try:
                keep = []

                def alloc(m):
                    import ctypes as C
                    ptr = L.cfd_b200_host_alloc(m * 8)
                    keep.append(ptr)
                    return np.frombuffer((C.c_double * m).from_address(ptr), dtype=np.float64)

                g = abi.HostGrid(n, n, 1, 0.0, 1.0, 0.0, 1.0)
                fh = c4_host_field(n, alloc=alloc)
                api.set_poisson(pc)
                api.set_device_count(world)
                steps = max(2, min(a.steps, 5))
                rc = api.solve_method_gpu("projection", fh, g, prm)
                assert rc == 0, lib.last_error()
                t0 = time.perf_counter()
                for _ in range(steps):
                    rc = api.solve_method_gpu("projection", fh, g, prm)
                    assert rc == 0, lib.last_error()
                dt = time.perf_counter() - t0
                e2e = {"value": n * n * steps / dt / 1e6, "unit": UNIT, "h2d_bytes_per_step": 5 * 8 * n * n,
                       "d2h_bytes_per_step": 5 * 8 * n * n, "steps": steps, "ms_per_step": dt / steps * 1e3,
                       "api": "solve_projection_method_gpu (C-ABI drop-in), pinned host flow_field, "
                              "cfd_b200_set_device_count(%d)" % world}
                api.set_device_count(0)
                L.cfd_b200_release_cached_sessions()
                for ptr in keep:
                    L.cfd_b200_host_free(ptr)
            finally:
                if store is not None:
                    store.set("cfd_b200_e2e_done", "1")
        elif store is not None:
            import datetime
            store.wait(["cfd_b200_e2e_done"], datetime.timedelta(seconds=900))
    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
                "ms_per_step": ms / a.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "f64", "data": "synthetic",
                "config": {"workload": "C4 natural convection 2-D %d x %d (energy + buoyancy + thermal BCs), projection step + "
                                       "%s x%d iterations (fixed), y-slab decomposition over %d GPU(s), FIXED global grid"
                                       % (n, n, method, a.iters, world),
                           "grid_global": [n, n, 1], "poisson": method, "poisson_iterations": a.iters, "dt": prm.dt,
                           "parallelism": "yslab%d" % world, "transport": transport,
                           "l2": "fields of %.0f MB per rank against 126 MB of L2: %s" %
                                 (8 * n * nloc / 1e6, "larger than L2" if 8 * n * nloc * 9 > 126e6 else "cache resident")},
                "roofline": roofline, "cpu_baseline": None, "e2e": e2e, "gpu_launches": launches, "clocks": clk.summary()}
        if parity is not None:
            line["parity"] = parity
        real_out.write(json.dumps(line) + "\n")
    if td is not None:
        td.barrier()
        td.destroy_process_group()
    return 0


# ------------------------------------------------------------------------------------------
# CPU arm: the reference's own code on the host cores (oracle/_ref, see oracle/Makefile):
# reference Poisson solver objects + the restated OpenMP stencil loops of solver_projection_omp.c
# ------------------------------------------------------------------------------------------
def cpu_arm(method, iters, steps, warmup, budget_s, n_force=0):
    import numpy as np
    import oracle
    from cfd_b200 import abi
    kind = "reference"
    try:
        orc = oracle.Oracle("ref")
    except (FileNotFoundError, OSError):
        orc, kind = oracle.Oracle(), "port"  # scalar restatement, 1 thread
    # (OMP_PROC_BIND pins the calling thread once libgomp runs, so the affinity mask is not the thread count)
    cores = int(os.environ.get("OMP_NUM_THREADS") or host_cores()) if kind == "reference" else 1
    meth = {"jacobi": abi.POISSON_METHOD_JACOBI, "rbsor": abi.POISSON_METHOD_REDBLACK_SOR,
            "cg": abi.POISSON_METHOD_CG, "pcg": abi.POISSON_METHOD_CG, "bicgstab": abi.POISSON_METHOD_BICGSTAB}[method]
    pc, p = poisson_cfg(method, iters), make_params()
    names = {1: "scalar", 2: "omp", 3: "simd(avx2+omp)"}

    def run(n, backend, nsteps):
        if kind == "reference":
            orc.lib.orc_set_ref_backend(backend)
        g, f = synthetic_host_field(n)
        t0 = time.perf_counter()
        rc, st = orc.projection(f, g, p, pc, nsteps)
        dt = time.perf_counter() - t0
        if rc != 0:
            raise RuntimeError("CPU projection failed rc=%d" % rc)
        return dt

    # Size first (a scalar 64^3 step gives the cost per cell), then probe every backend the reference offers for
    # this method AT THAT SIZE and keep the fastest: its AVX2+OpenMP "SIMD" BiCGSTAB loses to the scalar one on a
    # 64^3 box (thread start-up per BLAS-1 call) and wins from ~160^3 up, where the timing is taken.
    t64 = run(64, 1, 1)
    per_cell_step = t64 / 64 ** 3
    n = n_force
    if not n:
        n = 160
        for cand in (192, 224, 256, 320, 384):
            # probes (<= 3 backends) + warm-up + timed steps, the scalar rate as the pessimistic estimate
            if per_cell_step * cand ** 3 * (steps + warmup + 2) <= budget_s:
                n = cand
    backend, probes = 1, {}
    if kind == "reference":
        for b in (3, 2, 1):
            if not orc.lib.orc_ref_backend_available(meth, b):
                continue
            probes[names[b]] = round(run(n, b, 1), 3)
        backend = min((t, b) for b, t in ((bb, probes[names[bb]]) for bb in (1, 2, 3) if names[bb] in probes))[1]
    for _ in range(max(0, warmup - (1 if kind == "reference" else 0))):  # the probe of the winner was its first warm-up
        run(n, backend, 1)
    dt = run(n, backend, steps)
    mlups = n ** 3 * steps / dt / 1e6
    sample = ("%d^3 box, %d projection steps (+%d warm-up), %s x%d iterations, reference Poisson backend %s "
              "(fastest of the backends probed at this size: %s s/step), restated OpenMP stencil loops, %d threads, "
              "OMP_PROC_BIND=%s OMP_PLACES=%s" % (n, steps, warmup, method, iters,
                                                   names.get(backend, "?") if kind == "reference" else "scalar port",
                                                   json.dumps(probes), cores, os.environ.get("OMP_PROC_BIND", "-"),
                                                   os.environ.get("OMP_PLACES", "-")))
    return {"value": mlups, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample, "n": n,
            "ms_per_step": dt / steps * 1e3, "backend": names.get(backend, "scalar port"), "probe_s_per_step": probes}


def reference_main(a, real_out):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    r = cpu_arm(a.method, a.iters, a.steps, a.warmup, budget_s=a.budget, n_force=a.cpu_n)
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": a.gpus,
            "steps": a.steps, "warmup": a.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "C5 synthetic 3-D box, projection step + %s x%d iterations (fixed), CPU sample %d^3"
                                   % (a.method, a.iters, r["n"]), "poisson": a.method, "poisson_iterations": a.iters,
                       "dt": DT, "nu": NU},
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample", "backend")},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    real_out.write(json.dumps(line) + "\n")
    return 0


# ------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------
FIELDS_RESIDENT = {"bicgstab": 17, "cg": 13, "pcg": 13, "rbsor": 9, "jacobi": 10}  # device arrays of one slab


def pick_n(method, want, free_bytes):
    for n in ([want] if want else [1024, 896, 768, 640, 512, 384, 256]):
        if (FIELDS_RESIDENT[method] + 0.5) * 8.0 * n ** 3 <= 0.92 * free_bytes:
            return n
    raise RuntimeError("not enough device memory for the smallest bench box")


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ours_main(a):
    import numpy as np
    import torch
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: cfd_b200 has no CPU fallback")
    torch.cuda.set_device(local)
    td = None
    if world > 1:
        import torch.distributed as td
        td.init_process_group("nccl", device_id=torch.device("cuda", local))
    from cfd_b200 import abi, api, dist, lib
    L = lib.load()
    free, _total = torch.cuda.mem_get_info()
    n = pick_n(a.method, a.n, free)
    if td is not None:  # every rank must use the same box
        t = torch.tensor([n], device="cuda", dtype=torch.int64)
        td.all_reduce(t, op=td.ReduceOp.MIN)
        n = int(t.item())
    parity = None
    if td is not None:
        # driver-visible decomposed-vs-oracle evidence at this rank count (the GPU test box has one GPU):
        # the headline method and the bitwise Jacobi twin on a box of 2 x 2 TMA tiles and >= 2 z-chunks per rank
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import multigpu_parity
        parity = multigpu_parity.bench_parity(multigpu_parity.Ctx(td, rank, world, local))
    h = 1.0 / (n - 1)
    gnz = (n - 2) * world + 2
    s = api.Session(n, n, n, h, h, h, 0.0, 0.0, rank * (n - 2) * h, rank=rank, nranks=world, device=local)
    dist.attach_comm(s, rank, world)
    transport = s.transport()
    s.init_synthetic(12345, rank * (n - 2), gnz)
    prm, pc = make_params(), poisson_cfg(a.method, a.iters)
    stream = torch.cuda.ExternalStream(s.stream())

    def barrier():
        torch.cuda.synchronize()
        if td is not None:
            td.barrier()
        torch.cuda.synchronize()

    def run_steps(k):
        rc, st = s.projection(prm, pc, k)
        if rc != 0:
            raise RuntimeError("projection failed rc=%d: %s" % (rc, lib.last_error()))
        return st

    run_steps(a.warmup)
    s.reset_kernel_stats()
    s.profile(True)
    l0 = L.cfd_b200_kernel_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    with Clocks(local) as clk:
        e0.record(stream)
        st = run_steps(a.steps)
        e1.record(stream)
        barrier()
    ms = e0.elapsed_time(e1)
    launches = int(L.cfd_b200_kernel_launch_count() - l0)
    s.profile(False)
    kstats = s.kernel_stats()
    if td is not None:
        t = torch.tensor([ms], device="cuda", dtype=torch.float64)
        td.all_reduce(t, op=td.ReduceOp.MAX)
        ms = float(t.item())
    cells_global = n * n * gnz
    value = cells_global * a.steps / (ms * 1e-3) / 1e6

    # ---- roofline of the dominant kernel (per-launch CUDA events inside the timed region) ----
    interior = (n - 2) ** 3
    peak, peak_src = peaks()
    kstats = [k for k in kstats if k["total_ms"] > 0 and k["bytes_per_cell"] > 0]
    kstats.sort(key=lambda k: -k["total_ms"])
    ksum = sum(k["total_ms"] for k in kstats) or 1.0
    table = []
    for k in kstats:
        avg = k["total_ms"] / k["launches"]
        gbs = k["bytes_per_cell"] * interior / (avg * 1e-3) / 1e9
        table.append({"kernel": k["name"], "launches": k["launches"], "avg_ms": round(avg, 4),
                      "share": round(k["total_ms"] / ksum, 4), "bytes_per_cell": k["bytes_per_cell"],
                      "GBps": round(gbs, 1), "frac": round(gbs / peak, 4)})
    top = table[0] if table else None
    roofline = None
    if top:
        traffic, traffic_src = ncu_traffic(top["kernel"], n)
        for row in table:
            row["traffic"] = ncu_traffic(row["kernel"], n)[0]
        roofline = {"bound": "hbm", "kernel": top["kernel"], "achieved": top["GBps"], "peak": peak, "unit": "GB/s",
                    "frac": top["frac"], "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                    "algorithmic_bytes_per_launch": top["bytes_per_cell"] * interior, "kernels": table}
    return dict(rank=rank, world=world, local=local, n=n, gnz=gnz, ms=ms, value=value, launches=launches,
                roofline=roofline, clocks=clk.summary(), session=s, prm=prm, pc=pc, td=td, L=L, stats=st,
                parity=parity, transport=transport)


NCU_LATEST = os.path.join(ROOT, "profiles", "ncu_headline_latest.csv")
# kernel-family name (csrc/session.cu kKernelNames) -> substring of the CUDA kernel name in the ncu report
NCU_KERNEL_OF = {"bicg_dir_matvec": "k_bicg_dir_tma", "bicg_s_matvec": "k_bicg_s_tma", "bicg_update": "k_bicg_update",
                 "predictor": "k_predictor", "divergence_rhs": "k_divergence", "corrector": "k_corrector",
                 "krylov_init": "k_krylov_init", "cg_dir_matvec": "k_cg_dir_tma", "cg_update": "k_cg_update",
                 "jacobi_sweep": "k_jacobi_tma", "rbsor_fused_iteration": "k_rbsor_fused"}


def ncu_traffic(family, n):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `family` from profiles/ncu_headline_latest.csv
    (tools/ncu_summary.py over one `ncu --set full` capture of THIS bench command, regenerated in the round's
    final GPU pass; its title line names the box).  Returns (bytes or None, source)."""
    import csv
    if not os.path.exists(NCU_LATEST) or family not in NCU_KERNEL_OF:
        return None, None
    with open(NCU_LATEST) as fh:
        title = fh.readline()
        rows = list(csv.reader(fh))
    if "box=%d" % n not in title:
        return None, os.path.relpath(NCU_LATEST, ROOT) + " (other box: not applicable)"
    hdr = rows[0]
    try:
        kn, tm, rd, wr = (hdr.index(c) for c in ("Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum",
                                                 "dram__bytes_write.sum"))
    except ValueError:
        return None, None
    unit = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}.get(rows[1][rd], 1e9)
    best = None
    for r in rows[2:]:
        if len(r) <= max(kn, tm, rd, wr) or NCU_KERNEL_OF[family] not in r[kn]:
            continue
        try:
            t, bts = float(r[tm]), (float(r[rd]) + float(r[wr])) * unit
        except ValueError:
            continue
        if best is None or t > best[0]:  # the full-size launch of the family (decide-only launches are tiny)
            best = (t, bts)
    return (best[1] if best else None), os.path.relpath(NCU_LATEST, ROOT) + ": " + title.strip("# \n")[:160]


def e2e_arm(a, r):
    """The same step through the reference-facing boundary with HOST buffers: the drop-in symbol
    solve_projection_method_gpu(field, grid, params{max_iter=1}, config) -- what the reference's registry wrapper
    gpu_projection_step calls (solver_registry.c:1121-1135) -- with the caller's flow_field in pinned host memory;
    every step copies u,v,w,p H2D and back.  N > 1: the device-resident ranks release their slabs and RANK 0 alone
    makes the same call with cfd_b200_set_device_count(N): one host field of n x n x ((n-2)N+2) cells, N slabs, one
    host thread per GPU inside that process (csrc/gang.cu) -- the reference's process model."""
    import ctypes as C
    import numpy as np
    import psutil
    import torch
    from cfd_b200 import abi, api, lib
    L, s, n, world, td, rank = r["L"], r["session"], r["n"], r["world"], r["td"], r["rank"]
    s.close()  # the drop-in path owns its own cached runner; both do not fit at 1024^3
    torch.cuda.empty_cache()
    store = None
    if td is not None:
        td.barrier()  # every slab is released before rank 0 allocates on every device
        store = td.distributed_c10d._get_default_store()
        if rank != 0:
            import datetime
            store.wait(["cfd_b200_e2e_done"], datetime.timedelta(seconds=1800))  # host-side wait: the GPUs are rank 0's
            return None
    try:
        return _e2e_rank0(a, r, C, np, psutil, torch, abi, api, lib)
    finally:
        if store is not None:
            store.set("cfd_b200_e2e_done", "1")


def _e2e_rank0(a, r, C, np, psutil, torch, abi, api, lib):
    L, n, world = r["L"], r["n"], r["world"]
    # global box: n x n x nzg; bounded by what this host can hold (6 arrays, 4 of them pinned) and pin in
    # reasonable time -- stated in "box"
    avail = psutil.virtual_memory().available
    cap = min(0.55 * avail, 215e9)
    nzg = (n - 2) * world + 2
    while 6.1 * 8.0 * n * n * nzg > cap and nzg > 66:
        nzg = (nzg - 2) * 3 // 4 + 2
    keep, count = [], [0]

    def alloc(m):
        count[0] += 1
        if count[0] > 4:  # rho, T: never transferred by the projection path without energy/buoyancy
            return abi.aligned_zeros(m)
        p = L.cfd_b200_host_alloc(m * 8)
        if not p:
            raise MemoryError("cfd_b200_host_alloc(%d bytes) failed" % (m * 8))
        keep.append(p)
        return np.frombuffer((C.c_double * m).from_address(p), dtype=np.float64)

    h = 1.0 / (n - 1)
    g = abi.HostGrid(n, n, nzg, 0.0, 1.0, 0.0, 1.0, 0.0, (nzg - 1) * h)
    f = abi.HostField(n, n, nzg, alloc=alloc)
    # the C5 state, filled by the device generator slab by slab (same generator as the device-resident arm)
    mask = api.FM_U | api.FM_V | api.FM_W | api.FM_P
    chunk = min(nzg, 64)
    gen = api.Session(n, n, chunk, h, h, h)
    view = abi.HostField(n, n, chunk, alloc=lambda m: np.empty(m))
    for k0 in range(0, nzg, chunk):
        k0 = min(k0, nzg - chunk)
        gen.init_synthetic(12345, k0, nzg)
        gen.download(view, mask)
        for nm in ("u", "v", "w", "p"):
            getattr(f, nm)[k0:k0 + chunk] = getattr(view, nm)
    gen.close()
    del view
    f.rho[...] = 1.0
    f.T[...] = 300.0
    api.set_poisson(r["pc"])
    api.set_device_count(world)
    prm = r["prm"]
    steps = max(1, min(a.steps, 2))
    cells = n * n * nzg

    def timed(call):
        rc = call()  # warm-up: allocations, cached runner / first full upload
        if rc != 0:
            raise RuntimeError("e2e projection failed rc=%d: %s" % (rc, lib.last_error()))
        for d in range(torch.cuda.device_count()):
            torch.cuda.synchronize(d)
        t0 = time.perf_counter()
        for _ in range(steps):
            rc = call()
            if rc != 0:
                raise RuntimeError("e2e projection failed rc=%d: %s" % (rc, lib.last_error()))
        for d in range(torch.cuda.device_count()):
            torch.cuda.synchronize(d)
        return time.perf_counter() - t0

    dt = timed(lambda: api.solve_method_gpu("projection", f, g, prm))
    out = {"value": cells * steps / dt / 1e6, "unit": UNIT, "h2d_bytes_per_step": 4 * 8 * cells,
           "d2h_bytes_per_step": 4 * 8 * cells, "steps": steps, "ms_per_step": dt / steps * 1e3,
           "api": "solve_projection_method_gpu (C-ABI drop-in), pinned host flow_field"
                  + ("" if world == 1 else ", cfd_b200_set_device_count(%d): %d z-slabs driven by this one call" % (world, world)),
           "box": "%d x %d x %d global (%s)" % (n, n, nzg, "%d^3 per GPU" % n if nzg == (n - 2) * world + 2 else
                                                 "z extent reduced to what this host can pin; %d planes per GPU" % ((nzg - 2) // world))}
    # ---- opt-in host-sync policy: the same step through the persistent ns_solver_t with
    #      CFD_B200_SYNC_UPLOAD_BOUNDARY -- the caller edits boundary cells only, so only the six faces go up
    try:
        L.cfd_b200_release_cached_sessions()
        slv = api.GpuSolver("projection_gpu")
        slv.set_host_sync(api.SYNC_UPLOAD_BOUNDARY)
        dtb = timed(lambda: slv.step(f, g, prm)[0])
        slv.close()
        faces = 2 * (n * n + n * nzg + n * nzg)
        out["e2e_boundary_only"] = {"value": cells * steps / dtb / 1e6, "unit": UNIT, "h2d_bytes_per_step": 4 * 8 * faces,
                                    "d2h_bytes_per_step": 4 * 8 * cells, "ms_per_step": dtb / steps * 1e3,
                                    "api": "cfd_b200_create_projection_gpu_solver() + cfd_b200_solver_set_host_sync("
                                           "CFD_B200_SYNC_UPLOAD_BOUNDARY); solver->step() per step"}
    except Exception as ex:  # the headline e2e stands on its own
        out["e2e_boundary_only"] = {"value": None, "note": str(ex)[:200]}
    api.set_device_count(0)
    L.cfd_b200_release_cached_sessions()
    for p in keep:
        L.cfd_b200_host_free(p)
    return out


def main():
    a = parse()
    # libraries (NCCL prints its version banner) write to fd 1; the contract is ONE JSON line on
    # stdout, so everything else goes to stderr and the line is written to the saved descriptor
    sys.stdout.flush()
    real_out = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    try:
        return _main(a, real_out)
    finally:
        real_out.flush()


def _main(a, real_out):
    if a.impl == "reference":
        # all host cores for the reference's OpenMP / AVX2 backends (set before libgomp loads)
        os.environ.setdefault("OMP_NUM_THREADS", str(host_cores()))
        os.environ.setdefault("OMP_PROC_BIND", "close")
        os.environ.setdefault("OMP_PLACES", "cores")
        return reference_main(a, real_out)
    if a.config == "c4":
        return c4_main(a, real_out)
    r = ours_main(a)
    e2e = None
    if not a.no_e2e:
        try:
            e2e = e2e_arm(a, r)
        except MemoryError as ex:
            e2e = {"value": None, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0, "note": str(ex)}
    e2e_b = e2e.pop("e2e_boundary_only", None) if isinstance(e2e, dict) else None
    cpu = None
    if r["rank"] == 0 and r["world"] == 1 and not a.no_cpu:
        # a fresh process: torch's OpenMP runtime and thread affinity stay out of the CPU measurement
        cmd = [sys.executable, os.path.abspath(__file__), "--impl", "reference", "--steps", "1", "--warmup", "1",
               "--budget", "24", "--method", a.method, "--iters", str(a.iters), "--cpu-n", str(a.cpu_n)]
        env = {k: v for k, v in os.environ.items() if not k.startswith("OMP_")}
        out = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=600)
        try:
            cpu = json.loads(out.stdout.strip().splitlines()[-1])["cpu_baseline"]
        except Exception:
            cpu = {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": "failed: " + out.stderr[-300:]}
    if r["rank"] == 0:
        n, world = r["n"], r["world"]
        line = {"metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": world, "steps": a.steps,
                "warmup": a.warmup, "ms_per_step": r["ms"] / a.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": "C5 synthetic 3-D box %d^3 per GPU, z-slab decomposition, projection step + "
                                       "%s x%d iterations (fixed: cap reached => continue)" % (n, a.method, a.iters),
                           "grid_global": [n, n, r["gnz"]], "poisson": a.method, "poisson_iterations": a.iters,
                           "dt": DT, "nu": NU, "parallelism": "zslab%d" % world, "transport": r["transport"],
                           "l2": "inputs larger than L2 (each field %.1f GB vs 126 MB L2); no flush needed" % (8 * n ** 3 / 1e9)},
                "roofline": r["roofline"], "cpu_baseline": cpu, "e2e": e2e, "e2e_boundary_only": e2e_b,
                "gpu_launches": r["launches"], "clocks": r["clocks"]}
        if r["parity"] is not None:
            line["parity"] = r["parity"]
        real_out.write(json.dumps(line) + "\n")
    if r["td"] is not None:
        r["td"].barrier()
        r["td"].destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
