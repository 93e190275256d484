#!/usr/bin/env python
"""C3 on z-slabs: n^3 box, projection + Red-Black SOR x iters, strong scaling over the ranks of one node.

    python tools/c3_slabs.py --size 256                                        # 1 GPU
    python -m torch.distributed.run --nproc-per-node 2 tools/c3_slabs.py    # 2 GPUs (peer transport)

Runs the same steps twice -- with the one-pass sweep (poisson_tma3.inc; on slabs: two-layer exchange once per
iteration) and with the half-sweeps it replaces (CFD_B200_NO_RB_FUSED=1) -- and prints one JSON line per path:
ms per step (device time, max over ranks) and the kernel table of rank 0.  Companion of bench.py --config c4;
same timing method.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", dest="n", type=int, default=256)
    ap.add_argument("--iters", type=int, default=30)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    a = ap.parse_args()
    import torch
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    td = None
    if world > 1:
        import torch.distributed as td
        td.init_process_group("nccl", device_id=torch.device("cuda", local))
    from bench import peaks, slab_partition
    from cfd_b200 import abi, api, dist, lib
    n = a.n
    h = 1.0 / (n - 1)
    lo, nloc = slab_partition(n, world)[rank]
    prm = abi.ns_params_default()
    prm.dt, prm.mu = 1e-4, 0.01
    prm.source_amplitude_u = prm.source_amplitude_v = 0.0
    pc = abi.poisson_config("rbsor", max_iterations=a.iters, tolerance=0.0, absolute_tolerance=0.0,
                            on_nonconvergence=abi.CFD_B200_NONCONV_CONTINUE)
    peak, _ = peaks()

    def barrier():
        torch.cuda.synchronize()
        if td is not None:
            td.barrier()
        torch.cuda.synchronize()

    for path in ("one_pass", "half_sweeps"):
        if path == "half_sweeps":
            os.environ["CFD_B200_NO_RB_FUSED"] = "1"
        else:
            os.environ.pop("CFD_B200_NO_RB_FUSED", None)
        s = api.Session(n, n, nloc, h, h, h, 0.0, 0.0, lo * h, rank=rank, nranks=world, device=local)
        if world > 1:
            s.set_slab(n, lo)
        dist.attach_comm(s, rank, world)
        s.init_synthetic(12345, lo, n)
        stream = torch.cuda.ExternalStream(s.stream())

        def run(k):
            rc, _ = s.projection(prm, pc, k)
            if rc != 0:
                raise RuntimeError("projection failed rc=%d: %s" % (rc, lib.last_error()))

        run(a.warmup)
        s.reset_kernel_stats()
        s.profile(True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record(stream)
        run(a.steps)
        e1.record(stream)
        barrier()
        ms = e0.elapsed_time(e1)
        s.profile(False)
        if td is not None:
            t = torch.tensor([ms], device="cuda", dtype=torch.float64)
            td.all_reduce(t, op=td.ReduceOp.MAX)
            ms = float(t.item())
        interior = (n - 2) * (n - 2) * (nloc - 2)
        rows = []
        for k in sorted(s.kernel_stats(), key=lambda k: -k["total_ms"]):
            if k["total_ms"] <= 0:
                continue
            avg = k["total_ms"] / k["launches"]
            gbs = k["bytes_per_cell"] * interior / (avg * 1e-3) / 1e9
            rows.append({"kernel": k["name"], "launches": k["launches"], "avg_us": round(avg * 1e3, 2),
                         "GBps": round(gbs, 1), "frac": round(gbs / peak, 4)})
        transport = s.transport()
        s.close()
        if rank == 0:
            print(json.dumps({"workload": "C3-shaped: %d^3 box, projection + rbsor x %d" % (n, a.iters), "path": path,
                              "n_gpus": world, "transport": transport,
                              "slab_planes": nloc - 2, "steps": a.steps, "ms_per_step": round(ms / a.steps, 4),
                              "MLUPS": round(n ** 3 * a.steps / (ms * 1e-3) / 1e6, 1), "kernels": rows[:6]}), flush=True)
    if td is not None:
        td.barrier()
        td.destroy_process_group()


if __name__ == "__main__":
    main()
