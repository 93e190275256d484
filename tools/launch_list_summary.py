#!/usr/bin/env python
"""Condense an ncu launch list (`ncu --metrics gpu__time_duration.sum --clock-control none --csv`) of
bench.py into a per-kernel table and set its kernel SHARES beside the shares bench.py measured with
CUDA events in the un-profiled run (the ncu per-launch times are cold-cache and serialised: the
share must agree, not the absolute).

    python tools/launch_list_summary.py launches.csv bench.json > summary.md
"""
import csv
import json
import re
import sys

# kernel symbol -> bench.py family name (cfd_b200/csrc/session.cu: kKernelNames)
FAMILY = [("k_bicg_update", "bicg_update"), ("k_bicg_dir", "bicg_dir_matvec"), ("k_bicg_s", "bicg_s_matvec"),
          ("k_predictor", "predictor"), ("k_divergence", "divergence_rhs"), ("k_corrector", "corrector"),
          ("k_krylov_init", "krylov_init"), ("k_copy", "copy"), ("k_field_stats", "field_stats"),
          ("k_boundary", "boundary_cells"), ("k_neumann", "neumann_fill"), ("k_zero_ctl", "ctl_reset"),
          ("k_synth_init", "synth_init")]


def family(sym):
    for key, fam in FAMILY:
        if key in sym:
            return fam
    return re.sub(r"\(.*", "", sym)[:40]


def main():
    launches, bench = sys.argv[1], sys.argv[2]
    rows = [r for r in csv.reader(l for l in open(launches) if l.startswith('"'))]
    hdr = rows[0]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = {}
    for r in rows[1:]:
        ns = float(r[vi].replace(",", "")) * {"ns": 1.0, "us": 1e3, "ms": 1e6, "s": 1e9}.get(r[ui], 1.0)
        a = agg.setdefault(family(r[ki]), [0, 0.0])
        a[0] += 1
        a[1] += ns
    b = json.loads([l for l in open(bench) if l.strip().startswith("{")][-1])
    ev = {k["kernel"]: k for k in b["roofline"]["kernels"]}
    step_fams = set(ev)
    tot = sum(v[1] for f, v in agg.items() if f in step_fams)
    print("| kernel family | ncu launches | ncu avg ms | ncu share of step kernels | bench.py (CUDA events) avg ms | bench.py share |")
    print("|---|---|---|---|---|---|")
    for f, (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        if f in step_fams:
            print("| %s | %d | %.3f | %.4f | %.3f | %.4f |" % (f, n, ns / n / 1e6, ns / tot, ev[f]["avg_ms"], ev[f]["share"]))
        else:
            print("| %s (outside the step / not timed per family) | %d | %.3f | - | - | - |" % (f, n, ns / n / 1e6))


if __name__ == "__main__":
    main()
