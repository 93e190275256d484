"""bench.py contract checks that need no GPU: the reference arm (`--impl reference`) prints one JSON
line with the keys the driver reads, and under torchrun only rank 0 works."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref", "liboracle_ref.so")


def _run(env_extra=None):
    env = dict(os.environ)
    env.update(env_extra or {})
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                        "--warmup", "1", "--budget", "2"], capture_output=True, text=True, env=env, timeout=300)
    assert r.returncode == 0, r.stderr[-800:]
    return [ln for ln in r.stdout.splitlines() if ln.strip()]


def test_reference_arm_prints_one_contract_line():
    lines = _run()
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "projection_step_mlups_fp64" and d["unit"] == "MLUPS"
    assert d["higher_is_better"] is True and d["value"] > 0 and d["dtype"] == "f64"
    cb = d["cpu_baseline"]
    assert cb["kind"] in ("reference", "port") and cb["cores"] >= 1 and cb["value"] == d["value"]
    if os.path.exists(REF):
        assert cb["kind"] == "reference"
    assert d["e2e"] == {"value": d["value"], "unit": "MLUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "workload" in d["config"] and "model" not in d["config"]


def test_reference_arm_other_ranks_exit_quietly():
    assert _run({"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"}) == []


def test_slab_partition_matches_the_gangs_rule():
    """bench.py --config c4 cuts the FIXED grid like csrc/gang.cu does: interior layers dealt out as evenly as
    possible, the first (interior % G) ranks one layer more, one boundary / halo layer each side."""
    sys.path.insert(0, ROOT)
    import bench
    for layers, world in ((4096, 1), (4096, 2), (4096, 4), (131, 3), (29, 4), (1024, 8)):
        parts = bench.slab_partition(layers, world)
        assert len(parts) == world and parts[0][0] == 0
        own = [n - 2 for _, n in parts]
        assert sum(own) == layers - 2 and max(own) - min(own) <= 1 and own == sorted(own, reverse=True)
        for (lo, n), (lo2, _) in zip(parts, parts[1:]):
            assert lo2 == lo + n - 2  # the next slab starts on my last owned layer's successor minus its halo
        assert parts[-1][0] + parts[-1][1] == layers


def test_roofline_traffic_comes_from_the_committed_ncu_summary():
    """roofline.traffic is read from profiles/ncu_headline_latest.csv (no number typed into bench.py): the file
    parses, names its box, and the DRAM traffic of the dominant kernel is within 0.9 .. 1.3 of its algorithmic bytes."""
    sys.path.insert(0, ROOT)
    import bench
    if not os.path.exists(bench.NCU_LATEST):
        pytest.skip("no committed ncu summary yet")
    with open(bench.NCU_LATEST) as fh:
        title = fh.readline()
    assert "box=" in title and "ncu --set full" in title
    n = int(title.split("box=")[1].split(";")[0].split()[0])
    got, src = bench.ncu_traffic("bicg_update", n)
    assert got and src and "ncu_headline_latest.csv" in src
    algorithmic = 56.0 * (n - 2) ** 3
    assert 0.9 * algorithmic <= got <= 1.3 * algorithmic, (got, algorithmic)
    assert bench.ncu_traffic("bicg_update", n + 2) [0] is None  # another box: not applicable
