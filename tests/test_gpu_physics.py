"""The reference's long physics gates on the GPU path (tools/physics_gates.py holds the drivers):

  * de Vahl Davis natural convection Ra = 1e3, 41 x 41, marched to steady state
    (tests/validation/test_natural_convection.c:139-307): u_max*, v_max*, Nu within 10 % of the benchmark, the same
    number of steps to the kinetic-energy criterion as the oracle, values equal to the oracle's to 6 digits;
  * Ghia Re = 100 lid-driven cavity at the FULL validation size, 129 x 129, 50 000 steps, dt = 2e-4
    (tests/validation/test_cavity_backends.c:57-59, gate RMS < 0.10), device resident (DOWNLOAD_LAZY); the oracle's
    numbers for the same run are kept in tests/golden/physics_gates_oracle.json (made here by
    `python tools/physics_gates.py --impl oracle`, ~15 min of CPU) and must agree to 3 significant figures.
"""
import json
import os
import sys

import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(os.path.dirname(HERE), "tools"))
GOLD = os.path.join(HERE, "golden", "physics_gates_oracle.json")


@pytest.fixture
def clean():
    from cfd_b200 import api
    yield
    api.set_poisson("cg")


def test_de_vahl_davis_ra1e3_gates_and_oracle_agreement(clean):
    import physics_gates
    g = physics_gates.dvd("gpu")
    o = physics_gates.dvd("oracle")
    assert g["converged"] and g["pass"], g
    assert max(g["rel_err"].values()) < 0.10
    assert g["steps"] == o["steps"], (g["steps"], o["steps"])
    for k in ("u_max", "v_max", "Nu"):
        assert abs(g[k] - o[k]) <= 1e-6 * abs(o[k]), (k, g[k], o[k])


def test_ghia_129_full_validation_run(clean):
    import physics_gates
    g = physics_gates.ghia129("gpu", 50000)
    assert g["pass"], g
    if os.path.exists(GOLD):
        with open(GOLD) as fh:
            o = json.load(fh)["ghia129"]
        assert o["steps"] == g["steps"]
        for k in ("rms_u", "rms_v", "u_center"):
            assert abs(g[k] - o[k]) <= 5e-3 * abs(o[k]), (k, g[k], o[k])  # 3 significant figures
