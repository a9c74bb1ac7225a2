"""The compiled C / OpenMP CPU restatement (oracle/cpu_step.c — bench.py's cpu_baseline and `--impl reference` arm)
pinned against the NumPy oracle on the same seeded inputs: tendencies bit-exact (same association order, no FMA
contraction), fields after full RK3 steps to FFT rounding (radix-2 here, pocketfft there)."""
import numpy as np
import pytest

import oracle as o
from oracle.model import update_state
from oracle.cpu_step import CpuStepModel


def _pair(N, L, seed, project):
    rng = np.random.default_rng(seed)
    ics = {n: rng.uniform(-1, 1, N) for n in "uvw"}
    og = o.RectilinearGrid(np.float64, size=N, extent=L, topology=(o.Periodic,) * 3)
    om = o.NonhydrostaticModel(og, advection=o.WENO(np.float64, 5))
    o.set_model(om, enforce_incompressibility=project, **ics)
    cm = CpuStepModel(N, L)
    cm.set(ics["u"], ics["v"], ics["w"], enforce_incompressibility=project)
    return om, cm


@pytest.mark.parametrize("N", [(16, 8, 32), (8, 8, 8)])
def test_tendencies_bit_exact(N):
    om, cm = _pair(N, (1.0, 2.0, 3.0), 1234, project=False)
    update_state(om)
    cm.compute_tendencies()
    for n in "uvw":
        np.testing.assert_array_equal(om.Gn[n].interior, cm.get("G" + n))


def test_ten_rk3_steps_match_the_numpy_oracle():
    om, cm = _pair((16, 16, 16), (1.0, 1.0, 1.0), 7, project=True)
    for _ in range(10):
        o.time_step(om, 1e-3)
    cm.time_step(1e-3, 10)
    for a, n in enumerate("uvw"):
        y = om.velocities[a].interior
        assert np.linalg.norm(cm.get(n) - y) / np.linalg.norm(y) < 1e-12
    y = om.pNHS.interior
    assert np.linalg.norm(cm.get("p") - y) / np.linalg.norm(y) < 1e-11   # p ~ 1e-13 · (|u|²/Δt-scale conditioning)


def test_rejects_non_power_of_two():
    with pytest.raises(ValueError):
        CpuStepModel((12, 16, 16), (1.0, 1.0, 1.0))
