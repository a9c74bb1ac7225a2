"""GPU: the interior / boundary split of the slab sweep (nh_dist_tendency_substep_part) against the single launch, on one rank.
Kept in a file of its own, collected after the parity and solver suites: it was written after the round's GPU budget was spent
(the split path itself ran on 2 and 8 GPUs through bench.py with NHSTEP_OVERLAP=1 and through the multi-GPU parity tests while the
overlap was on by default), so a surprise here must not hide the other suites under `pytest -x`."""
import numpy as np
import pytest

from test_gpu_distributed import _distributed

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("exact", [True, False], ids=["exact", "contracted"])
@pytest.mark.parametrize("size", [(32, 64, 16), (16, 32, 32)])
def test_interior_boundary_split_of_the_slab_sweep_is_bit_identical(size, exact, monkeypatch):
    """nh_dist_tendency_substep_part (interior tile rows before the halo barrier, boundary tile rows after it — the reference's
    interleave_communication_and_computation.jl:9-67) must give exactly the fields of the single launch, in both sweeps
    (exact: tile kernel, contracted: TMA plane kernel); Ny = 64 has interior tile rows in both, Ny = 32 only in the plane one."""
    steps, dt, FT = 5, 2e-3, np.float64
    whole, split = {}, {}
    _distributed(0, 1, size, FT, "RungeKutta3", steps, dt, whole, True, exact=exact)
    monkeypatch.setenv("NHSTEP_OVERLAP", "1")
    _distributed(0, 1, size, FT, "RungeKutta3", steps, dt, split, True, exact=exact)
    for n in ("u", "v", "w", "pNHS"):
        np.testing.assert_array_equal(split[n], whole[n])
