"""The reference's halo-region tests (test/test_halo_regions.jl:1-70) restated against the oracle (all of the reference's sizes,
incl. the one-cell directions) and against the CUDA library through the C ABI (GPU; the three-dimensional sizes)."""
import numpy as np
import pytest

import oracle as o
from oracle.fields import Field, fill_halo_regions

P, B = "Periodic", "Bounded"
NS = [(8, 8, 8), (8, 8, 4), (10, 7, 5), (1, 8, 8), (1, 9, 5), (8, 1, 8), (5, 1, 9), (8, 8, 1), (5, 9, 1), (1, 1, 8)]


def _checks(data, N, H):
    """data: parent array; the four equalities of halo_regions_correctly_filled (periodic x, y; no-flux z)."""
    Nx, Ny, Nz = N
    Hx, Hy, Hz = H
    I = lambda lo, hi, h: slice(lo - 1 + h, hi + h)          # 1-based inclusive range → parent slice
    ok = np.array_equal(data[I(1 - Hx, 0, Hx), I(1, Ny, Hy), I(1, Nz, Hz)], data[I(Nx - Hx + 1, Nx, Hx), I(1, Ny, Hy), I(1, Nz, Hz)])
    ok &= np.array_equal(data[I(1, Nx, Hx), I(1 - Hy, 0, Hy), I(1, Nz, Hz)], data[I(1, Nx, Hx), I(Ny - Hy + 1, Ny, Hy), I(1, Nz, Hz)])
    ok &= np.array_equal(data[I(1, Nx, Hx), I(1, Ny, Hy), I(0, 0, Hz)], data[I(1, Nx, Hx), I(1, Ny, Hy), I(1, 1, Hz)])
    ok &= np.array_equal(data[I(1, Nx, Hx), I(1, Ny, Hy), I(Nz + 1, Nz + 1, Hz)], data[I(1, Nx, Hx), I(1, Ny, Hy), I(Nz, Nz, Hz)])
    return ok


@pytest.mark.parametrize("FT", [np.float32, np.float64])
@pytest.mark.parametrize("N", NS)
def test_oracle_halo_regions_initialized_and_filled_like_the_reference(N, FT):
    rng = np.random.default_rng(0)
    g = o.RectilinearGrid(FT, size=N, extent=(10, 20, 30), halo=(1, 1, 1))
    f = Field(g, ("c", "c", "c"), name="c")
    f.set(rng.random(N).astype(FT))
    interior = np.zeros(f.data.shape, dtype=bool)
    interior[1:-1, 1:-1, 1:-1] = True
    assert np.all(f.data[~interior] == 0)                     # halo_regions_initalized_correctly
    g = o.RectilinearGrid(FT, size=N, extent=(100, 200, 300), halo=(1, 1, 1), topology=(P, P, B))
    f = Field(g, ("c", "c", "c"), name="c")
    f.set(rng.random(N).astype(FT))
    fill_halo_regions(f)
    assert _checks(f.data, N, (1, 1, 1))                      # halo_regions_correctly_filled


@pytest.mark.gpu
@pytest.mark.parametrize("FT", [np.float32, np.float64])
@pytest.mark.parametrize("N", NS[:3])
def test_library_halo_regions_initialized_and_filled_like_the_reference(N, FT):
    import oceananigans_jl_b200 as oc
    rng = np.random.default_rng(0)
    g = oc.RectilinearGrid(FT, size=N, extent=(100, 200, 300), halo=(1, 1, 1), topology=(P, P, B))
    m = oc.NonhydrostaticModel(g, advection=None, tracers=("c",))
    c = m.tracers["c"]
    assert m.halo == (1, 1, 1)
    values = rng.random(N).astype(FT)
    c.set(values)
    data = c.parent()
    assert np.array_equal(data[1:-1, 1:-1, 1:-1], values)
    interior = np.zeros(data.shape, dtype=bool)
    interior[1:-1, 1:-1, 1:-1] = True
    assert np.all(data[~interior] == 0)
    oc.fill_halo_regions(c)
    assert _checks(c.parent(), N, (1, 1, 1))


# ---- test_field.jl:262-292: parent sizes of the four field locations on the four topologies ------------------------------------

_SIZES = {  # topology → +1 in which (location, direction)
    (P, P, P): set(), (P, P, B): {("w", 2)}, (P, B, B): {("v", 1), ("w", 2)}, (B, B, B): {("u", 0), ("v", 1), ("w", 2)}}


@pytest.mark.parametrize("FT", [np.float32, np.float64])
@pytest.mark.parametrize("topo", list(_SIZES), ids=lambda t: "".join(x[0] for x in t))
def test_oracle_field_initialization_sizes(topo, FT):
    N, H = (4, 6, 8), (1, 1, 1)
    g = o.RectilinearGrid(FT, size=N, extent=(2 * np.pi, 3 * np.pi, 5 * np.pi), halo=H, topology=topo)
    locs = {"c": ("c", "c", "c"), "u": ("f", "c", "c"), "v": ("c", "f", "c"), "w": ("c", "c", "f")}
    for name, loc in locs.items():
        want = tuple(N[d] + 2 * H[d] + ((name, d) in _SIZES[topo]) for d in range(3))
        f = Field(g, loc, name=name)
        assert f.data.shape == want and f.data.dtype == FT, (name, f.data.shape, want)
