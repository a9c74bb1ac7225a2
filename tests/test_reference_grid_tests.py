"""The reference's own RectilinearGrid tests (test/test_grids.jl:8-160, 296-342, 403-446) restated against the oracle — the
exact `==` known-answer checks (halo faces, first cells, end faces, Flat sizes / halos / extents) and the stretched-grid spacing
checks — and, for the z metrics the C ABI exposes, against the CUDA library (GPU)."""
import numpy as np
import pytest

import oracle as o

P, B, F = "Periodic", "Bounded", "Flat"
FTS = [np.float32, np.float64]


@pytest.mark.parametrize("FT", FTS)
def test_regular_rectilinear_domain_size_lengths_and_halo(FT):
    # test_grids.jl:8-68
    g = o.RectilinearGrid(FT, size=(4, 6, 8), extent=(2 * np.pi, 4 * np.pi, 9 * np.pi))
    assert (g.Nx, g.Ny, g.Nz) == (4, 6, 8)
    assert np.isclose(g.Lx, 2 * np.pi) and np.isclose(g.Ly, 4 * np.pi) and np.isclose(g.Lz, 9 * np.pi)
    g = o.RectilinearGrid(FT, size=(16, 16, 16), x=(-0.5, 0.5), y=(np.pi, 3 * np.pi), z=(-6, -2))
    assert np.isclose(g.Lx, 1) and np.isclose(g.Ly, 2 * np.pi) and np.isclose(g.Lz, 4)
    Nx, Ny, Nz, Hx, Hy, Hz = 2, 3, 4, 1, 1, 1
    g = o.RectilinearGrid(FT, size=(Nx, Ny, Nz), extent=(1, 1, 1), halo=(Hx, Hy, Hz), topology=(P, B, B))
    cx, cy, cz = g.coords
    assert len(cx.C) == Nx + 2 * Hx and len(cy.C) == Ny + 2 * Hy and len(cz.C) == Nz + 2 * Hz
    assert len(cx.F) == Nx + 2 * Hx and len(cy.F) == Ny + 2 * Hy + 1 and len(cz.F) == Nz + 2 * Hz + 1
    g = o.RectilinearGrid(FT, size=(4, 6, 8), extent=(2 * np.pi, 4 * np.pi, 9 * np.pi), halo=(1, 2, 3))
    assert (g.Hx, g.Hy, g.Hz) == (1, 2, 3)


@pytest.mark.parametrize("FT", FTS)
def test_regular_rectilinear_halo_faces_first_cells_end_faces_exactly(FT):
    # test_grids.jl:70-120: `==`, not ≈
    N, H, L = 4, 1, 2.0
    d = L / N
    g = o.RectilinearGrid(FT, topology=(P, B, B), size=(N, N, N), x=(0, L), y=(0, L), z=(0, L), halo=(H, H, H))
    for c in g.coords:
        assert c.F[0] == -H * d
    assert g.coords[0].F[N + 1] == L                       # Periodic: the last face of the parent array
    assert g.coords[1].F[N + 2] == L + H * d and g.coords[2].F[N + 2] == L + H * d
    L = 4.0
    d = L / N
    g = o.RectilinearGrid(FT, size=(N, N, N), x=(0, L), y=(0, L), z=(0, L), halo=(H, H, H))
    for c in g.coords:
        assert c.C[1] == d / 2
    L = 2.0
    d = L / N
    g = o.RectilinearGrid(FT, size=(N, N, N), x=(0, L), y=(0, L), z=(0, L), halo=(1, 1, 1), topology=(P, B, B))
    assert g.coords[0].F[N + 1] == L and g.coords[1].F[N + 2] == L + d and g.coords[2].F[N + 2] == L + d


@pytest.mark.parametrize("FT", FTS)
def test_flat_sizes_halos_and_extents(FT):
    # test_grids.jl:296-342
    size = lambda **kw: o.RectilinearGrid(FT, **kw).N
    halo = lambda **kw: o.RectilinearGrid(FT, **kw).H
    ext = lambda **kw: tuple(float(L) for L in (lambda g: (g.Lx, g.Ly, g.Lz))(o.RectilinearGrid(FT, **kw)))
    assert size(topology=(F, P, P), size=(2, 3), extent=(1, 1)) == (1, 2, 3)
    assert size(topology=(P, F, B), size=(2, 3), extent=(1, 1)) == (2, 1, 3)
    assert size(topology=(P, B, F), size=(2, 3), extent=(1, 1)) == (2, 3, 1)
    assert size(topology=(P, F, F), size=2, extent=1) == (2, 1, 1)
    assert size(topology=(F, P, F), size=2, extent=1) == (1, 2, 1)
    assert size(topology=(F, F, B), size=2, extent=1) == (1, 1, 2)
    assert halo(topology=(F, P, P), size=(3, 3), extent=(1, 1)) == (0, 3, 3)
    assert halo(topology=(P, F, B), size=(3, 3), extent=(1, 1)) == (3, 0, 3)
    assert halo(topology=(P, B, F), size=(3, 3), extent=(1, 1)) == (3, 3, 0)
    assert halo(topology=(F, P, P), size=(3, 3), extent=(1, 1), halo=(2, 3)) == (0, 2, 3)
    assert halo(topology=(P, F, B), size=(3, 3), extent=(1, 1), halo=(2, 3)) == (2, 0, 3)
    assert halo(topology=(P, B, F), size=(3, 3), extent=(1, 1), halo=(2, 3)) == (2, 3, 0)
    assert halo(topology=(P, F, F), size=2, extent=1, halo=2) == (2, 0, 0)
    assert halo(topology=(F, F, B), size=2, extent=1, halo=2) == (0, 0, 2)
    assert ext(topology=(F, P, P), size=(2, 3), extent=(1, 1)) == (1, 1, 1)
    assert ext(topology=(P, F, P), size=(2, 3), extent=(1, 1)) == (1, 1, 1)
    assert ext(topology=(P, P, F), size=(2, 3), extent=(1, 1)) == (1, 1, 1)
    assert ext(topology=(P, F, F), size=2, extent=1) == (1, 1, 1)


def _tanh_faces(N, S=3):
    return lambda k: np.tanh(S * (2 * (k - 1) / N - 1)) / np.tanh(S)


@pytest.mark.parametrize("N", [16, 17])
@pytest.mark.parametrize("FT", FTS)
def test_stretched_z_spacings_as_the_reference_checks_them(FT, N):
    # test_rectilinear_grid_correct_spacings, test_grids.jl:403-446 (the z half: x regular, z tanh-spaced)
    zf = _tanh_faces(N)
    g = o.RectilinearGrid(FT, size=(N, N, N), x=(0, N), y=(0, N), z=zf, topology=(P, P, B))
    cz = g.coords[2]
    assert g.coords[0].dc == 1 and g.coords[0].df == 1
    k = np.arange(1, N + 2)
    zc = lambda q: (zf(q) + zf(q + 1)) / 2
    assert np.allclose(cz.F.r(1, N + 1), zf(k))
    assert np.allclose(cz.C.r(1, N), zc(k[:-1]))
    assert np.allclose(cz.dc.r(1, N), zf(k[:-1] + 1) - zf(k[:-1]))
    # Δzᵃᵃᶠ[1] involves a halo point, which the user-supplied faces do not determine: 2 … N
    assert np.allclose(cz.df.r(2, N), zc(np.arange(2, N + 1)) - zc(np.arange(1, N)))
    assert np.isclose(cz.dc.r(1, N).min(), (zf(k[:-1] + 1) - zf(k[:-1])).min())


@pytest.mark.gpu
@pytest.mark.parametrize("N", [16, 17])
@pytest.mark.parametrize("FT", FTS)
def test_library_stretched_z_spacings_equal_the_oracle_bit_for_bit(FT, N):
    import oceananigans_jl_b200 as oc
    zf = _tanh_faces(N)
    g1 = oc.RectilinearGrid(FT, size=(N, N, N), x=(0, N), y=(0, N), z=zf, topology=(P, P, B))
    g2 = o.RectilinearGrid(FT, size=(N, N, N), x=(0, N), y=(0, N), z=zf, topology=(P, P, B))
    dx, dy, dzc, dzf = g1.spacings()
    cz = g2.coords[2]
    H = g2.Hz
    assert dx == g2.coords[0].dc and dy == g2.coords[1].dc
    # the ranges the reference's test checks (Δzᶠ[1] involves a halo point); the arrays start at k = 1 - H
    assert np.array_equal(dzc[H: H + N], cz.dc.r(1, N))
    assert np.array_equal(dzf[H + 1: H + N], cz.df.r(2, N))
