"""The reference's NonhydrostaticModel constructor / set! tests (test/test_nonhydrostatic_models.jl) restated over the oracle
(CPU).  Left out: what is out of this path's scope — forcing / background_fields type errors, ScalarBiharmonicDiffusivity,
schemes of order > 5."""
import numpy as np
import pytest

import oracle as o

P, B = "Periodic", "Bounded"
FTS = [np.float32, np.float64]


@pytest.mark.parametrize("FT", FTS)
@pytest.mark.parametrize("topo", [(P, P, P), (P, P, B), (P, B, B), (B, B, B)], ids=lambda t: "".join(x[0] for x in t))
def test_model_construction(topo, FT):
    """test_nonhydrostatic_models.jl:15-32."""
    grid = o.RectilinearGrid(FT, topology=topo, size=(16, 16, 2), extent=(1, 2, 3))
    m = o.NonhydrostaticModel(grid)
    assert m.grid.FT is FT and m.clock.iteration == 0


def _halo(m):
    return (m.grid.Hx, m.grid.Hy, m.grid.Hz)


def test_adjustment_of_halos_in_the_constructor():
    """test_nonhydrostatic_models.jl:34-62."""
    minimal = lambda: o.RectilinearGrid(np.float64, size=(4, 4, 4), extent=(1, 2, 3), halo=(1, 1, 1))
    funny = lambda: o.RectilinearGrid(np.float64, size=(4, 4, 4), extent=(1, 2, 3), halo=(1, 3, 4))
    FT = np.float64
    assert _halo(o.NonhydrostaticModel(minimal())) == (1, 1, 1)
    assert _halo(o.NonhydrostaticModel(funny())) == (1, 3, 4)
    for scheme in (o.Centered(FT, 4), o.UpwindBiased(FT, 3)):
        assert _halo(o.NonhydrostaticModel(minimal(), advection=scheme)) == (2, 2, 2)
        assert _halo(o.NonhydrostaticModel(funny(), advection=scheme)) == (2, 3, 4)
    for scheme in (o.WENO(FT, 5), o.UpwindBiased(FT, 5)):
        assert _halo(o.NonhydrostaticModel(minimal(), advection=scheme)) == (3, 3, 3)
        assert _halo(o.NonhydrostaticModel(funny(), advection=scheme)) == (3, 3, 4)


def test_adjustment_of_advection_schemes_in_the_constructor():
    """test_nonhydrostatic_models.jl:64-72: on a 4 × 2 × 4 grid WENO() keeps its buffer 3 in x and z and drops to 2 in y."""
    small = o.RectilinearGrid(np.float64, size=(4, 2, 4), extent=(1, 2, 3), halo=(1, 1, 1))
    adv = o.advection.adapt_advection_order(o.WENO(np.float64, 5), small)
    assert [s.B for s in adv] == [3, 2, 3] and all(s.kind == "WENO" for s in adv)
    adv = o.advection.adapt_advection_order(o.Centered(np.float64, 4), small)
    assert [s.B for s in adv] == [2, 2, 2]


def test_halo_is_inflated_per_direction_when_the_adapted_orders_differ():
    """test_nonhydrostatic_models.jl:64-72 with inflate_halo_size (automatic_halo_sizing.jl:71-83): on the 4 × 2 × 4 grid WENO() needs
    (3, 2, 3) — each direction its own adapted scheme's buffer — and Hy = 2 ≤ Ny = 2 keeps the Periodic y direction valid."""
    small = o.RectilinearGrid(np.float64, size=(4, 2, 4), extent=(1, 2, 3), halo=(1, 1, 1))
    m = o.NonhydrostaticModel(small, advection=o.WENO(np.float64, 5))
    assert _halo(m) == (3, 2, 3)
    assert m.velocities[0].data.shape == (4 + 6, 2 + 4, 4 + 6)
    o.time_step(m, 1e-3)
    funny = o.RectilinearGrid(np.float64, size=(4, 2, 4), extent=(1, 2, 3), halo=(1, 2, 4), topology=(P, B, B))
    assert _halo(o.NonhydrostaticModel(funny, advection=o.Centered(np.float64, 4))) == (2, 2, 4)


def test_model_construction_with_single_tracer_and_nothing_tracer():
    """test_nonhydrostatic_models.jl:95-106."""
    grid = o.RectilinearGrid(np.float64, size=(1, 1, 1), extent=(1, 2, 3))
    assert list(o.NonhydrostaticModel(grid, tracers="c", buoyancy=None).tracers) == ["c"]
    assert list(o.NonhydrostaticModel(grid, tracers=(), buoyancy=None).tracers) == []


@pytest.mark.parametrize("FT", FTS)
def test_setting_model_fields(FT):
    """test_nonhydrostatic_models.jl:108-187 (set! with arrays; functions are evaluated at the field's nodes as set! does)."""
    N, L = (4, 4, 4), (2 * np.pi, 3 * np.pi, 5 * np.pi)
    grid = o.RectilinearGrid(FT, size=N, extent=L)
    m = o.NonhydrostaticModel(grid, buoyancy=o.SeawaterBuoyancy(), tracers=("T", "S"))
    u, v, w = m.velocities
    T, S = m.tracers["T"], m.tracers["S"]
    rng = np.random.default_rng(0)
    T0 = rng.random(N).astype(FT)
    o.set_model(m, enforce_incompressibility=False, T=T0)
    assert np.array_equal(T.interior, T0)

    funcs = dict(u=lambda x, y, z: 1 + x + y + z, v=lambda x, y, z: 2 + np.sin(x * y * z), w=lambda x, y, z: 3 + y * z,
                 T=lambda x, y, z: 4 + np.tanh(x + y - z), S=lambda x, y, z: 5 + 0 * x * y * z)

    def answer(name):
        f = dict(u=u, v=v, w=w, T=T, S=S)[name]
        xs = [(c.F if f.loc[d] == "f" else c.C).r(1, f.size[d]).astype(np.float64) for d, c in enumerate(grid.coords)]
        return funcs[name](xs[0][:, None, None], xs[1][None, :, None], xs[2][None, None, :]).astype(FT)

    o.set_model(m, enforce_incompressibility=False, **{n: answer(n) for n in funcs})
    Nx, Ny, Nz = N
    tol = dict(rtol=float(np.sqrt(np.finfo(FT).eps)), atol=0)                 # Julia's ≈
    assert np.allclose(u.interior, answer("u"), **tol) and np.allclose(v.interior, answer("v"), **tol)
    assert np.allclose(w.interior[:, :, 1:Nz], answer("w")[:, :, 1:Nz], **tol)
    assert np.allclose(T.interior, answer("T"), **tol) and np.allclose(S.interior, answer("S"), **tol)
    # set! fills the halos: periodic in x and y, free slip at bottom and top
    at = lambda i, j, k: u.at((i, i), (j, j), (k, k))[0, 0, 0]
    assert at(1, 1, 1) == at(Nx + 1, 1, 1) and at(1, 1, 1) == at(1, Ny + 1, 1)
    assert np.array_equal(u.at((1, Nx), (1, Ny), (1, 1)), u.at((1, Nx), (1, Ny), (0, 0)))
    assert np.array_equal(u.at((1, Nx), (1, Ny), (Nz, Nz)), u.at((1, Nx), (1, Ny), (Nz + 1, Nz + 1)))
    # enforce_incompressibility: a uniform w = 1 between two walls is projected out
    z = np.zeros(N, dtype=FT)
    o.set_model(m, u=z, v=z, w=np.ones((Nx, Ny, Nz + 1), dtype=FT), T=z, S=z)
    assert np.all(np.abs(w.interior) < 10 * np.finfo(FT).eps)


# ---- the same tests against the library through the C ABI ---------------------------------------------------------------------

@pytest.mark.gpu
def test_library_adjustment_of_halos_in_the_constructor():
    """test_nonhydrostatic_models.jl:34-72 on nh_model_create (model.halo = the inflated halo of nh_model_grid_info)."""
    import oceananigans_jl_b200 as oc
    FT = np.float64
    minimal = lambda: oc.RectilinearGrid(FT, size=(4, 4, 4), extent=(1, 2, 3), halo=(1, 1, 1))
    funny = lambda: oc.RectilinearGrid(FT, size=(4, 4, 4), extent=(1, 2, 3), halo=(1, 3, 4))
    assert oc.NonhydrostaticModel(minimal()).halo == (1, 1, 1)
    assert oc.NonhydrostaticModel(funny()).halo == (1, 3, 4)
    for scheme in (oc.Centered(FT, 4), oc.UpwindBiased(FT, 3)):
        assert oc.NonhydrostaticModel(minimal(), advection=scheme).halo == (2, 2, 2)
        assert oc.NonhydrostaticModel(funny(), advection=scheme).halo == (2, 3, 4)
    for scheme in (oc.WENO(FT, 5), oc.UpwindBiased(FT, 5)):
        assert oc.NonhydrostaticModel(minimal(), advection=scheme).halo == (3, 3, 3)
        assert oc.NonhydrostaticModel(funny(), advection=scheme).halo == (3, 3, 4)
    small = oc.RectilinearGrid(FT, size=(4, 2, 4), extent=(1, 2, 3), halo=(1, 1, 1))
    m = oc.NonhydrostaticModel(small, advection=oc.WENO(FT, 5))
    assert m.halo == (3, 2, 3) and m.velocities.u.parent().shape == (10, 6, 10)
    oc.time_step(m, 1e-3)


@pytest.mark.gpu
@pytest.mark.parametrize("FT", FTS)
@pytest.mark.parametrize("topo", [(P, P, P), (P, P, B), (P, B, B), (B, B, B)], ids=lambda t: "".join(x[0] for x in t))
def test_library_model_construction(topo, FT):
    import oceananigans_jl_b200 as oc
    m = oc.NonhydrostaticModel(oc.RectilinearGrid(FT, topology=topo, size=(16, 16, 2), extent=(1, 2, 3)))
    assert m.clock.iteration == 0
    grid = oc.RectilinearGrid(FT, size=(1, 1, 1), extent=(1, 2, 3))
    assert list(oc.NonhydrostaticModel(grid, tracers="c", buoyancy=None).tracers) == ["c"]
    assert list(oc.NonhydrostaticModel(grid, tracers=(), buoyancy=None).tracers) == []


@pytest.mark.gpu
@pytest.mark.parametrize("FT", FTS)
def test_library_setting_model_fields(FT):
    """test_nonhydrostatic_models.jl:108-187: set! with an array, with functions of (x, y, z), the halos set! leaves, and
    enforce_incompressibility."""
    import oceananigans_jl_b200 as oc
    N, L = (4, 4, 4), (2 * np.pi, 3 * np.pi, 5 * np.pi)
    grid = oc.RectilinearGrid(FT, size=N, extent=L)
    m = oc.NonhydrostaticModel(grid, buoyancy=oc.SeawaterBuoyancy(), tracers=("T", "S"))
    u, v, w = (m.velocities[n] for n in "uvw")
    T, S = m.tracers["T"], m.tracers["S"]
    rng = np.random.default_rng(0)
    T0 = rng.random(N).astype(FT)
    oc.set_model(m, enforce_incompressibility=False, T=T0)
    assert np.array_equal(T.interior(), T0)
    funcs = dict(u=lambda x, y, z: 1 + x + y + z, v=lambda x, y, z: 2 + np.sin(x * y * z), w=lambda x, y, z: 3 + y * z,
                 T=lambda x, y, z: 4 + np.tanh(x + y - z), S=lambda x, y, z: 5.0)
    oc.set_model(m, enforce_incompressibility=False, **funcs)
    og = o.RectilinearGrid(FT, size=N, extent=L)

    def answer(name, f):
        xs = [(c.F if f.loc[d] == "f" else c.C).r(1, f.size[d]).astype(np.float64) for d, c in enumerate(og.coords)]
        return (funcs[name](xs[0][:, None, None], xs[1][None, :, None], xs[2][None, None, :]) * np.ones(f.size)).astype(FT)

    Nx, Ny, Nz = N
    tol = dict(rtol=float(np.sqrt(np.finfo(FT).eps)), atol=0)
    assert np.allclose(u.interior(), answer("u", u), **tol) and np.allclose(v.interior(), answer("v", v), **tol)
    assert np.allclose(w.interior()[:, :, 1:Nz], answer("w", w)[:, :, 1:Nz], **tol)
    assert np.allclose(T.interior(), answer("T", T), **tol) and np.allclose(S.interior(), answer("S", S), **tol)
    up, H = u.parent(), m.halo
    at = lambda i, j, k: up[i - 1 + H[0], j - 1 + H[1], k - 1 + H[2]]
    assert at(1, 1, 1) == at(Nx + 1, 1, 1) and at(1, 1, 1) == at(1, Ny + 1, 1)
    assert np.array_equal(up[H[0]:H[0] + Nx, H[1]:H[1] + Ny, H[2]], up[H[0]:H[0] + Nx, H[1]:H[1] + Ny, H[2] - 1])
    assert np.array_equal(up[H[0]:H[0] + Nx, H[1]:H[1] + Ny, H[2] + Nz - 1], up[H[0]:H[0] + Nx, H[1]:H[1] + Ny, H[2] + Nz])
    z = np.zeros(N, dtype=FT)
    oc.set_model(m, u=z, v=z, w=np.ones((Nx, Ny, Nz + 1), dtype=FT), T=z, S=z)
    assert np.all(np.abs(w.interior()) < 10 * np.finfo(FT).eps)
