"""Pins the oracle against the reference's own property / analytic / known-answer tests.

Each test restates one reference test and runs it on ``oracle/``:
  test/test_poisson_solvers.jl:58-106 + test/dependencies_for_poisson_solvers.jl:13-141
  test/test_poisson_solvers_stretched_grids.jl:12-50
  test/test_batched_tridiagonal_solver.jl:7-42
  test/test_time_stepping.jl:89-115,125-158,166-200,433-461
  test/test_dynamics.jl:215-260
"""
import itertools

import numpy as np
import pytest

import oracle as o
from oracle.operators import divergence_ccc, laplacian_ccc, interior_region
from oracle.fields import Field, fill_halo_regions

P, B, F = o.Periodic, o.Bounded, o.Flat


def isapprox(a, b, rtol=None):
    """Julia's isapprox for arrays: norm(a-b) <= rtol*max(norm(a), norm(b)), rtol = sqrt(eps)."""
    a, b = np.asarray(a), np.asarray(b)
    if rtol is None:
        rtol = np.sqrt(np.finfo(a.dtype).eps)
    return np.linalg.norm((a - b).ravel()) <= rtol * max(np.linalg.norm(a.ravel()), np.linalg.norm(b.ravel()))


def random_velocities(grid, rng):
    U = [Field(grid, loc, name=n) for n, loc in zip("uvw", (("f", "c", "c"), ("c", "f", "c"), ("c", "c", "f")))]
    for f in U:
        f.set(rng.random(f.size))
        fill_halo_regions(f)
    return U


def poisson_residual_ok(grid, rng):
    """divergence_free_poisson_solution (dependencies_for_poisson_solvers.jl:79-97)."""
    U = random_velocities(grid, rng)
    R = interior_region(grid)
    rhs = np.array(divergence_ccc(grid, *U, R))
    solver = o.FFTBasedPoissonSolver(grid)
    phi = Field(grid, ("c", "c", "c"), prognostic=False)
    solver.storage[...] = rhs / 1            # solve_for_pressure!(ϕ, solver, Δt=1, U)
    solver.solve(phi)
    fill_halo_regions(phi)
    lap = np.array(laplacian_ccc(grid, phi, R))
    return isapprox(lap, rhs.astype(grid.FT))


def make_grid(FT, topo, size3):
    size = tuple(n for n, t in zip(size3, topo) if t != F)
    kw = {}
    for name, t in zip("xyz", topo):
        if t != F:
            kw[name] = (0, 1)
    return o.RectilinearGrid(FT, size=size, topology=topo, **kw)


TOPOS3 = list(itertools.product((P, B), repeat=3))
TOPOS2 = [(F, P, P), (F, P, B), (F, B, B), (P, F, P), (P, F, B), (B, F, B), (P, P, F), (P, B, F), (B, B, F)]


@pytest.mark.parametrize("topo", TOPOS3)
@pytest.mark.parametrize("N", [7, 16])
def test_poisson_divergence_free_3d(topo, N):
    rng = np.random.default_rng(N)
    assert poisson_residual_ok(make_grid(np.float64, topo, (N, N, N)), rng)
    for size in [(1, N, N), (N, 1, N), (N, N, 1)]:
        assert poisson_residual_ok(make_grid(np.float64, topo, size), rng)


@pytest.mark.parametrize("topo", TOPOS2)
def test_poisson_divergence_free_2d_flat(topo):
    rng = np.random.default_rng(3)
    for N in (7, 16):
        assert poisson_residual_ok(make_grid(np.float64, topo, (N, N, N)), rng)


@pytest.mark.parametrize("topo", TOPOS3)
def test_poisson_rectangular_sizes(topo):
    rng = np.random.default_rng(5)
    for size in itertools.product((11, 16), repeat=3):
        assert poisson_residual_ok(make_grid(np.float64, topo, size), rng)


def test_poisson_float32():
    rng = np.random.default_rng(7)
    assert poisson_residual_ok(make_grid(np.float32, (P, B, B), (16, 16, 16)), rng)
    assert poisson_residual_ok(make_grid(np.float32, (B, P, P), (7, 11, 16)), rng)


def _analytic_error(N, topo, mode):
    two_pi = 2 * np.pi
    grid = o.RectilinearGrid(np.float64, size=(N, N, N), x=(0, two_pi), y=(0, two_pi), z=(0, two_pi), topology=topo)
    solver = o.FFTBasedPoissonSolver(grid)
    psi = {B: lambda x: np.cos(mode * x / 2), P: lambda x: np.cos(mode * x)}
    k2 = {B: (mode / 2) ** 2, P: mode ** 2}
    c = [grid.coords[d].C.r(1, N) for d in range(3)]
    Psi = (psi[topo[0]](c[0])[:, None, None] * psi[topo[1]](c[1])[None, :, None] * psi[topo[2]](c[2])[None, None, :])
    f = -(k2[topo[0]] + k2[topo[1]] + k2[topo[2]]) * Psi
    phi = solver.solve(None, f)
    return np.mean(np.abs(phi - Psi))


@pytest.mark.parametrize("topo", TOPOS3)
def test_poisson_convergence(topo):
    for N1, N2 in ((2 ** 6, 2 ** 7), (67, 131)):
        for mode in (1, 2):
            e1, e2 = _analytic_error(N1, topo, mode), _analytic_error(N2, topo, mode)
            rate = np.log(e1 / e2) / np.log(N2 / N1)
            assert abs(rate - 2) <= 5e-3 * 2, (topo, N1, N2, mode, rate)


# -- batched tridiagonal: test_batched_tridiagonal_solver.jl:7-42 -------------------------------
@pytest.mark.parametrize("direction", [0, 1, 2])
def test_batched_tridiagonal_vs_dense(direction):
    rng = np.random.default_rng(11)
    N = 8
    size = [3, 4, 5]
    size[direction] = N
    grid = o.RectilinearGrid(np.float64, size=tuple(size), extent=(1, 1, 1), topology=(P, P, B))
    a, c = rng.random(N - 1), rng.random(N - 1)
    b = 3 + rng.random(N)                      # diagonally dominant
    f = rng.random(size)
    solver = o.BatchedTridiagonalSolver(grid, a, b, c, direction=direction)
    phi = np.zeros(size)
    solver.solve(phi, f)
    M = np.diag(b) + np.diag(a, -1) + np.diag(c, 1)
    expect = np.moveaxis(np.linalg.solve(M, np.moveaxis(f, direction, 0).reshape(N, -1)).reshape(
        [N] + [s for d, s in enumerate(size) if d != direction]), 0, direction)
    assert np.allclose(phi, expect, rtol=1e-12, atol=0)


# -- Fourier-tridiagonal: test_poisson_solvers_stretched_grids.jl:12-50 -------------------------
FACES = [1, 2, 4, 7, 11, 16, 22, 29, 37]


def _w_from_continuity(grid, U):
    """compute_w_from_continuity!: w[k+1] = w[k] - Δz·(δx(Ax u)+δy(Ay v))/Az, w[1] = 0."""
    u, v, w = U
    from oracle.operators import area, diff, delta
    Nz = grid.N[2]
    w.data[...] = 0
    for k in range(1, Nz + (1 if grid.topology[2] == B else 0)):
        R = [(1, grid.N[0]), (1, grid.N[1]), (k, k)]
        flux = (diff(grid, lambda *r: area(grid, 0, ("f", "c", "c"), list(r)) * u.at(*r), 0, "c", R)
                + diff(grid, lambda *r: area(grid, 1, ("c", "f", "c"), list(r)) * v.at(*r), 1, "c", R))
        Rk1 = [R[0], R[1], (k + 1, k + 1)]
        w.at(*Rk1)[...] = w.at(*R) - flux / area(grid, 2, ("c", "c", "c"), R)
    return w


@pytest.mark.parametrize("topo", [(P, P, B), (P, B, B), (B, P, B), (B, B, B), (F, P, B), (P, F, B), (F, B, B), (B, F, B)])
def test_fourier_tridiagonal_stretched_z(topo):
    rng = np.random.default_rng(13)
    for N1, N2 in ((8, 8), (7, 11)):
        size3 = (N1, N2, len(FACES) - 1)
        size = tuple(n for n, t in zip(size3, topo) if t != F)
        kw = {n: (0, 1) for n, t in zip("xy", topo[:2]) if t != F}
        grid = o.RectilinearGrid(np.float64, size=size, topology=topo, z=np.array(FACES, float), **kw)
        U = random_velocities(grid, rng)
        _w_from_continuity(grid, U)
        fill_halo_regions(U[2])
        R = interior_region(grid)
        rhs = np.array(divergence_ccc(grid, *U, R))
        solver = o.FourierTridiagonalPoissonSolver(grid)
        phi = Field(grid, ("c", "c", "c"), prognostic=False)
        solver.solve(phi, rhs)
        fill_halo_regions(phi)
        lap = np.array(laplacian_ccc(grid, phi, R))
        assert isapprox(lap, rhs), (topo, N1, N2)


# -- time stepping ---------------------------------------------------------------------------------
@pytest.mark.parametrize("FT", [np.float64, np.float32])
def test_first_ab2_step_fields_stay_zero(FT):
    """run_first_AB2_time_step_tests minus the user forcing (forcings are out of scope): all fields stay 0."""
    grid = o.RectilinearGrid(FT, size=(13, 17, 19), extent=(1, 2, 3))
    m = o.NonhydrostaticModel(grid, timestepper="QuasiAdamsBashforth2", buoyancy=o.SeawaterBuoyancy(), tracers=("T", "S"))
    for G in m.Gn.values():
        assert np.all(G.interior == 0)
    o.time_step(m, 1, euler=True)
    for f in m.prognostic_fields():
        assert np.all(f.interior == 0)
    assert m.clock.iteration == 1 and m.clock.time == 1.0


def _tanh_faces(Nz, S=1.3):
    k = np.arange(1, Nz + 2)
    return np.tanh(S * (2 * (k - 1) / Nz - 1)) / np.tanh(S)


@pytest.mark.parametrize("timestepper", ["QuasiAdamsBashforth2", "RungeKutta3"])
@pytest.mark.parametrize("gridkind", ["regular", "tanh", "regular_vs"])
@pytest.mark.parametrize("FT", [np.float64, np.float32])
def test_incompressible_in_time(FT, gridkind, timestepper):
    N = 32
    z = {"regular": (-1, 1), "tanh": _tanh_faces(N), "regular_vs": np.linspace(0, 1, N + 1)}[gridkind]
    grid = o.RectilinearGrid(FT, size=(N, N, N), x=(0, 1), y=(0, 1), z=z)
    m = o.NonhydrostaticModel(grid, timestepper=timestepper, buoyancy=o.SeawaterBuoyancy(), tracers=("T", "S"))
    m.tracers["T"].interior[7:24, 7:24, 7:24] += FT(0.01)
    from oracle.model import update_state
    update_state(m)
    for n in range(10):                              # reference runs 1, 10 and 100; 1 and 10 here (100 in the slow suite)
        o.time_step(m, 0.05)
        if n == 0:
            u, v, w = m.velocities
            assert np.abs(divergence_ccc(m.grid, u, v, w, interior_region(m.grid))).max() < 5e-8
    u, v, w = m.velocities
    assert np.abs(w.interior).max() > 0
    assert np.abs(divergence_ccc(m.grid, u, v, w, interior_region(m.grid))).max() < 5e-8


@pytest.mark.parametrize("FT", [np.float64, np.float32])
def test_tracer_conserved_in_channel(FT):
    """tracer_conserved_in_channel with one isotropic ScalarDiffusivity (the anisotropic pair is out of scope)."""
    Nx, Ny, Nz = 16, 32, 16
    grid = o.RectilinearGrid(FT, size=(Nx, Ny, Nz), extent=(160e3, 320e3, 1024), topology=(P, B, B))
    m = o.NonhydrostaticModel(grid, closure=o.ScalarDiffusivity(nu=1e-2, kappa=1e-2),
                              buoyancy=o.SeawaterBuoyancy(), tracers=("T", "S"))
    rng = np.random.default_rng(17)
    y = grid.coords[1].C.r(1, Ny)[None, :, None]
    z = grid.coords[2].C.r(1, Nz)[None, None, :]
    o.set_model(m, T=10 + 1e-4 * y + 5e-3 * z + 1e-4 * rng.random((Nx, Ny, Nz)))
    T0 = np.mean(m.tracers["T"].interior.astype(np.float64))
    for _ in range(10):
        o.time_step(m, 600)
    T1 = np.mean(m.tracers["T"].interior.astype(np.float64))
    assert abs(T1 - T0) <= Nx * Ny * Nz * np.finfo(FT).eps


@pytest.mark.parametrize("timestepper", ["QuasiAdamsBashforth2", "RungeKutta3"])
def test_taylor_green_vortex(timestepper):
    N, Nt, nu = 64, 10, 1
    dx = 1 / N
    dt = (1 / (10 * np.pi)) * dx ** 2 / nu
    grid = o.RectilinearGrid(np.float64, size=(N, N, 2), extent=(1, 1, 1))
    m = o.NonhydrostaticModel(grid, timestepper=timestepper, closure=o.ScalarDiffusivity(nu=1))
    ua = lambda y, t: -np.sin(2 * np.pi * y) * np.exp(-4 * np.pi ** 2 * nu * t)
    va = lambda x, t: np.sin(2 * np.pi * x) * np.exp(-4 * np.pi ** 2 * nu * t)
    xF, xC = grid.coords[0].F.r(1, N), grid.coords[0].C.r(1, N)
    yF, yC = grid.coords[1].F.r(1, N), grid.coords[1].C.r(1, N)
    u0 = np.broadcast_to(ua(yC, 0)[None, :, None], (N, N, 2))
    v0 = np.broadcast_to(va(xC, 0)[:, None, None], (N, N, 2))
    o.set_model(m, u=u0, v=v0)
    for _ in range(Nt):
        o.time_step(m, dt)
    t = m.clock.time
    ue = np.broadcast_to(ua(yC, t)[None, :, None], (N, N, 2))
    ve = np.broadcast_to(va(xC, t)[:, None, None], (N, N, 2))
    assert np.max(np.abs((m.velocities[0].interior - ue) / ue)) < 5e-6
    assert np.max(np.abs((m.velocities[1].interior - ve) / ve)) < 5e-6


# ---------------------------------------------------------------------------------------------
# Flux / Value / Gradient boundary conditions — test/test_boundary_conditions_integration.jl:28-54,384-433
# ---------------------------------------------------------------------------------------------

_BUDGET = ([((P, B, B), n, side) for n in ("u", "c") for side in ("north", "south", "top", "bottom")]
           + [((B, P, B), n, side) for n in ("v", "c") for side in ("east", "west", "top", "bottom")]
           + [((B, B, P), n, side) for n in ("w", "c") for side in ("east", "west", "north", "south")])


@pytest.mark.parametrize("topo,name,side", _BUDGET, ids=[f"{''.join(x[0] for x in t)}-{n}-{s}" for t, n, s in _BUDGET])
def test_nonhydrostatic_flux_budget(topo, name, side):
    """test_nonhydrostatic_flux_budget: a constant boundary flux π into a field at rest raises its mean by flux·t/L."""
    FT = np.float64
    Lx, Ly, Lz = 0.3, 0.4, 0.5
    L = {"west": Lx, "east": Lx, "south": Ly, "north": Ly, "bottom": Lz, "top": Lz}[side]
    flux = FT(np.pi)
    direction = 1 if side in ("west", "south", "bottom") else -1
    g = o.RectilinearGrid(FT, size=(2, 2, 2), x=(0, Lx), y=(0, Ly), z=(0, Lz), topology=topo)
    m = o.NonhydrostaticModel(g, tracers=("c",), boundary_conditions={name: {side: ("flux", flux * direction)}})
    o.time_step(m, 1.0)
    f = dict(zip(m.prognostic_names(), m.prognostic_fields()))[name]
    assert np.isclose(f.interior.mean(), flux * m.clock.time / L, rtol=np.sqrt(np.finfo(FT).eps), atol=0)


@pytest.mark.parametrize("kind", ["value", "gradient", "flux"])
@pytest.mark.parametrize("side", ["east", "south", "top"])
def test_array_boundary_condition_time_steps(kind, side):
    """test_boundary_condition(..., array_bc) of test_boundary_conditions_integration.jl:10-28,355-358: a 1×1 array condition
    on T of a (1, 1, 1) Bounded box steps (Δt = 1e-16) without error; utils_for_runtests.jl:218."""
    g = o.RectilinearGrid(np.float64, size=(1, 1, 1), extent=(1, np.pi, 42), topology=(B, B, B))
    arr = np.random.default_rng(0).random((1, 1))
    m = o.NonhydrostaticModel(g, tracers=("T", "S"), buoyancy=o.SeawaterBuoyancy(), boundary_conditions={"T": {side: (kind, arr)}})
    o.time_step(m, 1e-16)
    assert np.all(np.isfinite(m.tracers["T"].interior))


def _wall_G(kind, cb, ct):
    from oracle.model import update_state
    g = o.RectilinearGrid(np.float64, size=(4, 3, 5), extent=(1, 1, 2), topology=(P, P, B))
    m = o.NonhydrostaticModel(g, tracers=("c",), closure=o.ScalarDiffusivity(1e-2, 1e-2),
                              boundary_conditions={"c": {"bottom": (kind, cb), "top": (kind, ct)}})
    m.tracers["c"].set(np.broadcast_to(np.linspace(0.1, 0.9, 5), (4, 3, 5)).copy())
    update_state(m)
    return m.Gn["c"].interior.copy()


@pytest.mark.parametrize("kind", ["value", "gradient", "flux"])
def test_array_boundary_condition_is_the_number_condition_pointwise(kind):
    """getbc(bc::BC{<:AbstractArray}, i, j) = condition[i, j] (boundary_condition.jl:118): a wall array acts at each wall
    point like the number it holds there — a constant array reproduces the number condition bit for bit and a two-valued
    array reproduces, column by column, the tendencies of the two number-condition models (horizontally uniform state)."""
    lo = np.full((4, 3), 0.7); lo[2:, :] = -0.2
    hi = np.full((4, 3), -0.3); hi[2:, :] = 0.5
    G1, G2 = _wall_G(kind, 0.7, -0.3), _wall_G(kind, -0.2, 0.5)
    np.testing.assert_array_equal(_wall_G(kind, np.full((4, 3), 0.7), np.full((4, 3), -0.3)), G1)
    Gm = _wall_G(kind, lo, hi)
    np.testing.assert_array_equal(Gm[:2], G1[:2])
    np.testing.assert_array_equal(Gm[2:], G2[2:])
    assert not np.array_equal(G1, G2)


def test_array_flux_budget():
    """The budget of test_nonhydrostatic_flux_budget with a non-uniform wall array: the mean rises by mean(flux)·t/L and
    each bottom cell by its own flux·t/Δz."""
    FT = np.float64
    q = np.random.default_rng(1).uniform(0.5, 1.5, (4, 3))
    g = o.RectilinearGrid(FT, size=(4, 3, 2), x=(0, 0.3), y=(0, 0.4), z=(0, 0.5), topology=(P, P, B))
    m = o.NonhydrostaticModel(g, tracers=("c",), boundary_conditions={"c": {"bottom": ("flux", q)}})
    o.time_step(m, 1.0)
    assert np.isclose(m.tracers["c"].interior.mean(), q.mean() * m.clock.time / 0.5, rtol=1e-7, atol=0)
    np.testing.assert_allclose(m.tracers["c"].interior[:, :, 0], q * m.clock.time / 0.25, rtol=1e-12)


def test_array_boundary_condition_shape_is_checked():
    g = o.RectilinearGrid(np.float64, size=(4, 3, 2), extent=(1, 1, 1), topology=(P, P, B))
    with pytest.raises(ValueError, match="shape"):
        o.NonhydrostaticModel(g, tracers=("c",), boundary_conditions={"c": {"top": ("flux", np.zeros((3, 4)))}})


@pytest.mark.parametrize("kind", ["value", "gradient", "flux"])
def test_value_gradient_halo_definitions(kind):
    """fill_halo_regions_value_gradient.jl:10-15: the wall value is the mean of halo and first interior cell; the wall
    gradient is their difference over Δ; Flux leaves the no-flux mirror (the flux enters the tendency instead)."""
    FT = np.float64
    g = o.RectilinearGrid(FT, size=(4, 4, 4), extent=(1, 1, 2), topology=(P, P, B))
    m = o.NonhydrostaticModel(g, tracers=("c",), boundary_conditions={"c": {"bottom": (kind, 0.7), "top": (kind, -0.3)}})
    rng = np.random.default_rng(3)
    c = m.tracers["c"]
    c.set(rng.uniform(-1, 1, (4, 4, 4)))
    fill_halo_regions(c)
    lo_h, lo_i = c.at((1, 4), (1, 4), (0, 0)), c.at((1, 4), (1, 4), (1, 1))
    hi_h, hi_i = c.at((1, 4), (1, 4), (5, 5)), c.at((1, 4), (1, 4), (4, 4))
    dz = 0.5
    if kind == "value":
        np.testing.assert_allclose((lo_h + lo_i) / 2, 0.7, rtol=1e-14)
        np.testing.assert_allclose((hi_h + hi_i) / 2, -0.3, rtol=1e-14)
    elif kind == "gradient":
        np.testing.assert_allclose((lo_i - lo_h) / dz, 0.7, rtol=1e-13)
        np.testing.assert_allclose((hi_h - hi_i) / dz, -0.3, rtol=1e-13)
    else:
        np.testing.assert_array_equal(lo_h, lo_i)
        np.testing.assert_array_equal(hi_h, hi_i)


def test_diffusive_flux_through_a_gradient_wall_matches_the_budget():
    """fluxes_with_diffusivity_boundary_conditions_are_correct (test_boundary_conditions_integration.jl:56-108) with a
    constant κ instead of the AMD closure: b = z·bz with a bottom Gradient(bz) condition loses κ·bz·t/Lz of its mean
    through the (insulated by default) top, to 1e-6."""
    FT = np.float64
    Lz, kappa, bz = 1.0, FT(np.exp(-3)), FT(np.pi)
    g = o.RectilinearGrid(FT, size=(16, 16, 16), extent=(1, 1, Lz), topology=(P, P, B))
    m = o.NonhydrostaticModel(g, tracers=("b",), buoyancy=o.BuoyancyTracer(), closure=o.ScalarDiffusivity(kappa, kappa),
                              timestepper="QuasiAdamsBashforth2", boundary_conditions={"b": {"bottom": ("gradient", bz)}})
    zc = (np.arange(16) + 0.5) / 16 * Lz - Lz
    o.set_model(m, b=np.broadcast_to(zc * bz, (16, 16, 16)))
    b = m.tracers["b"]
    mean0 = b.interior.mean()
    dt = 1e-6 * Lz ** 2 / kappa
    for n in range(10):
        o.time_step(m, dt, euler=(n == 0))
    assert abs((b.interior.mean() - mean0) - (-kappa * bz) * m.clock.time / Lz) < 1e-6


# ---------------------------------------------------------------------------------------------
# Smagorinsky closures — analytic eddy viscosity (the reference only smoke-tests these closures:
# test_time_stepping.jl:244-257, test_turbulence_closures.jl:300-320; its golden LES files are remote DataDeps)
# ---------------------------------------------------------------------------------------------

@pytest.mark.parametrize("FT", [np.float64, np.float32])
def test_smagorinsky_viscosity_of_a_linear_shear(FT):
    """u = S·z: Σ₁₃ = S/2, 2ΣᵢⱼΣᵢⱼ = S², so νₑ = (C·Δᶠ)²·|S| (smagorinsky.jl:46-50); with b = N²·z the Lilly factor is
    √(1 - Cb·N²/Σ²) = √(1 - 2·Cb·N²/S²) (lilly_coefficient.jl:13-16,125-130), clipped at 0 and unity for N² ≤ 0."""
    from oracle.model import update_state
    S, C = 3.0, 0.16
    N = (8, 8, 16)
    g = o.RectilinearGrid(FT, size=N, extent=(1, 1, 2), topology=(P, P, B))
    zc = -2 + (np.arange(16) + 0.5) * 2 / 16
    u = np.broadcast_to(S * zc, N)
    delta_f = np.cbrt((1 / 8) * (1 / 8) * (2 / 16))
    rtol = 1e-12 if FT == np.float64 else 2e-6
    for Cb, N2, expect in ((1.0, 0.0, 1.0), (1.0, 2.0, np.sqrt(1 - 2 * 2.0 / S ** 2)), (1.0, 100.0, 0.0), (1.0, -5.0, 1.0),
                           (0.0, 2.0, 1.0)):
        m = o.NonhydrostaticModel(g, tracers=("b",), buoyancy=o.BuoyancyTracer(), closure=o.SmagorinskyLilly(C=C, Cb=Cb))
        m.velocities[0].set(u)
        m.tracers["b"].set(np.broadcast_to(N2 * zc, N))
        update_state(m, compute_tendencies=False)
        nu = m.nu_e.interior[:, :, 2:-2]                  # away from the walls (mirror halos flatten the profiles there)
        np.testing.assert_allclose(nu, (C * delta_f) ** 2 * S * expect, rtol=rtol, atol=1e-30)
    m = o.NonhydrostaticModel(g, closure=o.Smagorinsky(coefficient=C))
    m.velocities[0].set(u)
    update_state(m, compute_tendencies=False)
    np.testing.assert_allclose(m.nu_e.interior[:, :, 2:-2], (C * delta_f) ** 2 * S, rtol=rtol)


def test_smagorinsky_time_stepping_stays_incompressible_and_dissipates():
    """time_stepping_works_with_closure (test_time_stepping.jl:36-49) for ConstantSmagorinsky and SmagorinskyLilly, plus
    the physics the closure is there for: kinetic energy of a random field decays faster than without it."""
    rng = np.random.default_rng(11)
    N = (8, 8, 8)
    ics = {n: rng.uniform(-1, 1, N) for n in "uvw"}
    ke = {}
    for name, cl in (("none", None), ("smag", o.Smagorinsky()), ("lilly", o.SmagorinskyLilly())):
        g = o.RectilinearGrid(np.float64, size=N, extent=(1, 1, 1), topology=(P, P, P))
        m = o.NonhydrostaticModel(g, advection=o.Centered(np.float64, 2), closure=cl, tracers=("b",) if cl else (),
                                  buoyancy=o.BuoyancyTracer() if cl else None)
        o.set_model(m, **ics)
        for _ in range(5):
            o.time_step(m, 1e-3)
        u, v, w = m.velocities
        R = interior_region(g)
        assert np.abs(divergence_ccc(g, u, v, w, R)).max() < 1e-10
        ke[name] = sum((f.interior ** 2).sum() for f in m.velocities)
    assert ke["smag"] < ke["none"] and ke["lilly"] < ke["none"]


def test_amd_analytic_eddy_viscosity_and_diffusivity():
    """AnisotropicMinimumDissipation (anisotropic_minimum_dissipation.jl:143-189) on linear flows, where every discrete
    gradient is exact: a pure shear u = S·z has uᵢₐuⱼₐΣᵢⱼ = 0, so νₑ = 0 (the closure switches off in laminar shear);
    the strain u = a·x, v = b·y, w = -(a+b)·z has r = a³ + b³ - (a+b)³ and q = a² + b² + (a+b)², so
    νₑ = max(0, -Cν·δ²·r/q) with δ² = 3/(1/Δᶠx² + 1/Δᶠy² + 1/Δᶠz²), Δᶠ = 2Δ; a tracer c = γ·z in it has
    κₑ = max(0, Cκ·δ²·(a+b))."""
    from oracle.model import update_state
    FT = np.float64
    N = (8, 8, 8)
    g = o.RectilinearGrid(FT, size=N, x=(0, 1), y=(0, 2), z=(0, 0.5), topology=(B, B, B))
    d = [1 / 8, 2 / 8, 0.5 / 8]
    d2 = 3 / sum(1 / (2 * x) ** 2 for x in d)
    C = 1 / 12
    xf, yf, zf = (np.arange(9) * d[0]), (np.arange(9) * d[1]), (np.arange(9) * d[2])
    zc = (np.arange(8) + 0.5) * d[2]
    inner = (slice(2, -2),) * 3
    # pure shear
    m = o.NonhydrostaticModel(g, closure=o.AnisotropicMinimumDissipation())
    m.velocities[0].set(np.broadcast_to(3.0 * zc, (9, 8, 8)))
    update_state(m, compute_tendencies=False)
    assert np.abs(m.nu_e.interior[inner]).max() == 0
    # three-dimensional strain, both signs of a·b·(a+b)
    for a, b in ((1.0, 2.0), (-1.0, -0.5)):
        m = o.NonhydrostaticModel(g, tracers=("c",), closure=o.AnisotropicMinimumDissipation())
        m.velocities[0].set(np.broadcast_to((a * xf)[:, None, None], (9, 8, 8)))
        m.velocities[1].set(np.broadcast_to((b * yf)[None, :, None], (8, 9, 8)))
        m.velocities[2].set(np.broadcast_to((-(a + b) * zf)[None, None, :], (8, 8, 9)))
        m.tracers["c"].set(np.broadcast_to(0.7 * zc, N))
        update_state(m, compute_tendencies=False)
        r, q = a ** 3 + b ** 3 - (a + b) ** 3, a ** 2 + b ** 2 + (a + b) ** 2
        np.testing.assert_allclose(m.nu_e.interior[inner], max(0.0, -C * d2 * r / q), rtol=1e-12, atol=1e-18)
        np.testing.assert_allclose(m.kappa_e["c"].interior[inner], max(0.0, C * d2 * (a + b)), rtol=1e-12, atol=1e-18)


# ---------------------------------------------------------------------------------------------
# Dynamics — test/test_dynamics.jl:17-86,176-207,404-460,495-600,620-624 (explicit ScalarDiffusivity cases)
# ---------------------------------------------------------------------------------------------

def _field(m, name):
    return dict(zip(m.prognostic_names(), m.prognostic_fields()))[name]


@pytest.mark.parametrize("vitd", [False, True], ids=["explicit", "vertically-implicit"])
@pytest.mark.parametrize("timestepper", ["QuasiAdamsBashforth2", "RungeKutta3"])
@pytest.mark.parametrize("name", ["u", "v", "c"])
def test_diffusion_simple(name, timestepper, vitd):
    """test_diffusion_simple: a constant field stays constant under ν = κ = 1 diffusion, 10 steps of Δt = 1 on (1, 1, 16),
    for ExplicitTimeDiscretization and VerticallyImplicitTimeDiscretization (test_dynamics.jl:404-414)."""
    from oracle.model import update_state
    g = o.RectilinearGrid(np.float64, size=(1, 1, 16), extent=(1, 1, 1), topology=(P, P, B))
    m = o.NonhydrostaticModel(g, timestepper=timestepper, closure=o.ScalarDiffusivity(1, 1, vertically_implicit=vitd),
                              tracers=("c",))
    f = _field(m, name)
    f.interior[...] = np.pi
    update_state(m)
    for _ in range(10):
        o.time_step(m, 1.0)
    assert np.allclose(f.interior, np.pi, rtol=np.sqrt(np.finfo(np.float64).eps), atol=0)


_BUDGET_TOPOS = [(P, P, P), (P, P, B), (P, B, B), (B, B, B)]


@pytest.mark.parametrize("vitd", [False, True], ids=["explicit", "vertically-implicit"])
@pytest.mark.parametrize("timestepper", ["QuasiAdamsBashforth2", "RungeKutta3"])
@pytest.mark.parametrize("topo", _BUDGET_TOPOS, ids=["".join(t[0] for t in tp) for tp in _BUDGET_TOPOS])
def test_scalar_diffusivity_budget(topo, timestepper, vitd):
    """test_ScalarDiffusivity_budget: the mean of a random field is conserved by isotropic diffusion (c everywhere, a
    velocity component along each Periodic direction), 10 steps of Δt = 1e-4·Δ²/κ on 4³; the vertically implicit
    discretization on the vertically Bounded topologies."""
    from oracle.model import update_state
    if vitd and topo[2] != B:
        with pytest.raises(ValueError):
            o.NonhydrostaticModel(o.RectilinearGrid(np.float64, size=(4, 4, 4), extent=(1, 1, 1), topology=topo),
                                  closure=o.ScalarDiffusivity(1, 1, vertically_implicit=True))
        return
    rng = np.random.default_rng(0)
    names = ["c"] + [n for n, t in zip("uvw", topo) if t == P]
    for name in names:
        g = o.RectilinearGrid(np.float64, size=(4, 4, 4), extent=(1, 1, 1), topology=topo)
        m = o.NonhydrostaticModel(g, timestepper=timestepper, closure=o.ScalarDiffusivity(1, 1, vertically_implicit=vitd),
                                  tracers=("c",))
        f = _field(m, name)
        o.set_model(m, **{name: rng.random(f.size)})
        init = f.interior.mean()
        update_state(m)
        dt = 1e-4 * 0.25 ** 2 / 1
        for _ in range(10):
            o.time_step(m, dt)
        assert np.isclose(init, f.interior.mean(), rtol=np.sqrt(np.finfo(np.float64).eps), atol=0), (name, topo)


_COSINE = [((128, 1, 1), (B, P, P), 0, ("v", "w", "c")), ((1, 128, 1), (P, B, P), 1, ("u", "w", "c")),
           ((1, 1, 128), (P, P, B), 2, ("u", "v", "c")), ((128,), (B, F, F), 0, ("v", "w", "c")),
           ((128,), (F, F, B), 2, ("u", "v", "c")),
           ((1, 1, 128), (P, P, B), 2, ("u", "v", "c"), True), ((128,), (F, F, B), 2, ("u", "v", "c"), True)]


@pytest.mark.parametrize("case", _COSINE, ids=["x", "y", "z", "x-flat", "z-flat", "z-implicit", "z-flat-implicit"])
def test_diffusion_cosine(case):
    """test_diffusion_cosine: cos(2ξ) on ξ ∈ (0, π/2) between no-flux walls decays like exp(-κ m² t), 5 RK3 steps of
    Δt = 1e-6·Lz²/κ, to atol = rtol = 1e-6 (ScalarDiffusivity(ν = 1, κ = 1), explicit or vertically implicit;
    test_dynamics.jl:495-520)."""
    size, topo, dim, names = case[:4]
    vitd = len(case) > 4
    from oracle.model import update_state
    L = np.pi / 2
    kw = {}
    full = []
    for d, (axis, t) in enumerate(zip("xyz", topo)):
        if t != F:
            kw[axis] = (0, L) if d == dim else (0, 1)
            full.append(d)
    for name in names:
        g = o.RectilinearGrid(np.float64, size=size, topology=topo, **kw)
        m = o.NonhydrostaticModel(g, closure=o.ScalarDiffusivity(1, 1, vertically_implicit=vitd), tracers=("c",),
                                  buoyancy=None)
        f = _field(m, name)
        xi = (np.arange(128) + 0.5) * L / 128
        shape = [1, 1, 1]
        shape[dim] = 128
        f.interior[...] = np.cos(2 * xi).reshape(shape)
        update_state(m)
        dt = 1e-6 * float(g.Lz) ** 2 / 1
        for _ in range(5):
            o.time_step(m, dt)
        expect = (np.exp(-4 * m.clock.time) * np.cos(2 * xi)).reshape(shape)
        assert np.allclose(f.interior, np.broadcast_to(expect, f.interior.shape), rtol=1e-6, atol=1e-6), name


def test_passive_tracer_advection():
    """passive_tracer_advection_test(:QuasiAdamsBashforth2): a Gaussian carried by (U, V) = (0.5, 0.8) on 128×128×2 for
    100 steps of Δt = 0.05·Δx/|U| stays within 1e-4 (mean-square relative error) of the translated Gaussian."""
    N, L, U, V = 128, 1.0, 0.5, 0.8
    delta, x0, y0 = L / 15, L / 2, L / 2
    dt = 0.05 * L / N / np.sqrt(U ** 2 + V ** 2)
    g = o.RectilinearGrid(np.float64, size=(N, N, 2), extent=(L, L, L), topology=(P, P, B))
    m = o.NonhydrostaticModel(g, closure=o.ScalarDiffusivity(1e-12, 1e-12), timestepper="QuasiAdamsBashforth2",
                              buoyancy=o.SeawaterBuoyancy(), tracers=("T", "S"))
    xc = (np.arange(N) + 0.5) * L / N

    def T(t):
        return np.exp(-(((xc - U * t - x0) ** 2)[:, None] + ((xc - V * t - y0) ** 2)[None, :]) / (2 * delta ** 2))
    o.set_model(m, u=np.full((N, N, 2), U), v=np.full((N, N, 2), V), T=np.broadcast_to(T(0)[:, :, None], (N, N, 2)))
    for _ in range(100):
        o.time_step(m, dt)
    ans = np.broadcast_to(T(m.clock.time)[:, :, None], (N, N, 2))
    num = m.tracers["T"].interior
    assert ((num - ans) ** 2).mean() / (ans ** 2).mean() < 1e-4
