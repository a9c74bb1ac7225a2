"""More of the reference's own tests for this path restated over the oracle (CPU): the internal-wave test of test_dynamics.jl,
and the time-stepping smoke tests of test_time_stepping.jl that tests/test_oracle_pins.py does not already hold.  They pin the
restatement on the reference's analytic answers where the reference has them; the CUDA library is pinned on the oracle
(tests/test_gpu_parity.py runs these physics combinations bit for bit)."""
import numpy as np
import pytest

import oracle as o

P, B, F = "Periodic", "Bounded", "Flat"


# ---------------------------------------------------------------------------------------------
# Internal wave — test_dynamics.jl:627-685, test_internal_wave_dynamics.jl:4-87
# ---------------------------------------------------------------------------------------------

def _internal_wave_solution(L):
    nu = kappa = 1e-9
    z0, delta, a0, m, k, f, N = -L / 3, L / 20, 1e-3, 16, 1, 0.2, 1.0
    sigma = np.sqrt((N ** 2 * k ** 2 + f ** 2 * m ** 2) / (k ** 2 + m ** 2))
    dt = 0.01 * 1 / sigma
    cg = m * sigma / (k ** 2 + m ** 2) * (f ** 2 / sigma ** 2 - 1)
    U = a0 * k * sigma / (sigma ** 2 - f ** 2)
    V = a0 * k * f / (sigma ** 2 - f ** 2)
    W = a0 * m * sigma / (sigma ** 2 - N ** 2)
    Bb = a0 * m * N ** 2 / (sigma ** 2 - N ** 2)
    a = lambda x, z, t: np.exp(-(z - cg * t - z0) ** 2 / (2 * delta) ** 2)
    sol = dict(u=lambda x, z, t: a(x, z, t) * U * np.cos(k * x + m * z - sigma * t),
               v=lambda x, z, t: a(x, z, t) * V * np.sin(k * x + m * z - sigma * t),
               w=lambda x, z, t: a(x, z, t) * W * np.cos(k * x + m * z - sigma * t),
               b=lambda x, z, t: a(x, z, t) * Bb * np.sin(k * x + m * z - sigma * t) + N ** 2 * z)
    kwargs = dict(closure=o.ScalarDiffusivity(nu=nu, kappa=kappa), buoyancy=o.BuoyancyTracer(), tracers=("b",),
                  coriolis=o.FPlane(f=f))
    return sol, kwargs, dt


def _internal_wave_grids():
    Nx = Nz = 128
    Lx = Lz = 2 * np.pi
    regular = o.RectilinearGrid(np.float64, topology=(P, P, B), size=(Nx, 1, Nz), x=(0, Lx), y=(0, Lx), z=(-Lz, 0))
    flat = o.RectilinearGrid(np.float64, topology=(P, F, B), size=(Nx, Nz), x=(0, Lx), z=(-Lz, 0))
    z_faces = regular.coords[2].F.r(1, Nz + 1).astype(np.float64)
    stretched = o.RectilinearGrid(np.float64, topology=(P, P, B), size=(Nx, 1, Nz), x=(0, Lx), y=(0, Lx), z=z_faces)
    flat_stretched = o.RectilinearGrid(np.float64, topology=(P, F, B), size=(Nx, Nz), x=(0, Lx), z=z_faces)
    return {"y_periodic_regular": regular, "y_flat_regular": flat, "y_periodic_stretched": stretched,
            "y_flat_stretched": flat_stretched}


@pytest.mark.parametrize("which", ["y_periodic_regular", "y_flat_regular", "y_periodic_stretched", "y_flat_stretched"])
def test_internal_wave_nonhydrostatic(which):
    grid = _internal_wave_grids()[which]
    sol, kwargs, dt = _internal_wave_solution(2 * np.pi)
    m = o.NonhydrostaticModel(grid, **kwargs)
    Nx, Nz = grid.Nx, grid.Nz
    cx, cz = grid.coords[0], grid.coords[2]
    node = {"c": lambda c, n: c.C.r(1, n), "f": lambda c, n: c.F.r(1, n)}

    def on(name, t):
        f = dict(zip("uvwb", m.prognostic_fields()))[name]
        x = node[f.loc[0]](cx, f.size[0])[:, None, None]
        z = node[f.loc[2]](cz, f.size[2])[None, None, :]
        return np.broadcast_to(sol[name](x, z, t), f.size)

    o.set_model(m, **{n: on(n, 0.0) for n in "uvwb"})
    for _ in range(10):
        o.time_step(m, dt)
    u, ua = m.velocities[0].interior, on("u", m.clock.time)
    assert m.clock.iteration == 10
    assert np.mean((u - ua) ** 2) / np.mean(ua ** 2) < 1e-4               # test_internal_wave_dynamics.jl:84, relative_error


# ---------------------------------------------------------------------------------------------
# test_time_stepping.jl: smoke tests (the model must construct, step once and stay finite)
# ---------------------------------------------------------------------------------------------

def _finite(m):
    return all(np.all(np.isfinite(f.data)) for f in m.prognostic_fields())


@pytest.mark.parametrize("n", [1, 2])
@pytest.mark.parametrize("topology", [(F, P, P), (P, F, P), (P, P, F), (F, F, B),                 # the reference's four
                                      (F, P, B), (P, F, B), (P, B, F), (F, B, B), (B, F, B), (B, B, F)])
def test_time_stepping_works_with_flat_dimensions(topology, n):
    """test_time_stepping.jl:10-17,332-344 (one cell per non-Flat direction there; two as well here)."""
    size = tuple(n for t in topology if t != F)
    extent = tuple(1 for t in topology if t != F)
    grid = o.RectilinearGrid(np.float64, size=size, extent=extent, topology=topology)
    m = o.NonhydrostaticModel(grid)
    o.time_step(m, 1.0)
    assert m.clock.iteration == 1 and _finite(m)


def test_euler_time_stepping_doesnt_propagate_nans():
    """test_time_stepping.jl:19-29 (there on the HydrostaticFreeSurfaceModel; the same QuasiAdamsBashforth2 step here): G⁻ full of
    NaN must not reach u in an Euler step (χ = −1/2 zeroes its coefficient)."""
    grid = o.RectilinearGrid(np.float64, size=(1, 1, 1), extent=(1, 2, 3))
    m = o.NonhydrostaticModel(grid, timestepper="QuasiAdamsBashforth2", buoyancy=o.BuoyancyTracer(), tracers=("b",))
    m.Gm["u"].data[...] = np.nan
    o.time_step(m, 1.0, euler=True)
    assert not np.isnan(m.velocities[0].interior).any()


@pytest.mark.parametrize("FT", [np.float32, np.float64])
def test_time_stepping_works_with_coriolis(FT):
    """test_time_stepping.jl:31-37,346-351 for the Coriolis type on this path (FPlane)."""
    grid = o.RectilinearGrid(FT, size=(1, 1, 1), extent=(1, 2, 3))
    m = o.NonhydrostaticModel(grid, coriolis=o.FPlane(latitude=45))
    assert abs(float(m.coriolis.f) - 2 * 7.292115e-5 * np.sqrt(0.5)) < 1e-10        # f = 2Ω sind(φ), f_plane.jl:33-43
    o.time_step(m, 1.0)
    assert m.clock.iteration == 1 and _finite(m)


_ADVECTION = [("none", lambda FT: None), ("Centered2", lambda FT: o.Centered(FT, 2)), ("Centered4", lambda FT: o.Centered(FT, 4)),
              ("UpwindBiased1", lambda FT: o.UpwindBiased(FT, 1)), ("UpwindBiased3", lambda FT: o.UpwindBiased(FT, 3)),
              ("UpwindBiased5", lambda FT: o.UpwindBiased(FT, 5)), ("WENO3", lambda FT: o.WENO(FT, 3)), ("WENO5", lambda FT: o.WENO(FT, 5))]


@pytest.mark.parametrize("name,scheme", _ADVECTION, ids=[a[0] for a in _ADVECTION])
def test_time_stepping_works_with_advection_scheme(name, scheme):
    """test_time_stepping.jl:53-59,353-358 for the schemes on this path; the 3³ grid makes every scheme drop to the order the
    size allows (adapt_advection_order)."""
    grid = o.RectilinearGrid(np.float64, size=(3, 3, 3), extent=(1, 2, 3))
    m = o.NonhydrostaticModel(grid, advection=scheme(np.float64))
    rng = np.random.default_rng(0)
    o.set_model(m, u=rng.standard_normal((3, 3, 3)), v=rng.standard_normal((3, 3, 3)))
    o.time_step(m, 1e-3)
    assert m.clock.iteration == 1 and _finite(m)


_CLOSURES = [("nothing", lambda: None), ("ScalarDiffusivity", lambda: o.ScalarDiffusivity(nu=1e-2, kappa=1e-2)),
             ("ScalarDiffusivity-VITD", lambda: o.ScalarDiffusivity(nu=1e-2, kappa=1e-2, vertically_implicit=True)),
             ("SmagorinskyLilly", lambda: o.SmagorinskyLilly()), ("AnisotropicMinimumDissipation", lambda: o.AnisotropicMinimumDissipation())]


@pytest.mark.parametrize("name,closure", _CLOSURES, ids=[c[0] for c in _CLOSURES])
@pytest.mark.parametrize("FT", [np.float32, np.float64])
def test_time_stepping_works_with_closure(FT, name, closure):
    """test_time_stepping.jl:39-51,69-74,382-403 for the closures on this path: SeawaterBuoyancy, tracers T, S, one step."""
    grid = o.RectilinearGrid(FT, size=(3, 3, 3), extent=(1, 2, 3))
    m = o.NonhydrostaticModel(grid, closure=closure(), tracers=("T", "S"), buoyancy=o.SeawaterBuoyancy())
    o.time_step(m, 1.0)
    assert m.clock.iteration == 1 and _finite(m)


def test_time_stepping_works_with_amd_without_buoyancy():
    """test_time_stepping.jl:401-402: AnisotropicMinimumDissipation can depend on buoyancy — also without one."""
    grid = o.RectilinearGrid(np.float64, size=(3, 3, 3), extent=(1, 2, 3))
    m = o.NonhydrostaticModel(grid, closure=o.AnisotropicMinimumDissipation(), tracers=("T", "S"), buoyancy=None)
    o.time_step(m, 1.0)
    assert m.clock.iteration == 1 and _finite(m)


# ---- the same tests against the library through the C ABI ---------------------------------------------------------------------

def _lib_finite(m):
    return all(np.all(np.isfinite(f.parent())) for f in m.prognostic_fields())


@pytest.mark.gpu
@pytest.mark.parametrize("which", ["y_periodic_regular", "y_flat_regular", "y_periodic_stretched", "y_flat_stretched"])
def test_library_internal_wave_nonhydrostatic(which):
    """test_dynamics.jl:673-684 on the CUDA library: the analytic inertia-gravity wave packet after 10 RK3 steps."""
    import oceananigans_jl_b200 as oc
    og = _internal_wave_grids()[which]
    Nx = Nz = 128
    Lx = Lz = 2 * np.pi
    flat = og.topology[1] == F
    z = (-Lz, 0) if og.regular_dim(2) else og.coords[2].F.r(1, Nz + 1).astype(np.float64)
    kw = dict(topology=og.topology, size=(Nx, Nz) if flat else (Nx, 1, Nz), x=(0, Lx), z=z)
    if not flat:
        kw["y"] = (0, Lx)
    grid = oc.RectilinearGrid(np.float64, **kw)
    sol, _, dt = _internal_wave_solution(2 * np.pi)
    m = oc.NonhydrostaticModel(grid, closure=oc.ScalarDiffusivity(nu=1e-9, kappa=1e-9), buoyancy=oc.BuoyancyTracer(), tracers=("b",),
                               coriolis=oc.FPlane(f=0.2))
    oc.set_model(m, **{n: (lambda x, y, z, n=n: sol[n](x, z, 0.0)) for n in "uvwb"})
    oc.time_step(m, dt, n_steps=10)
    t = m.clock.time
    u = m.velocities.u
    xs, zs = grid.nodes(0, "f")[:, None, None], grid.nodes(2, "c")[None, None, :]
    ua = np.broadcast_to(sol["u"](xs, zs, t), u.size)
    un = u.interior()
    assert m.clock.iteration == 10
    assert np.mean((un - ua) ** 2) / np.mean(ua ** 2) < 1e-4


@pytest.mark.gpu
@pytest.mark.parametrize("n", [1, 2])
@pytest.mark.parametrize("topology", [(F, P, P), (P, F, P), (P, P, F), (F, F, B), (F, P, B), (P, F, B), (P, B, F), (F, B, B), (B, F, B),
                                      (B, B, F)])
def test_library_time_stepping_works_with_flat_dimensions(topology, n):
    import oceananigans_jl_b200 as oc
    size = tuple(n for t in topology if t != F)
    grid = oc.RectilinearGrid(np.float64, size=size, extent=tuple(1 for _ in size), topology=topology)
    m = oc.NonhydrostaticModel(grid)
    oc.time_step(m, 1.0)
    assert m.clock.iteration == 1 and _lib_finite(m)


@pytest.mark.gpu
def test_library_euler_time_stepping_doesnt_propagate_nans():
    import oceananigans_jl_b200 as oc
    grid = oc.RectilinearGrid(np.float64, size=(1, 1, 1), extent=(1, 2, 3))
    m = oc.NonhydrostaticModel(grid, timestepper="QuasiAdamsBashforth2", buoyancy=oc.BuoyancyTracer(), tracers=("b",))
    G = m.timestepper.Gm["u"]
    G.set_parent(np.full(G.parent_size, np.nan))
    oc.time_step(m, 1.0, euler=True)
    assert not np.isnan(m.velocities.u.interior()).any()


@pytest.mark.gpu
@pytest.mark.parametrize("name,scheme", _ADVECTION, ids=[a[0] for a in _ADVECTION])
def test_library_time_stepping_works_with_advection_scheme(name, scheme):
    import oceananigans_jl_b200 as oc
    lib_scheme = {"none": None, "Centered2": oc.Centered(np.float64, 2), "Centered4": oc.Centered(np.float64, 4),
                  "UpwindBiased1": oc.UpwindBiased(np.float64, 1), "UpwindBiased3": oc.UpwindBiased(np.float64, 3),
                  "UpwindBiased5": oc.UpwindBiased(np.float64, 5), "WENO3": oc.WENO(np.float64, 3), "WENO5": oc.WENO(np.float64, 5)}[name]
    grid = oc.RectilinearGrid(np.float64, size=(3, 3, 3), extent=(1, 2, 3))
    m = oc.NonhydrostaticModel(grid, advection=lib_scheme)
    rng = np.random.default_rng(0)
    oc.set_model(m, u=rng.standard_normal((3, 3, 3)), v=rng.standard_normal((3, 3, 3)))
    oc.time_step(m, 1e-3)
    assert m.clock.iteration == 1 and _lib_finite(m)


@pytest.mark.gpu
@pytest.mark.parametrize("name", [c[0] for c in _CLOSURES])
@pytest.mark.parametrize("FT", [np.float32, np.float64])
def test_library_time_stepping_works_with_closure(FT, name):
    import oceananigans_jl_b200 as oc
    closure = {"nothing": None, "ScalarDiffusivity": oc.ScalarDiffusivity(nu=1e-2, kappa=1e-2),
               "ScalarDiffusivity-VITD": oc.ScalarDiffusivity(nu=1e-2, kappa=1e-2, vertically_implicit=True),
               "SmagorinskyLilly": oc.SmagorinskyLilly(), "AnisotropicMinimumDissipation": oc.AnisotropicMinimumDissipation()}[name]
    grid = oc.RectilinearGrid(FT, size=(3, 3, 3), extent=(1, 2, 3))
    m = oc.NonhydrostaticModel(grid, closure=closure, tracers=("T", "S"), buoyancy=oc.SeawaterBuoyancy())
    oc.time_step(m, 1.0)
    assert m.clock.iteration == 1 and _lib_finite(m)
    m = oc.NonhydrostaticModel(oc.RectilinearGrid(FT, size=(1, 1, 1), extent=(1, 2, 3)), coriolis=oc.FPlane(latitude=45))
    oc.time_step(m, 1.0)
    assert m.clock.iteration == 1 and _lib_finite(m)
