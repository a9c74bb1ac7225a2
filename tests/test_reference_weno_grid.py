"""WENO(grid) — stretched reconstruction coefficients (src/Advection/reconstruction_coefficients.jl:74-115,
237-324, weno_reconstruction.jl:98-124, weno_interpolants.jl:124-127).  The reference has no test of its own for these tables; they
are pinned here by what they must satisfy — exact reconstruction of polynomials from cell averages on the stretched grid, the uniform
coefficients on an evenly spaced "stretched" grid, last = 1 − Σ — and the two independent generators (oracle / host mirror) and the
kernels (GPU) by each other, bit for bit."""
import numpy as np
import pytest

import oracle as o
from oracle.advection import stencil_coefficients, stretched_weno_tables

P, B_ = "Periodic", "Bounded"
N = 12
TANH = -1 + np.tanh(np.linspace(0, 1.5, N + 1)) / np.tanh(1.5)


def _grid(mod, FT, z=TANH, size=(8, 8, N)):
    return mod.RectilinearGrid(FT, size=size, x=(0, 1), y=(0, 1), z=z, topology=(P, P, B_))


@pytest.mark.parametrize("B", [2, 3])
def test_tables_reconstruct_polynomials_exactly(B):
    """A polynomial of degree B − 1 is recovered at every face / centre from its B cell averages, for every stencil r = −1 … B−1."""
    g = _grid(o, np.float64)
    t = stretched_weno_tables(g, np.float64, B)
    gh = g.with_halo((4, 4, 4))
    F, C = gh.coords[2].F, gh.coords[2].C
    coef = [0.3, 1.7, -2.1][:B]
    poly = lambda z: sum(c * z ** n for n, c in enumerate(coef))
    prim = lambda z: sum(c * z ** (n + 1) / (n + 1) for n, c in enumerate(coef))
    for key, X, last in (("f", F, N + 1), ("c", C, N)):
        avg = lambda k: (prim(X[k + 1]) - prim(X[k])) / (X[k + 1] - X[k])
        for i in range(1, last + 1):
            for r in range(-1, B):
                value = sum(t[key][r][i][q] * avg(i - 1 - r + q) for q in range(B))
                assert abs(value - poly(X[i])) < 2e-14, (key, i, r)


@pytest.mark.parametrize("FT", [np.float32, np.float64])
@pytest.mark.parametrize("B", [2, 3])
def test_tables_reduce_to_the_uniform_coefficients_on_even_spacing(B, FT):
    g = _grid(o, FT, z=np.linspace(0, 2, N + 1))
    t = stretched_weno_tables(g, FT, B)
    eps = np.finfo(FT).eps
    for r in range(B):
        uniform = np.array(stencil_coefficients(FT, r, B), dtype=np.float64)
        for i in range(N + 2):
            row = np.array(t["f"][r][i], dtype=np.float64)
            assert np.abs(row - uniform).max() < 40 * eps
            assert abs(row.sum() - 1) < 4 * eps                              # last = 1 − Σ of the others


@pytest.mark.parametrize("FT", [np.float32, np.float64])
@pytest.mark.parametrize("B", [2, 3])
def test_host_mirror_generator_equals_the_oracle_bit_for_bit(B, FT):
    import oceananigans_jl_b200.stretched_weno as sw
    t = stretched_weno_tables(_grid(o, FT), FT, B)
    mine = sw.weno_z_tables(TANH, FT, B, 3)
    for location, key in ((0, "f"), (1, "c")):
        assert mine[location].shape == (N + 2, B + 1, B) and mine[location].dtype == FT
        for r in range(-1, B):
            for i in range(N + 2):
                assert tuple(mine[location][i, r + 1, :]) == tuple(t[key][r][i]), (key, r, i)


def _ics(FT):
    rng = np.random.default_rng(0)
    return dict(u=rng.standard_normal((8, 8, N)).astype(FT), v=rng.standard_normal((8, 8, N)).astype(FT),
                w=rng.standard_normal((8, 8, N + 1)).astype(FT), c=rng.standard_normal((8, 8, N)).astype(FT))


def test_oracle_weno_with_grid_changes_the_z_reconstruction_on_a_stretched_grid_only():
    FT = np.float64
    out = {}
    for name, z in (("stretched", TANH), ("regular", (-1, 0))):
        g = _grid(o, FT, z=z)
        for adv in ("uniform", "grid"):
            m = o.NonhydrostaticModel(g, advection=o.WENO(FT, 5, grid=g if adv == "grid" else None), tracers=("c",))
            o.set_model(m, **_ics(FT))
            for _ in range(3):
                o.time_step(m, 1e-3)
            out[name, adv] = m.tracers["c"].interior.copy()
    assert np.array_equal(out["regular", "uniform"], out["regular", "grid"])          # no stretched direction: nothing changes
    d = np.abs(out["stretched", "uniform"] - out["stretched", "grid"]).max()
    assert 1e-6 < d < 0.1                                                               # the same scheme to truncation error


# ---- the library -----------------------------------------------------------------------------------------------------------------

@pytest.mark.gpu
@pytest.mark.parametrize("physics", ["plain", "full"])
@pytest.mark.parametrize("FT", [np.float32, np.float64])
def test_library_weno_grid_tendencies_bit_exact(FT, physics):
    import oceananigans_jl_b200 as oc
    from oracle.model import update_state
    g1, g2 = _grid(oc, FT), _grid(o, FT)
    kw1, kw2 = dict(tracers=("c",)), dict(tracers=("c",))
    if physics == "full":
        kw1 = dict(tracers=("b", "c"), buoyancy=oc.BuoyancyTracer(), coriolis=oc.FPlane(f=0.3), closure=oc.SmagorinskyLilly())
        kw2 = dict(tracers=("b", "c"), buoyancy=o.BuoyancyTracer(), coriolis=o.FPlane(f=0.3), closure=o.SmagorinskyLilly())
    m1 = oc.NonhydrostaticModel(g1, advection=oc.WENO(g1), **kw1)
    m2 = o.NonhydrostaticModel(g2, advection=o.WENO(FT, 5, grid=g2), **kw2)
    ics = _ics(FT)
    if physics == "full":
        ics["b"] = ics["c"][::-1].copy()
    oc.set_model(m1, enforce_incompressibility=False, **ics)
    o.set_model(m2, enforce_incompressibility=False, **ics)
    oc.update_state(m1, True)
    update_state(m2)
    for n in ["u", "v", "w"] + list(kw1["tracers"]):
        np.testing.assert_array_equal(m1.timestepper.Gn[n].parent(), m2.Gn[n].data, err_msg=n)
    # and the uniform-coefficient scheme on the same grid is a different discretisation
    m3 = oc.NonhydrostaticModel(g1, advection=oc.WENO(FT, 5), **kw1)
    oc.set_model(m3, enforce_incompressibility=False, **ics)
    oc.update_state(m3, True)
    assert not np.array_equal(m3.timestepper.Gn["c"].parent(), m1.timestepper.Gn["c"].parent())


@pytest.mark.gpu
@pytest.mark.parametrize("timestepper", ["RungeKutta3", "QuasiAdamsBashforth2"])
def test_library_weno_grid_100_steps_vs_oracle(timestepper):
    import oceananigans_jl_b200 as oc
    FT = np.float64
    g1, g2 = _grid(oc, FT), _grid(o, FT)
    m1 = oc.NonhydrostaticModel(g1, advection=oc.WENO(g1), tracers=("c",), timestepper=timestepper,
                                closure=oc.ScalarDiffusivity(nu=1e-3, kappa=1e-3))
    m2 = o.NonhydrostaticModel(g2, advection=o.WENO(FT, 5, grid=g2), tracers=("c",), timestepper=timestepper,
                               closure=o.ScalarDiffusivity(nu=1e-3, kappa=1e-3))
    ics = {k: 0.3 * v for k, v in _ics(FT).items()}
    oc.set_model(m1, **ics)
    o.set_model(m2, **ics)
    oc.time_step(m1, 2e-3, n_steps=100)
    for _ in range(100):
        o.time_step(m2, 2e-3)
    for f1, f2 in zip(m1.prognostic_fields(), m2.prognostic_fields()):
        a, b = f1.interior(), f2.interior
        assert np.linalg.norm((a - b).ravel()) / np.linalg.norm(b.ravel()) <= 1e-12, f1.name
    assert m1.max_abs_divergence() < 1e-11


@pytest.mark.gpu
def test_library_weno_grid_on_a_regular_grid_and_errors():
    import oceananigans_jl_b200 as oc
    import ctypes as C
    from oceananigans_jl_b200._lib import lib, NhError, check
    FT = np.float64
    g = _grid(oc, FT, z=(-1, 0))
    ics = _ics(FT)
    res = []
    for adv in (oc.WENO(g), oc.WENO(FT, 5)):
        m = oc.NonhydrostaticModel(g, advection=adv, tracers=("c",))
        oc.set_model(m, **ics)
        oc.time_step(m, 1e-3, n_steps=2)
        res.append(m.tracers["c"].parent())
    np.testing.assert_array_equal(res[0], res[1])                 # no stretched direction: the uniform coefficients
    gs = _grid(oc, FT)
    m = oc.NonhydrostaticModel(gs, advection=oc.WENO(gs), tracers=("c",))
    bad = np.zeros(5)
    with pytest.raises(NhError, match="count must be"):
        check(lib().nh_model_set_weno_z_coefficients(m._h, 3, 0, bad.ctypes.data_as(C.c_void_p), bad.size, 0))
    with pytest.raises(NhError, match="buffer must be"):
        check(lib().nh_model_set_weno_z_coefficients(m._h, 4, 0, bad.ctypes.data_as(C.c_void_p), bad.size, 0))
    mc = oc.NonhydrostaticModel(gs, advection=oc.Centered(FT, 2))
    with pytest.raises(NhError, match="not a WENO scheme"):
        check(lib().nh_model_set_weno_z_coefficients(mc._h, 2, 0, bad.ctypes.data_as(C.c_void_p), bad.size, 0))
    other = _grid(oc, FT, z=np.linspace(-1, 0, N + 1) ** 3)
    with pytest.raises(ValueError, match="another vertical coordinate"):
        oc.NonhydrostaticModel(gs, advection=oc.WENO(other))
