"""The reference's boundary-condition integration tests (test/test_boundary_conditions_integration.jl:28-54,384-433: the flux budget
of a field at rest; :10-28,355-358: array conditions on a one-cell box) against the library through the C ABI.  The oracle versions
are tests/test_oracle_pins.py::test_nonhydrostatic_flux_budget and ::test_array_boundary_condition_time_steps."""
import numpy as np
import pytest

P, B = "Periodic", "Bounded"
pytestmark = pytest.mark.gpu

_BUDGET = ([((P, B, B), n, side) for n in ("u", "c") for side in ("north", "south", "top", "bottom")]
           + [((B, P, B), n, side) for n in ("v", "c") for side in ("east", "west", "top", "bottom")]
           + [((B, B, P), n, side) for n in ("w", "c") for side in ("east", "west", "north", "south")])


@pytest.mark.parametrize("timestepper", ["RungeKutta3", "QuasiAdamsBashforth2"])
@pytest.mark.parametrize("topo,name,side", _BUDGET, ids=[f"{''.join(x[0] for x in t)}-{n}-{s}" for t, n, s in _BUDGET])
def test_library_nonhydrostatic_flux_budget(topo, name, side, timestepper):
    """A constant boundary flux π into a field at rest raises its mean by flux · t / L."""
    import oceananigans_jl_b200 as oc
    FT = np.float64
    Lx, Ly, Lz = 0.3, 0.4, 0.5
    L = {"west": Lx, "east": Lx, "south": Ly, "north": Ly, "bottom": Lz, "top": Lz}[side]
    flux = np.pi
    direction = 1 if side in ("west", "south", "bottom") else -1
    g = oc.RectilinearGrid(FT, size=(2, 2, 2), x=(0, Lx), y=(0, Ly), z=(0, Lz), topology=topo)
    bcs = {name: oc.FieldBoundaryConditions(**{side: oc.FluxBoundaryCondition(flux * direction)})}
    m = oc.NonhydrostaticModel(g, tracers=("c",), boundary_conditions=bcs, timestepper=timestepper)
    oc.time_step(m, 1.0)
    f = dict(zip(["u", "v", "w", "c"], m.prognostic_fields()))[name]
    assert np.isclose(f.interior().mean(), flux * m.clock.time / L, rtol=np.sqrt(np.finfo(FT).eps), atol=0)


@pytest.mark.parametrize("kind", ["Value", "Gradient", "Flux"])
@pytest.mark.parametrize("side", ["east", "south", "top"])
def test_library_array_boundary_condition_time_steps(kind, side):
    """A 1 × 1 array condition on T of a (1, 1, 1) Bounded box steps (Δt = 1e-16) without error."""
    import oceananigans_jl_b200 as oc
    g = oc.RectilinearGrid(np.float64, size=(1, 1, 1), extent=(1, np.pi, 42), topology=(B, B, B))
    arr = np.random.default_rng(0).random((1, 1))
    bc = getattr(oc, kind + "BoundaryCondition")(arr)
    m = oc.NonhydrostaticModel(g, tracers=("T", "S"), buoyancy=oc.SeawaterBuoyancy(),
                               boundary_conditions={"T": oc.FieldBoundaryConditions(**{side: bc})})
    oc.time_step(m, 1e-16)
    assert np.all(np.isfinite(m.tracers["T"].interior()))
