"""GPU parity: the CUDA path (through the C ABI of include/nhstep.h) against the oracle on identical seeded inputs.

Tolerances: Float64 rel. L2 ≤ 1e-12, Float32 ≤ 1e-5 (BASELINE.json north_star); single kernels in the parity build
(-fmad=false, reference expression order) are held to bit-exactness where no FFT is involved.
"""
import itertools

import numpy as np
import pytest

import oracle as o
import parity_log
from oracle.model import (update_state as o_update_state, compute_tendencies_ as o_compute_tendencies,
                          calculate_pressure_correction as o_calc_pc, pressure_correct_velocities as o_pcv,
                          rk3_substep as o_rk3, ab2_step as o_ab2)
from oracle.fields import fill_halo_regions as o_fill

pytestmark = pytest.mark.gpu

P, B, F = "Periodic", "Bounded", "Flat"


@pytest.fixture(scope="module")
def oc():
    import oceananigans_jl_b200 as oc
    assert oc.device_count() >= 1, "no CUDA device: the product has no CPU fallback"
    return oc


def rel_l2(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    nb = np.linalg.norm(b.ravel())
    return np.linalg.norm((a - b).ravel()) / (nb if nb > 0 else 1.0)


def make_pair(oc, FT, size, topo, z=None, advection=("WENO", 5), timestepper="RungeKutta3", buoyancy=None, coriolis=None,
              closure=None, tracers=(), extent=None, x=None, y=None, exact_arithmetic=True, bcs=None):
    """The same model twice: (CUDA, oracle).  bcs = {field: {side: (kind, number)}} (Value / Gradient / Flux)."""
    kw = {}
    if extent is not None:
        kw["extent"] = extent
    else:
        if topo[0] != F:
            kw["x"] = x or (0, 1)
        if topo[1] != F:
            kw["y"] = y or (0, 1)
        if topo[2] != F:
            kw["z"] = z if z is not None else (0, 1)
    g1 = oc.RectilinearGrid(FT, size=size, topology=topo, **kw)
    g2 = o.RectilinearGrid(FT, size=size, topology=topo, **kw)

    def adv(mod, FTarg):
        if advection is None:
            return None
        kind, order = advection
        return getattr(mod, kind)(FTarg, order=order) if mod is o else getattr(mod, kind)(order=order)
    b1 = b2 = None
    if buoyancy == "seawater":
        b1, b2 = oc.SeawaterBuoyancy(), o.SeawaterBuoyancy()
    elif buoyancy == "tracer":
        b1, b2 = oc.BuoyancyTracer(), o.BuoyancyTracer()
    c1 = oc.FPlane(coriolis) if coriolis is not None else None
    c2 = o.FPlane(coriolis) if coriolis is not None else None
    if closure and closure[0] == "lilly":          # ("lilly", C, Cb, Pr)
        k1, k2 = oc.SmagorinskyLilly(*closure[1:]), o.SmagorinskyLilly(*closure[1:])
    elif closure and closure[0] == "amd":          # ("amd", C, Cnu, Ckappa)
        k1, k2 = oc.AnisotropicMinimumDissipation(*closure[1:]), o.AnisotropicMinimumDissipation(*closure[1:])
    elif closure and closure[0] == "smagorinsky":  # ("smagorinsky", coefficient, Pr)
        k1, k2 = oc.Smagorinsky(*closure[1:]), o.Smagorinsky(*closure[1:])
    else:
        k1 = oc.ScalarDiffusivity(*closure) if closure else None
        k2 = o.ScalarDiffusivity(*closure) if closure else None
    mk = {"value": oc.ValueBoundaryCondition, "gradient": oc.GradientBoundaryCondition, "flux": oc.FluxBoundaryCondition}
    bcs1 = {n: oc.FieldBoundaryConditions(**{side: mk[k](v) for side, (k, v) in sides.items()})
            for n, sides in (bcs or {}).items()}
    m1 = oc.NonhydrostaticModel(g1, advection=adv(oc, None), timestepper=timestepper, buoyancy=b1, coriolis=c1, closure=k1,
                                tracers=tracers, exact_arithmetic=exact_arithmetic, boundary_conditions=bcs1)
    m2 = o.NonhydrostaticModel(g2, advection=adv(o, FT), timestepper=timestepper, buoyancy=b2, coriolis=c2, closure=k2,
                               tracers=tracers, boundary_conditions=bcs)
    return m1, m2


def random_ics(m2, rng, amp=1.0):
    ics = {}
    for n, f in zip(m2.prognostic_names(), m2.prognostic_fields()):
        ics[n] = (amp * rng.uniform(-1, 1, f.size)).astype(m2.grid.FT)
    return ics


def names_fields(m1, m2):
    for n, f2 in zip(m2.prognostic_names(), m2.prognostic_fields()):
        f1 = m1.velocities[n] if n in "uvw" else m1.tracers[n]
        yield n, f1, f2


BCS_LES = {"u": {"top": ("flux", -1e-4), "bottom": ("value", 0.0)}, "v": {"bottom": ("value", 0.1)},
           "T": {"top": ("flux", 1e-3), "bottom": ("gradient", 0.01)}, "S": {"top": ("flux", -2e-4)}}
BCS_BOX = {"b": {"west": ("value", 1.0), "east": ("gradient", -0.5), "south": ("flux", 0.2), "north": ("value", 0.0),
                 "bottom": ("gradient", 0.3), "top": ("flux", -0.1)},
           "u": {"south": ("value", 0.5), "north": ("flux", 0.05), "top": ("gradient", 0.2)},
           "v": {"west": ("value", -0.25), "bottom": ("flux", 0.01)},
           "w": {"west": ("flux", 0.02), "east": ("value", 0.0), "north": ("gradient", 1.0)}}
BCS_STRETCHED = {"c": {"bottom": ("value", 2.0), "top": ("gradient", -1.0)}, "u": {"bottom": ("value", 0.0), "top": ("flux", 1e-3)},
                 "v": {"top": ("value", 0.2), "bottom": ("flux", -1e-3)}}

def _array_bcs(size, which):
    """Wall arrays of shape (N_e1, N_e2), seeded per wall."""
    rng = np.random.default_rng(2024)
    Nx, Ny, Nz = size

    def A(shape, lo=-0.5, hi=0.5):
        return rng.uniform(lo, hi, shape)
    if which == "z":
        return {"c": {"bottom": ("value", A((Nx, Ny))), "top": ("gradient", A((Nx, Ny)))},
                "u": {"bottom": ("value", A((Nx, Ny))), "top": ("flux", A((Nx, Ny), -1e-3, 1e-3))},
                "v": {"top": ("value", A((Nx, Ny))), "bottom": ("flux", A((Nx, Ny), -1e-3, 1e-3))}}
    if which == "les":
        return {"u": {"top": ("flux", A((Nx, Ny), -1e-4, 1e-4)), "bottom": ("value", 0.0)}, "v": {"bottom": ("value", A((Nx, Ny)))},
                "T": {"top": ("flux", A((Nx, Ny), 0, 1e-3)), "bottom": ("gradient", A((Nx, Ny), 0, 0.02))},
                "S": {"top": ("flux", -2e-4)}}
    return {"b": {"west": ("value", A((Ny, Nz))), "east": ("gradient", A((Ny, Nz))), "south": ("flux", A((Nx, Nz))),
                  "north": ("value", A((Nx, Nz))), "bottom": ("gradient", A((Nx, Ny))), "top": ("flux", A((Nx, Ny)))},
            "u": {"south": ("value", A((Nx, Nz))), "north": ("flux", A((Nx, Nz))), "top": ("gradient", A((Nx, Ny)))},
            "v": {"west": ("value", A((Ny, Nz))), "bottom": ("flux", A((Nx, Ny)))},
            "w": {"west": ("flux", A((Ny, Nz))), "east": ("value", A((Ny, Nz))), "north": ("gradient", A((Nx, Nz)))}}


BCS_STRETCHED_B = {("b" if k == "c" else k): v for k, v in BCS_STRETCHED.items()}

CONFIGS = [
    # size, topo, z, advection, kwargs
    ((12, 10, 9), (P, P, P), None, ("WENO", 5), {}),
    ((12, 10, 9), (P, P, B), None, ("WENO", 5), dict(buoyancy="seawater", tracers=("T", "S"), coriolis=1e-2, closure=(1e-2, 2e-2))),
    ((9, 8, 10), (B, B, B), None, ("WENO", 5), dict(buoyancy="tracer", tracers=("b",), closure=(1e-2, 1e-2))),
    ((9, 8, 10), (P, B, B), None, ("Centered", 4), dict(coriolis=0.5, tracers=("c",))),
    ((9, 8, 10), (B, P, B), None, ("UpwindBiased", 5), dict(tracers=("c",))),
    ((9, 8, 10), (P, P, B), None, ("WENO", 3), {}),
    ((9, 8, 10), (P, P, B), None, ("UpwindBiased", 3), {}),
    ((9, 8, 10), (P, P, B), None, ("Centered", 2), dict(closure=(1e-2, 1e-2), tracers=("c",))),
    ((16, 16), (P, P, F), None, ("WENO", 5), {}),
    ((16, 12), (P, F, B), None, ("WENO", 5), dict(buoyancy="tracer", tracers=("b",))),
    ((10, 9, 8), (P, P, B), "stretched", ("WENO", 5), dict(buoyancy="seawater", tracers=("T", "S"), closure=(1e-3, 1e-3))),
    ((9, 8, 2), (P, P, B), None, ("WENO", 5), {}),     # order adapted in z (N < buffer)
    # Value / Gradient / Flux boundary conditions with number conditions (the ocean-LES style set, then every wall pair)
    ((12, 10, 9), (P, P, B), None, ("WENO", 5), dict(buoyancy="seawater", tracers=("T", "S"), coriolis=1e-2, closure=(1e-2, 2e-2), bcs=BCS_LES)),
    ((9, 8, 10), (B, B, B), None, ("WENO", 5), dict(buoyancy="tracer", tracers=("b",), closure=(1e-2, 1e-2), bcs=BCS_BOX)),
    ((10, 9, 8), (P, P, B), "stretched", ("Centered", 2), dict(tracers=("c",), closure=(1e-3, 1e-3), bcs=BCS_STRETCHED)),
    # Smagorinsky closures: eddy viscosity field νₑ, variable-coefficient viscous / diffusive fluxes
    ((12, 10, 9), (P, P, B), None, ("WENO", 5), dict(buoyancy="tracer", tracers=("b",), closure=("lilly", 0.16, 1.0, 1.0))),
    ((10, 9, 8), (P, P, B), "stretched", ("WENO", 5), dict(buoyancy="seawater", tracers=("T", "S"), coriolis=1e-2,
                                                           closure=("lilly", 0.23, 1.0, {"T": 1.0, "S": 2.0}), bcs=BCS_LES)),
    ((9, 8, 10), (B, B, B), None, ("Centered", 2), dict(tracers=("c",), closure=("smagorinsky", 0.16, 0.7))),
    ((12, 10, 9), (P, P, P), None, ("WENO", 5), dict(closure=("lilly", 0.16, 1.0, 1.0))),
    ((16, 12), (P, F, B), None, ("WENO", 5), dict(buoyancy="tracer", tracers=("b",), closure=("lilly", 0.16, 1.0, 1.0))),
    # AnisotropicMinimumDissipation: νₑ and one κₑ per tracer
    ((12, 10, 9), (P, P, B), None, ("WENO", 5), dict(buoyancy="tracer", tracers=("b",), closure=("amd", 1 / 12, None, None))),
    ((10, 9, 8), (P, P, B), "stretched", ("WENO", 5), dict(buoyancy="seawater", tracers=("T", "S"), coriolis=1e-2,
                                                           closure=("amd", 1 / 12, 0.1, {"T": 1 / 12, "S": 1 / 6}), bcs=BCS_LES)),
    ((9, 8, 10), (B, B, B), None, ("Centered", 2), dict(tracers=("c",), closure=("amd", 1 / 12, None, None))),
    ((12, 10, 9), (P, P, P), None, ("WENO", 5), dict(closure=("amd", 1 / 12, None, None))),
    ((16, 12), (P, F, B), None, ("WENO", 5), dict(buoyancy="tracer", tracers=("b",), closure=("amd", 1 / 12, None, None))),
    # ScalarDiffusivity(VerticallyImplicitTimeDiscretization(), ν, κ): explicit tendencies keep the wall fluxes and -(ν ∂x,y w)
    ((12, 10, 9), (P, P, B), None, ("WENO", 5), dict(buoyancy="seawater", tracers=("T", "S"), coriolis=1e-2, closure=(1e-2, 2e-2, True), bcs=BCS_LES)),
    ((9, 8, 10), (B, B, B), None, ("Centered", 2), dict(tracers=("b",), closure=(1e-2, 1e-2, True), bcs=BCS_BOX)),
    ((10, 9, 8), (P, P, B), "stretched", ("WENO", 5), dict(buoyancy="tracer", tracers=("b",), closure=(1e-3, 2e-3, True), bcs=BCS_STRETCHED_B)),
    ((16, 12), (P, F, B), None, ("WENO", 5), dict(buoyancy="tracer", tracers=("b",), closure=(1e-2, 1e-2, True))),
    # array-valued conditions (condition[i, j] over the wall's transverse grid range) on every wall pair
    ((9, 8, 10), (B, B, B), None, ("WENO", 5), dict(buoyancy="tracer", tracers=("b",), closure=(1e-2, 1e-2), bcs=_array_bcs((9, 8, 10), "box"))),
    ((10, 9, 8), (P, P, B), "stretched", ("Centered", 2), dict(tracers=("c",), closure=(1e-3, 1e-3), bcs=_array_bcs((10, 9, 8), "z"))),
    ((12, 10, 9), (P, P, B), None, ("WENO", 5), dict(buoyancy="seawater", tracers=("T", "S"), closure=("lilly", 0.16, 1.0, 1.0), bcs=_array_bcs((12, 10, 9), "les"))),
]


def _z(zkind, size):
    if zkind == "stretched":
        Nz = size[-1]
        k = np.arange(1, Nz + 2)
        return np.tanh(1.3 * (2 * (k - 1) / Nz - 1)) / np.tanh(1.3)
    return None


@pytest.mark.parametrize("FT", [np.float64, np.float32])
@pytest.mark.parametrize("cfg", CONFIGS, ids=[f"{c[1]}-{c[3]}-{c[2]}-{c[0]}" + ("-bcs" if "bcs" in c[4] else "") for c in CONFIGS])
def test_tendencies_and_halos_bit_exact(oc, FT, cfg):
    size, topo, zkind, adv, kw = cfg
    m1, m2 = make_pair(oc, FT, size, topo, z=_z(zkind, size), advection=adv, **kw)
    rng = np.random.default_rng(42)
    ics = random_ics(m2, rng)
    for n, f1, f2 in names_fields(m1, m2):
        f1.set(ics[n])
        f2.set(ics[n])
    oc.update_state(m1, compute_tendencies=True)
    o_update_state(m2, compute_tendencies=True)
    if m2.nu_e is not None:
        np.testing.assert_array_equal(m1.diffusivity_fields.nu_e.parent(), m2.nu_e.data, err_msg="nu_e")
    if m2.kappa_e is not None:
        for n, f2 in m2.kappa_e.items():
            np.testing.assert_array_equal(m1.diffusivity_fields.kappa_e[n].parent(), f2.data, err_msg=f"kappa_e[{n}]")
    for n, f1, f2 in names_fields(m1, m2):
        np.testing.assert_array_equal(f1.parent(), f2.data, err_msg=f"parent({n}) after halo fill")
        G1 = m1.timestepper.Gn[n].interior()
        G2 = m2.Gn[n].interior
        np.testing.assert_array_equal(G1, G2, err_msg=f"G{n}")
    if m2.pHY is not None:
        np.testing.assert_array_equal(m1.pressures.pHY.parent(), m2.pHY.data, err_msg="pHY")


@pytest.mark.parametrize("FT", [np.float64, np.float32])
@pytest.mark.parametrize("topo", [(P, P, P), (P, P, B), (B, B, B)])
def test_substeps_and_normal_velocity_fill_bit_exact(oc, FT, topo):
    m1, m2 = make_pair(oc, FT, (8, 7, 6), topo, tracers=("c",))
    rng = np.random.default_rng(3)
    ics = random_ics(m2, rng)
    for n, f1, f2 in names_fields(m1, m2):
        f1.set(ics[n]); f2.set(ics[n])
    oc.update_state(m1); o_update_state(m2)
    oc.rk3_substep(m1, 0.01, m2.rk3["g1"], None); o_rk3(m2, 0.01, m2.rk3["g1"], None)
    oc.cache_previous_tendencies(m1)
    from oracle.model import cache_previous_tendencies as o_cache
    o_cache(m2)
    oc.update_state(m1); o_update_state(m2)
    oc.rk3_substep(m1, 0.01, m2.rk3["g2"], m2.rk3["z2"]); o_rk3(m2, 0.01, m2.rk3["g2"], m2.rk3["z2"])
    oc.ab2_step(m1, 0.02, 0.1); o_ab2(m2, 0.02, 0.1)
    oc.ab2_step(m1, 0.02, -0.5); o_ab2(m2, 0.02, -0.5)
    oc.fill_halo_regions(list(m1.velocities.values())); o_fill(list(m2.velocities))
    for n, f1, f2 in names_fields(m1, m2):
        np.testing.assert_array_equal(f1.parent(), f2.data, err_msg=n)


TOPOS3 = list(itertools.product((P, B), repeat=3))


@pytest.mark.parametrize("FT,tol", [(np.float64, 1e-12), (np.float32, 1e-5)])
@pytest.mark.parametrize("topo", TOPOS3 + [(P, P, F), (P, F, B), (F, B, B)])
def test_pressure_correction_matches_oracle(oc, FT, tol, topo):
    size3 = (16, 12, 10)
    size = tuple(n for n, t in zip(size3, topo) if t != F)
    m1, m2 = make_pair(oc, FT, size, topo)
    rng = np.random.default_rng(5)
    ics = random_ics(m2, rng)
    for n, f1, f2 in names_fields(m1, m2):
        f1.set(ics[n]); f2.set(ics[n])
    oc.calculate_pressure_correction(m1, 0.37); o_calc_pc(m2, 0.37)
    assert rel_l2(m1.pressures.pNHS.parent(), m2.pNHS.data) < tol * 10
    oc.pressure_correct_velocities(m1, 0.37); o_pcv(m2, 0.37)
    for n, f1, f2 in names_fields(m1, m2):
        assert rel_l2(f1.interior(), f2.interior) < tol, n
    oc.fill_halo_regions(list(m1.velocities.values()))
    div_tol = 1e-11 if FT == np.float64 else 5e-3
    assert m1.max_abs_divergence() < div_tol


def _run_steps(oc, m1, m2, ics, dt, nsteps, m3=None):
    """Step CUDA (m1) and oracle (m2) from the same ICs; returns rel. L2 differences per field.  With m3 (a second oracle
    model) also returns the oracle's own sensitivity to a 1-ulp perturbation of the ICs — the floor below which no two
    correct implementations with different FFT roundings can be expected to agree."""
    oc.set_model(m1, **ics)
    o.set_model(m2, **ics)
    for _ in range(nsteps):
        o.time_step(m2, dt)
    oc.time_step(m1, dt, n_steps=nsteps)
    out = {}
    for n, f1, f2 in names_fields(m1, m2):
        out[n] = float(rel_l2(f1.interior(), f2.interior))
    out["p"] = float(rel_l2(m1.pressures.pNHS.interior(), m2.pNHS.interior))
    if m3 is None:
        return out
    rng = np.random.default_rng(99)
    FT = m2.grid.FT
    pert = {n: (np.asarray(v, dtype=FT) * (1 + np.finfo(FT).eps * rng.choice([-1, 1], np.shape(v)))).astype(FT) for n, v in ics.items()}
    o.set_model(m3, **pert)
    for _ in range(nsteps):
        o.time_step(m3, dt)
    floor = {n: float(rel_l2(f3.interior, f2.interior)) for n, f3, f2 in
             zip(m2.prognostic_names(), m3.prognostic_fields(), m2.prognostic_fields())}
    floor["p"] = float(rel_l2(m3.pNHS.interior, m2.pNHS.interior))
    return out, floor


def check_parity(err, tol, floor=None):
    """rel. L2 ≤ tol on every field; where the oracle's own 1-ulp sensitivity exceeds tol/10 the bound is 10× that."""
    bad, bounds = {}, {}
    for k, e in err.items():
        bound = tol if floor is None else max(tol, 10 * floor[k])
        bounds[k] = bound
        if not e <= bound:
            bad[k] = (e, bound)
    parity_log.record(err, tol, floor, bounds)
    assert not bad, f"parity failed {bad}; errors {err}; 1-ulp floor {floor}"


def smooth_ics(m2, rng, amp=1.0, noise=1e-3):
    """Σ of random-phase low-wavenumber sinusoids + small noise."""
    g = m2.grid
    ics = {}
    for n, f in zip(m2.prognostic_names(), m2.prognostic_fields()):
        shape = f.size
        ax = [np.arange(s) / max(g.N[d], 1) for d, s in enumerate(shape)]
        v = np.zeros(shape)
        for _ in range(6):
            k = rng.integers(0, 3, 3)
            ph = rng.uniform(0, 2 * np.pi)
            v += rng.uniform(0.2, 1) * np.cos(2 * np.pi * (k[0] * ax[0][:, None, None] + k[1] * ax[1][None, :, None]
                                                          + k[2] * ax[2][None, None, :]) + ph)
        v = amp * v / 3 + noise * rng.uniform(-1, 1, shape)
        if n == "T":
            v = 20 + 0.01 * v
        if n == "S":
            v = 35 + 0.01 * v
        ics[n] = v.astype(g.FT)
    return ics


@pytest.mark.parametrize("FT,tol", [(np.float64, 1e-12), (np.float32, 1e-5)])
def test_100_steps_triply_periodic_weno5_rk3(oc, FT, tol):
    """BASELINE config 2 at an oracle-sized grid: u, v, w, p after 100 RK3 steps."""
    N = 24
    m1, m2 = make_pair(oc, FT, (N, N, N), (P, P, P), extent=(2 * np.pi,) * 3)
    ics = smooth_ics(m2, np.random.default_rng(1234))
    dt = 0.2 * (2 * np.pi / N) / 2.0
    err = _run_steps(oc, m1, m2, ics, dt, 100)
    check_parity(err, tol)
    assert m1.clock.iteration == 100 and abs(m1.clock.time - m2.clock.time) < 1e-12


@pytest.mark.parametrize("timestepper", ["RungeKutta3", "QuasiAdamsBashforth2"])
def test_field_set_then_time_step_without_set_model_on_the_fused_path(oc, timestepper):
    """set!(field, …) writes interiors only; time_step! calls update_state! at iteration 0 (runge_kutta_3.jl:81,
    quasi_adams_bashforth_2.jl:80), so the first step must see filled halos on the fused path too."""
    N = 16
    _, m2 = make_pair(oc, np.float64, (N, N, N), (P, P, P), extent=(1, 1, 1), timestepper=timestepper)
    ics = smooth_ics(m2, np.random.default_rng(7))
    # a divergence-free start (projected once by the oracle), then set field by field without the set! tail
    o.set_model(m2, **ics)
    start = {n: np.array(f.interior) for n, f in zip(m2.prognostic_names(), m2.prognostic_fields())}
    m1b, m2b = make_pair(oc, np.float64, (N, N, N), (P, P, P), extent=(1, 1, 1), timestepper=timestepper)
    for n in "uvw":
        m1b.velocities[n].set(start[n])
        m2b.prognostic_fields()["uvw".index(n)].set(start[n])
    for _ in range(3):
        o.time_step(m2b, 1e-3)
    oc.time_step(m1b, 1e-3, n_steps=3)
    for a, n in enumerate("uvw"):
        assert rel_l2(m1b.velocities[n].interior(), m2b.prognostic_fields()[a].interior) <= 1e-12, n


@pytest.mark.parametrize("FT,tol", [(np.float64, 1e-12), (np.float32, 1e-5)])
def test_100_steps_full_physics_bounded_z(oc, FT, tol):
    """u, v, w, T, S, p after 100 steps with SeawaterBuoyancy + FPlane + ScalarDiffusivity on (P,P,B)."""
    kw = dict(extent=(1, 1, 1), buoyancy="seawater", tracers=("T", "S"), coriolis=1e-4, closure=(1e-3, 1e-3))
    m1, m2 = make_pair(oc, FT, (16, 16, 16), (P, P, B), **kw)
    _, m3 = make_pair(oc, FT, (16, 16, 16), (P, P, B), **kw)
    ics = smooth_ics(m2, np.random.default_rng(1234), amp=0.5)
    err, floor = _run_steps(oc, m1, m2, ics, 1e-2, 100, m3)
    check_parity(err, tol, floor)


@pytest.mark.parametrize("timestepper", ["RungeKutta3", "QuasiAdamsBashforth2"])
@pytest.mark.parametrize("FT,tol", [(np.float64, 1e-12), (np.float32, 1e-5)])
def test_100_steps_with_value_gradient_flux_boundary_conditions(oc, FT, tol, timestepper):
    """Ocean-LES style walls (test/regression_tests/ocean_large_eddy_simulation_regression_test.jl:20-40: wind-stress
    flux on u, cooling flux on T, bottom T gradient; plus no-slip bottom) on (P,P,B) — 100 steps against the oracle."""
    kw = dict(extent=(1, 1, 1), buoyancy="seawater", tracers=("T", "S"), coriolis=1e-4, closure=(1e-3, 1e-3),
              timestepper=timestepper, bcs=BCS_LES)
    m1, m2 = make_pair(oc, FT, (16, 16, 16), (P, P, B), **kw)
    _, m3 = make_pair(oc, FT, (16, 16, 16), (P, P, B), **kw)
    ics = smooth_ics(m2, np.random.default_rng(99), amp=0.5)
    err, floor = _run_steps(oc, m1, m2, ics, 1e-2, 100, m3)
    check_parity(err, tol, floor)


@pytest.mark.parametrize("timestepper", ["RungeKutta3", "QuasiAdamsBashforth2"])
@pytest.mark.parametrize("FT,tol", [(np.float64, 1e-12), (np.float32, 1e-5)])
def test_100_steps_smagorinsky_lilly_les(oc, FT, tol, timestepper):
    """The ocean-LES regression set-up's physics (ocean_large_eddy_simulation_regression_test.jl:20-60: SmagorinskyLilly
    C = 0.23, wind stress + cooling + bottom gradient, SeawaterBuoyancy, FPlane) on 16³ — 100 steps against the oracle."""
    kw = dict(extent=(1, 1, 1), buoyancy="seawater", tracers=("T", "S"), coriolis=1e-4, closure=("lilly", 0.23, 1.0, 1.0),
              timestepper=timestepper, bcs=BCS_LES)
    m1, m2 = make_pair(oc, FT, (16, 16, 16), (P, P, B), **kw)
    _, m3 = make_pair(oc, FT, (16, 16, 16), (P, P, B), **kw)
    ics = smooth_ics(m2, np.random.default_rng(5), amp=0.5)
    err, floor = _run_steps(oc, m1, m2, ics, 1e-2, 100, m3)
    check_parity(err, tol, floor)


@pytest.mark.parametrize("zkind", [None, "stretched"])
@pytest.mark.parametrize("timestepper", ["RungeKutta3", "QuasiAdamsBashforth2"])
@pytest.mark.parametrize("FT,tol", [(np.float64, 1e-12), (np.float32, 1e-5)])
def test_100_steps_vertically_implicit_diffusion(oc, FT, tol, timestepper, zkind):
    """ScalarDiffusivity(VerticallyImplicitTimeDiscretization(), ν, κ) with the ocean-LES wall set on (P,P,B) at a vertical
    diffusion number ν·Δt/Δz² = 2.56 (unstable if explicit; horizontal 0.04) — 100 steps against the oracle (implicit_step!
    after every substep, vertically_implicit_diffusion_solver.jl:231-273).  The walls carry Flux and Gradient conditions only:
    the reference keeps the wall fluxes explicit, so a Value wall is bound by the explicit limit."""
    bcs = {"u": {"top": ("flux", -1e-4)}, "v": {"bottom": ("gradient", 0.1)},
           "T": {"top": ("flux", 1e-3), "bottom": ("gradient", 0.01)}, "S": {"top": ("flux", -2e-4)}}
    kw = dict(buoyancy="seawater", tracers=("T", "S"), coriolis=1e-4, closure=(1.0, {"T": 1.0, "S": 0.5}, True),
              timestepper=timestepper, bcs=bcs)
    size = (16, 16, 16)
    if zkind:
        kw.update(z=_z(zkind, size), x=(0, 4), y=(0, 4))
    else:
        kw.update(extent=(4, 4, 0.5))
    m1, m2 = make_pair(oc, FT, size, (P, P, B), **kw)
    _, m3 = make_pair(oc, FT, size, (P, P, B), **kw)
    ics = smooth_ics(m2, np.random.default_rng(17), amp=0.5)
    err, floor = _run_steps(oc, m1, m2, ics, 2.5e-3, 100, m3)
    check_parity(err, tol, floor)


@pytest.mark.parametrize("FT", [np.float64, np.float32])
@pytest.mark.parametrize("timestepper", ["RungeKutta3", "QuasiAdamsBashforth2"])
def test_vertically_implicit_substep_bit_exact(oc, FT, timestepper):
    """One substep + implicit_step! of every field (no FFT in between): bit-exact against the oracle, F64 and F32 (the F32 RK3
    coefficients are formed in Float64 from the caller's Float64 Δt, the AB2 ones in Float32)."""
    from oracle.model import rk3_substep, ab2_step
    kw = dict(extent=(1, 1, 1), tracers=("b",), closure=(1.0, 0.5, True), timestepper=timestepper, bcs=BCS_BOX)
    m1, m2 = make_pair(oc, FT, (9, 8, 10), (B, B, B), **kw)
    ics = random_ics(m2, np.random.default_rng(3))
    for n, f1, f2 in names_fields(m1, m2):
        f1.set(ics[n])
        f2.set(ics[n])
    oc.update_state(m1, compute_tendencies=True)
    o_update_state(m2, compute_tendencies=True)
    if timestepper == "RungeKutta3":
        c = m2.rk3
        oc.rk3_substep(m1, 0.013, c["g1"], None)
        rk3_substep(m2, 0.013, c["g1"], None)
        oc.rk3_substep(m1, 0.013, c["g2"], c["z2"])
        rk3_substep(m2, 0.013, c["g2"], c["z2"])
    else:
        oc.ab2_step(m1, 0.013, 0.1)
        ab2_step(m2, 0.013, 0.1)
    for n, f1, f2 in names_fields(m1, m2):
        np.testing.assert_array_equal(f1.interior(), f2.interior, err_msg=n)


@pytest.mark.parametrize("timestepper", ["RungeKutta3", "QuasiAdamsBashforth2"])
@pytest.mark.parametrize("FT,tol", [(np.float64, 1e-12), (np.float32, 1e-5)])
def test_100_steps_with_array_boundary_conditions(oc, FT, tol, timestepper):
    """The ocean-LES wall set with array conditions (a wind-stress pattern, a cooling pattern, a bottom-gradient map;
    boundary_condition.jl:118) on (P,P,B) — 100 steps against the oracle."""
    kw = dict(extent=(1, 1, 1), buoyancy="seawater", tracers=("T", "S"), coriolis=1e-4, closure=(1e-3, 1e-3),
              timestepper=timestepper, bcs=_array_bcs((16, 16, 16), "les"))
    m1, m2 = make_pair(oc, FT, (16, 16, 16), (P, P, B), **kw)
    _, m3 = make_pair(oc, FT, (16, 16, 16), (P, P, B), **kw)
    ics = smooth_ics(m2, np.random.default_rng(77), amp=0.5)
    err, floor = _run_steps(oc, m1, m2, ics, 1e-2, 100, m3)
    check_parity(err, tol, floor)


def test_array_boundary_condition_errors(oc):
    g = oc.RectilinearGrid(np.float64, size=(8, 6, 4), extent=(1, 1, 1), topology=(P, P, B))
    with pytest.raises(ValueError, match="shape"):
        oc.NonhydrostaticModel(g, tracers=("c",), boundary_conditions={
            "c": oc.FieldBoundaryConditions(top=oc.FluxBoundaryCondition(np.zeros((6, 8))))})


def test_vertically_implicit_needs_bounded_z(oc):
    g = oc.RectilinearGrid(np.float64, size=(8, 8, 8), extent=(1, 1, 1), topology=(P, P, P))
    with pytest.raises(Exception, match="Bounded in the z-direction"):
        oc.NonhydrostaticModel(g, closure=oc.ScalarDiffusivity(1e-2, 1e-2, vertically_implicit=True))


@pytest.mark.parametrize("FT,tol", [(np.float64, 1e-12), (np.float32, 1e-5)])
def test_100_steps_amd_les(oc, FT, tol):
    """AnisotropicMinimumDissipation with the ocean-LES wall set (the closure of the reference's
    ocean_large_eddy_simulation_regression_test.jl:14-16 AMD variant) on 16³ — 100 RK3 steps against the oracle."""
    kw = dict(extent=(1, 1, 1), buoyancy="seawater", tracers=("T", "S"), coriolis=1e-4, closure=("amd", 1 / 12, None, None),
              bcs=BCS_LES)
    m1, m2 = make_pair(oc, FT, (16, 16, 16), (P, P, B), **kw)
    _, m3 = make_pair(oc, FT, (16, 16, 16), (P, P, B), **kw)
    ics = smooth_ics(m2, np.random.default_rng(6), amp=0.5)
    err, floor = _run_steps(oc, m1, m2, ics, 1e-2, 100, m3)
    check_parity(err, tol, floor)


def test_boundary_condition_errors(oc):
    """Construction-time errors like the reference's: walls only in Bounded directions, wall-normal velocity stays Open."""
    g = oc.RectilinearGrid(np.float64, size=(8, 8, 8), extent=(1, 1, 1), topology=(P, P, B))
    with pytest.raises(oc.NhError):
        oc.NonhydrostaticModel(g, boundary_conditions={"u": oc.FieldBoundaryConditions(west=oc.ValueBoundaryCondition(0))})
    with pytest.raises(oc.NhError):
        oc.NonhydrostaticModel(g, boundary_conditions={"w": oc.FieldBoundaryConditions(top=oc.ValueBoundaryCondition(0))})
    with pytest.raises(ValueError):
        oc.NonhydrostaticModel(g, boundary_conditions={"T": oc.FieldBoundaryConditions(top=oc.FluxBoundaryCondition(1))})


@pytest.mark.parametrize("timestepper", ["RungeKutta3", "QuasiAdamsBashforth2"])
def test_steps_stretched_z_fourier_tridiagonal(oc, timestepper):
    Nz = 16
    k = np.arange(1, Nz + 2)
    z = np.tanh(1.3 * (2 * (k - 1) / Nz - 1)) / np.tanh(1.3)
    kw = dict(z=z, timestepper=timestepper, buoyancy="seawater", tracers=("T", "S"), closure=(1e-3, 1e-3))
    m1, m2 = make_pair(oc, np.float64, (16, 12, Nz), (P, P, B), **kw)
    _, m3 = make_pair(oc, np.float64, (16, 12, Nz), (P, P, B), **kw)
    ics = smooth_ics(m2, np.random.default_rng(7), amp=0.5)
    err, floor = _run_steps(oc, m1, m2, ics, 1e-2, 30, m3)
    check_parity(err, 1e-12, floor)
    oc.fill_halo_regions(list(m1.velocities.values()))
    assert m1.max_abs_divergence() < 5e-8


def test_thermal_bubble_regression_setup_ab2(oc):
    """test/regression_tests/thermal_bubble_regression_test.jl:4-35 set-up (16³, AB2, FPlane, SeawaterBuoyancy,
    ν = κ = 4e-2, Δt = 6, 10 steps) — oracle vs CUDA since the golden file is a remote DataDep."""
    N, L = 16, 100
    kw = dict(extent=(L, L, L), advection=("Centered", 2), timestepper="QuasiAdamsBashforth2", buoyancy="seawater",
              tracers=("T", "S"), coriolis=1e-4, closure=(4e-2, 4e-2))
    m1, m2 = make_pair(oc, np.float64, (N, N, N), (P, P, B), **kw)
    _, m3 = make_pair(oc, np.float64, (N, N, N), (P, P, B), **kw)
    T0 = np.full((N, N, N), 9.85)
    T0[N // 4:3 * N // 4, N // 4:3 * N // 4, N // 4:3 * N // 4] += 0.01
    ics = dict(T=T0, S=np.full((N, N, N), 35.0))
    err, floor = _run_steps(oc, m1, m2, ics, 6.0, 10, m3)
    check_parity(err, 1e-12, floor)


@pytest.mark.parametrize("FT", [np.float64, np.float32])
@pytest.mark.parametrize("timestepper", ["RungeKutta3", "QuasiAdamsBashforth2"])
@pytest.mark.parametrize("size", [(40, 24, 16), (33, 13, 9)])
def test_fused_fast_path_is_bit_identical_to_complete_path(oc, FT, timestepper, size, monkeypatch):
    """The fused tendency+substep / pressure-correction+halo kernels against the one-kernel-per-reference-kernel path."""
    import os
    rng = np.random.default_rng(8)
    ics = {n: rng.uniform(-1, 1, size).astype(FT) for n in "uvw"}
    out = []
    for no_fast in (False, True):
        if no_fast:
            monkeypatch.setenv("NHSTEP_NO_FAST", "1")
        else:
            monkeypatch.delenv("NHSTEP_NO_FAST", raising=False)
        g = oc.RectilinearGrid(FT, size=size, extent=(1, 2, 3), topology=(P, P, P))
        m = oc.NonhydrostaticModel(g, advection=oc.WENO(order=5), timestepper=timestepper)
        oc.set_model(m, **ics)
        oc.time_step(m, 2e-3, n_steps=3)
        out.append([m.velocities[n].parent() for n in "uvw"] + [m.pressures.pNHS.parent()]
                   + [m.timestepper.Gn[n].interior() for n in "uvw"] + [m.timestepper.Gm[n].interior() for n in "uvw"])
    for a, b in zip(*out):
        np.testing.assert_array_equal(a, b)


@pytest.mark.parametrize("exact", [True, False], ids=["exact", "contracted"])
@pytest.mark.parametrize("FT", [np.float64, np.float32])
@pytest.mark.parametrize("timestepper", ["RungeKutta3", "QuasiAdamsBashforth2"])
@pytest.mark.parametrize("size", [(40, 24, 16), (33, 13, 9), (64, 64, 40)])
def test_plane_marching_tma_sweep_is_bit_identical_to_the_tile_sweep(oc, FT, timestepper, size, exact, monkeypatch):
    """The two fused tendency + substep sweeps — the plane-marching kernel that stages u, v, w planes by TMA
    (csrc/tendency_plane.cuh, the default in contracted arithmetic) and the tile kernel (csrc/tendency_fast.cuh, the default in
    exact arithmetic) — evaluate the same flux functions once per face, so they must agree bit for bit in both arithmetic
    modes, on partial tiles and partial z segments too."""
    rng = np.random.default_rng(5)
    ics = {n: rng.uniform(-1, 1, size).astype(FT) for n in "uvw"}
    out = {}
    for which, var in (("plane", "NHSTEP_PLANE_SWEEP"), ("tile", "NHSTEP_TILE_SWEEP")):
        monkeypatch.setenv(var, "1")
        g = oc.RectilinearGrid(FT, size=size, extent=(1, 1, 1), topology=(P, P, P))
        m = oc.NonhydrostaticModel(g, advection=oc.WENO(order=5), timestepper=timestepper, exact_arithmetic=exact)
        monkeypatch.delenv(var)
        oc.set_model(m, **ics)
        oc.time_step(m, 1e-3, n_steps=3)
        out[which] = [m.velocities[n].interior() for n in "uvw"] + [m.pressures.pNHS.interior()] + \
                     [m.timestepper.Gn[n].interior() for n in "uvw"]
    for a, b in zip(out["plane"], out["tile"]):
        np.testing.assert_array_equal(a, b)


@pytest.mark.parametrize("FT", [np.float64, np.float32])
@pytest.mark.parametrize("cfg", [((37, 9, 11), (P, P, B), "stretched", ("WENO", 5),
                                  dict(buoyancy="seawater", tracers=("T", "S"), coriolis=1e-2, closure=("lilly", 0.16, 1.0, 1.0))),
                                 ((9, 8, 10), (B, B, B), None, ("UpwindBiased", 3), dict(tracers=("c",), closure=(1e-2, 1e-2))),
                                 ((33, 6, 1), (P, B, B), None, ("Centered", 4), {}),
                                 ((16, 12), (P, F, B), None, ("WENO", 5), dict(buoyancy="tracer", tracers=("b",)))],
                         ids=["ppb-les", "bbb", "nz1", "flat"])
def test_tiled_tendency_kernel_is_bit_identical_to_the_per_cell_kernel(oc, FT, cfg, monkeypatch):
    """Complete path: faces evaluated once into shared memory (default) against the reference-style kernel that evaluates
    every face from both neighbouring cells (NHSTEP_NO_TILED), partial tiles and walls included."""
    size, topo, zkind, adv, kw = cfg
    out = []
    rng = np.random.default_rng(3)
    ics = None
    for no_tiled in (False, True):
        if no_tiled:
            monkeypatch.setenv("NHSTEP_NO_TILED", "1")
        else:
            monkeypatch.delenv("NHSTEP_NO_TILED", raising=False)
        m1, m2 = make_pair(oc, FT, size, topo, z=_z(zkind, size), advection=adv, **kw)
        if ics is None:
            ics = random_ics(m2, rng)
        for n, f1, _ in names_fields(m1, m2):
            f1.set(ics[n])
        oc.update_state(m1, compute_tendencies=True)
        out.append([m1.timestepper.Gn[n].interior() for n in m2.prognostic_names()])
    for a, b in zip(*out):
        np.testing.assert_array_equal(a, b)


@pytest.mark.parametrize("FT,tol", [(np.float64, 1e-12), (np.float32, 1e-5)])
@pytest.mark.parametrize("size", [(32, 32, 32), (64, 16, 32)])
def test_fused_poisson_path_100_steps_vs_oracle(oc, FT, tol, size):
    """Power-of-two triply periodic grids take the fused real-to-complex FFT solve (poisson_fast.cuh)."""
    m1, m2 = make_pair(oc, FT, size, (P, P, P), extent=(2 * np.pi,) * 3)
    ics = smooth_ics(m2, np.random.default_rng(1234))
    dt = 0.2 * (2 * np.pi / max(size)) / 2.0
    err = _run_steps(oc, m1, m2, ics, dt, 100)
    check_parity(err, tol)
    oc.fill_halo_regions(list(m1.velocities.values()))
    assert m1.max_abs_divergence() < (1e-11 if FT == np.float64 else 5e-3)


@pytest.mark.parametrize("FT,tol", [(np.float64, 1e-12), (np.float32, 1e-5)])
@pytest.mark.parametrize("timestepper", ["RungeKutta3", "QuasiAdamsBashforth2"])
@pytest.mark.parametrize("size", [(24, 24, 24), (64, 16, 32), (40, 24, 16)])
def test_contracted_arithmetic_100_steps_vs_oracle(oc, FT, tol, timestepper, size):
    """exact_arithmetic=False (tendency_fma.cu: FMA contraction, regrouped WENO5 weights, the arithmetic a GPU build of the
    reference produces) stays within the north_star tolerance of the oracle after 100 steps."""
    m1, m2 = make_pair(oc, FT, size, (P, P, P), extent=(2 * np.pi,) * 3, timestepper=timestepper, exact_arithmetic=False)
    _, m3 = make_pair(oc, FT, size, (P, P, P), extent=(2 * np.pi,) * 3, timestepper=timestepper)
    ics = smooth_ics(m2, np.random.default_rng(1234))
    dt = 0.2 * (2 * np.pi / max(size)) / 2.0
    err, floor = _run_steps(oc, m1, m2, ics, dt, 100, m3)
    check_parity(err, tol, floor)
    oc.fill_halo_regions(list(m1.velocities.values()))
    assert m1.max_abs_divergence() < (1e-11 if FT == np.float64 else 5e-3)


@pytest.mark.parametrize("FT,tol", [(np.float64, 1e-12), (np.float32, 1e-5)])
def test_contracted_arithmetic_complete_path_100_steps_vs_oracle(oc, FT, tol):
    """exact_arithmetic = 0 on the complete path (the tiled all-fields tendency kernel with the quotient-free WENO5 weights on
    its interior tiles, compiled with FMA contraction): u, v, w, T, S, p after 100 steps of the full-physics (P,P,B) model stay
    within the north_star tolerance of the oracle (10x the oracle's own 1-ulp sensitivity where that is larger)."""
    kw = dict(extent=(1, 1, 1), buoyancy="seawater", tracers=("T", "S"), coriolis=1e-4, closure=(1e-3, 1e-3))
    size = (32, 16, 32)                       # tiles k0 = 9, 17 are interior: the fast flux stage runs
    m1, m2 = make_pair(oc, FT, size, (P, P, B), exact_arithmetic=False, **kw)
    _, m3 = make_pair(oc, FT, size, (P, P, B), **kw)
    ics = smooth_ics(m2, np.random.default_rng(1234), amp=0.5)
    err, floor = _run_steps(oc, m1, m2, ics, 5e-3, 100, m3)
    check_parity(err, tol, floor)


def test_contracted_arithmetic_is_a_different_kernel_within_rounding(oc):
    """One tendency evaluation in both modes: not bit-identical (the contracted build really ran), equal to ~1e-13."""
    size = (40, 24, 16)
    rng = np.random.default_rng(8)
    ics = {n: rng.uniform(-1, 1, size) for n in "uvw"}
    G = []
    for exact in (True, False):
        g = oc.RectilinearGrid(np.float64, size=size, extent=(1, 2, 3), topology=(P, P, P))
        m = oc.NonhydrostaticModel(g, advection=oc.WENO(order=5), exact_arithmetic=exact)
        oc.set_model(m, **ics)
        oc.time_step(m, 1e-3)
        G.append([m.timestepper.Gn[n].interior() for n in "uvw"])
    for a, b in zip(*G):
        assert not np.array_equal(a, b)
        assert rel_l2(b, a) < 1e-13


def test_fused_poisson_matches_complete_solver(oc, monkeypatch):
    """Same model with and without the fused solve: fields agree to rounding after 5 steps (different FFT factorisation)."""
    size = (64, 32, 16)
    rng = np.random.default_rng(8)
    ics = {n: rng.uniform(-1, 1, size) for n in "uvw"}
    out = []
    for no_fast in (False, True):
        if no_fast:
            monkeypatch.setenv("NHSTEP_NO_FAST_POISSON", "1")
        else:
            monkeypatch.delenv("NHSTEP_NO_FAST_POISSON", raising=False)
        g = oc.RectilinearGrid(np.float64, size=size, extent=(1, 2, 3), topology=(P, P, P))
        m = oc.NonhydrostaticModel(g, advection=oc.WENO(order=5))
        oc.set_model(m, **ics)
        oc.time_step(m, 2e-3, n_steps=5)
        out.append([m.velocities[n].interior() for n in "uvw"] + [m.pressures.pNHS.parent()])
    for a, b in zip(*out):
        assert rel_l2(a, b) < 1e-12


@pytest.mark.parametrize("stretched", [False, True], ids=["dct", "tridiagonal"])
@pytest.mark.parametrize("timestepper", ["RungeKutta3", "QuasiAdamsBashforth2"])
def test_bounded_z_models_on_fused_solver_passes(oc, stretched, timestepper, monkeypatch):
    """(Periodic, Periodic, Bounded) models with power-of-two horizontal sizes take the fused solver passes (DCT along a
    regular z, Thomas sweep along a stretched z) inside the complete path: against the oracle and against the complete
    solver, Centered + closure + Coriolis physics, 30 steps."""
    Nz = 16 if not stretched else 12
    k = np.arange(1, Nz + 2)
    z = np.tanh(1.3 * (2 * (k - 1) / Nz - 1)) / np.tanh(1.3) if stretched else (0, 1)
    kw = dict(z=z, timestepper=timestepper, advection=("Centered", 4), coriolis=1e-2, closure=(1e-3, 1e-3))
    size = (32, 16, Nz)
    m1, m2 = make_pair(oc, np.float64, size, (P, P, B), **kw)
    monkeypatch.setenv("NHSTEP_NO_FAST_POISSON", "1")
    m3, _ = make_pair(oc, np.float64, size, (P, P, B), **kw)
    monkeypatch.delenv("NHSTEP_NO_FAST_POISSON")
    counts = []
    for m in (m1, m3):
        base = m.launch_count()
        oc.calculate_pressure_correction(m, 1.0)
        counts.append(m.launch_count() - base)
    assert counts[0] < counts[1], f"the fused passes were not taken: {counts}"
    ics = smooth_ics(m2, np.random.default_rng(21), amp=0.5)
    err = _run_steps(oc, m1, m2, ics, 5e-3, 30)
    check_parity(err, 1e-12)
    oc.set_model(m3, **ics)
    oc.time_step(m3, 5e-3, n_steps=30)
    for n in "uvw":
        assert rel_l2(m1.velocities[n].interior(), m3.velocities[n].interior()) < 1e-12
    assert rel_l2(m1.pressures.pNHS.interior(), m3.pressures.pNHS.interior()) < 1e-11
