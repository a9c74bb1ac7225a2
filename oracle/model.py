"""NonhydrostaticModel, update_state!, time_step! (RK3 / AB2), pressure correction (oracle; test infrastructure only).

Follows the reference's
  src/Models/NonhydrostaticModels/nonhydrostatic_model.jl:114-257         (constructor: pHY′ iff buoyancy, halo inflation,
                                                                            order adaptation, solver choice, update_state!)
  src/Models/NonhydrostaticModels/NonhydrostaticModels.jl:25-62           (solver choice by regular dimensions)
  src/Models/NonhydrostaticModels/update_nonhydrostatic_model_state.jl:20-70
  src/Models/NonhydrostaticModels/pressure_correction.jl:8-50
  src/Models/NonhydrostaticModels/solve_for_pressure.jl:12-38,57-82
  src/Models/NonhydrostaticModels/set_nonhydrostatic_model.jl:33-60
  src/TimeSteppers/runge_kutta_3.jl:48-62,77-208
  src/TimeSteppers/quasi_adams_bashforth_2.jl:39,74-175
  src/TimeSteppers/store_tendencies.jl:6-22
  src/TimeSteppers/clock.jl:87-116
"""
from fractions import Fraction

import numpy as np

from .grid import Bounded, Flat, round_fraction
from .fields import Field, fill_halo_regions
from .advection import Centered, adapt_advection_order
from .operators import divergence_ccc, deriv, delta, interior_region, area, volume
from .solvers import FFTBasedPoissonSolver, FourierTridiagonalPoissonSolver
from .tendencies import (VEL_LOC, CCC, momentum_tendency, tracer_tendency, tendency_region,
                         update_hydrostatic_pressure, compute_smagorinsky_viscosity)


SIDES = {"west": (0, 0), "east": (0, 1), "south": (1, 0), "north": (1, 1), "bottom": (2, 0), "top": (2, 1)}


OMEGA_EARTH = 7.292115e-5          # src/Coriolis/Coriolis.jl:19


class FPlane:
    """FPlane(f) or FPlane(rotation_rate=Ω_Earth, latitude=φ): f = 2·rotation_rate·sind(latitude) — src/Coriolis/f_plane.jl:25-40."""

    def __init__(self, f=None, rotation_rate=OMEGA_EARTH, latitude=None):
        if (f is None) == (latitude is None):
            raise ValueError("Either both keywords rotation_rate and latitude must be specified, *or* only f must be specified.")
        if f is None:
            import math
            # sind: exact at multiples of 30° like Julia's; elsewhere sin of the degree → radian conversion
            lat = float(latitude)
            s = {0.0: 0.0, 30.0: 0.5, 90.0: 1.0, -30.0: -0.5, -90.0: -1.0}.get(lat, math.sin(math.radians(lat)))
            f = 2 * rotation_rate * s
        self.f = f


class BuoyancyTracer:
    kind = "tracer"
    required_tracers = ("b",)


class SeawaterBuoyancy:
    """SeawaterBuoyancy with a LinearEquationOfState (defaults: g = 9.80665, α = 1.67e-4, β = 7.80e-4 —
    src/BuoyancyFormulations/seawater_buoyancy.jl, linear_equation_of_state.jl:24-37)."""
    kind = "seawater"
    required_tracers = ("T", "S")

    def __init__(self, gravitational_acceleration=9.80665, thermal_expansion=1.67e-4, haline_contraction=7.80e-4):
        self.g, self.alpha, self.beta = gravitational_acceleration, thermal_expansion, haline_contraction


class ScalarDiffusivity:
    """ScalarDiffusivity([time_discretization,] ν, κ): isotropic; vertically_implicit = VerticallyImplicitTimeDiscretization()."""
    kind = "scalar"

    def __init__(self, nu=0.0, kappa=0.0, vertically_implicit=False):
        self.nu, self.kappa, self.vitd = nu, kappa, bool(vertically_implicit)

    def kappa_for(self, name):
        return self.kappa[name] if isinstance(self.kappa, dict) else self.kappa


class SmagorinskyLilly:
    """SmagorinskyLilly(FT; C = 0.16, Cb = 1, Pr = 1) — Smagorinskys/lilly_coefficient.jl:108-113 (C, Cb, Pr → FT)."""
    kind = "smagorinsky_lilly"

    def __init__(self, C=0.16, Cb=1.0, Pr=1.0):
        self.C, self.Cb, self.Pr = C, Cb, Pr

    def Pr_for(self, name):
        return self.Pr[name] if isinstance(self.Pr, dict) else self.Pr


class Smagorinsky(SmagorinskyLilly):
    """Smagorinsky(FT; coefficient = 0.16, Pr = 1) — Smagorinskys/smagorinsky.jl:78-84,105 (constant coefficient)."""
    kind = "smagorinsky"

    def __init__(self, coefficient=0.16, Pr=1.0):
        self.C, self.Cb, self.Pr = coefficient, 0.0, Pr


class AnisotropicMinimumDissipation:
    """AnisotropicMinimumDissipation(FT; C = 1/12, Cν = C, Cκ = C (number or by tracer), Cb = nothing) —
    anisotropic_minimum_dissipation.jl:110-121."""
    kind = "amd"

    def __init__(self, C=1 / 12, Cnu=None, Ckappa=None):
        self.Cnu = C if Cnu is None else Cnu
        self.Ckappa = C if Ckappa is None else Ckappa

    def Ckappa_for(self, name):
        return self.Ckappa[name] if isinstance(self.Ckappa, dict) else self.Ckappa

    def Pr_for(self, name):       # κₑ is its own field: the face value is ℑ(κₑ), i.e. "Pr = 1"
        return 1.0


class Clock:
    def __init__(self):
        self.time = 0.0
        self.iteration = 0
        self.stage = 1
        self.last_dt = np.inf
        self.last_stage_dt = np.inf


class NonhydrostaticModel:
    def __init__(self, grid, advection="default", buoyancy=None, coriolis=None, closure=None, tracers=(),
                 timestepper="RungeKutta3", hydrostatic_pressure_anomaly="default", chi=0.1, boundary_conditions=None):
        FT = grid.FT
        if advection == "default":
            advection = Centered(FT, 2)
        tracers = (tracers,) if isinstance(tracers, str) else tuple(tracers)
        if buoyancy is not None:
            for req in buoyancy.required_tracers:
                if req not in tracers:
                    raise ValueError(f"buoyancy requires tracer {req}")
        adv = adapt_advection_order(advection, grid)
        # inflate_grid_halo_size (nonhydrostatic_model.jl:243-257, automatic_halo_sizing.jl:71-83): each non-Flat direction needs
        # the buffer of ITS adapted scheme (required_halo_size_x/y/z of the FluxFormAdvection) and 1 for the closure
        need = []
        for d in range(3):
            if grid.topology[d] == Flat:
                need.append(0)
                continue
            h = max(1, adv[d].B if adv[d] is not None else 1)
            need.append(max(h, grid.H[d]))
        if tuple(need) != tuple(grid.H):
            grid = grid.with_halo(need)
        self.grid = grid
        self.advection = adv
        self.buoyancy, self.coriolis, self.closure = buoyancy, coriolis, closure
        if getattr(closure, "vitd", False) and grid.topology[2] != Bounded:
            raise ValueError("VerticallyImplicitTimeDiscretization can only be specified on grids that are Bounded in the z-direction.")
        self.velocities = tuple(Field(grid, VEL_LOC[a], name="uvw"[a]) for a in range(3))
        self.tracers = {n: Field(grid, CCC, name=n) for n in tracers}
        # boundary_conditions = {field name: {"west"|"east"|"south"|"north"|"bottom"|"top": (kind, number)}} —
        # FieldBoundaryConditions with Value / Gradient / FluxBoundaryCondition(number) entries
        for name, sides in (boundary_conditions or {}).items():
            f = dict(zip(["u", "v", "w"] + list(tracers), list(self.velocities) + list(self.tracers.values())))[name]
            for side, (kind, value) in sides.items():
                d, s = SIDES[side]
                f.set_boundary_condition(d, s, kind, value)
        self.pNHS = Field(grid, CCC, prognostic=False, name="pNHS")
        # build_diffusivity_fields (smagorinsky.jl:123-133): νₑ at ccc with default boundary conditions
        self.nu_e = Field(grid, CCC, prognostic=False, name="nu_e") if (closure is not None and closure.kind != "scalar") else None
        # AMD: one eddy diffusivity field per tracer (anisotropic_minimum_dissipation.jl:371-385)
        self.kappa_e = ({n: Field(grid, CCC, prognostic=False, name="kappa_e_" + n) for n in tracers}
                        if (closure is not None and closure.kind == "amd") else None)
        if hydrostatic_pressure_anomaly == "default":
            self.pHY = Field(grid, CCC, prognostic=False, name="pHY") if buoyancy is not None else None
        else:
            self.pHY = Field(grid, CCC, prognostic=False, name="pHY") if hydrostatic_pressure_anomaly else None
        # solver choice: NonhydrostaticModels.jl:25-62
        irregular = [d for d in range(3) if not grid.regular_dim(d)]
        if not irregular:
            self.pressure_solver = FFTBasedPoissonSolver(grid)
        elif len(irregular) == 1:
            self.pressure_solver = FourierTridiagonalPoissonSolver(grid)
        else:
            raise ValueError("no FFT-based pressure solver for a grid stretched in more than one direction")
        self.timestepper = timestepper
        names = ["u", "v", "w"] + list(tracers)
        self.Gn = {n: f.copy() for n, f in zip(names, self.prognostic_fields())}
        self.Gm = {n: f.copy() for n, f in zip(names, self.prognostic_fields())}
        for G in list(self.Gn.values()) + list(self.Gm.values()):
            G.data[...] = 0
        # RK3 coefficients stored as FT (runge_kutta_3.jl:53-62)
        self.rk3 = dict(g1=round_fraction(Fraction(8, 15), FT), g2=round_fraction(Fraction(5, 12), FT),
                        g3=round_fraction(Fraction(3, 4), FT), z2=round_fraction(Fraction(-17, 60), FT),
                        z3=round_fraction(Fraction(-5, 12), FT))
        self.chi = FT(chi)
        self.clock = Clock()
        update_state(self, compute_tendencies=False)

    def prognostic_fields(self):
        return list(self.velocities) + list(self.tracers.values())

    def prognostic_names(self):
        return ["u", "v", "w"] + list(self.tracers.keys())


# ---------------------------------------------------------------------------------------------

def update_state(model, compute_tendencies=True):
    """update_state!: halo fill (normal velocities not re-imposed) → hydrostatic pressure → tendencies."""
    fill_halo_regions(model.prognostic_fields(), fill_boundary_normal_velocities=False)
    if model.nu_e is not None:                     # compute_auxiliaries!: diffusivities, then hydrostatic pressure
        if model.closure.kind == "amd":
            compute_amd_diffusivities(model)
        else:
            compute_smagorinsky_viscosity(model)
    update_hydrostatic_pressure(model.grid, model.pHY, model.buoyancy, model.tracers)
    if model.nu_e is not None:
        fill_halo_regions([model.nu_e] + (list(model.kappa_e.values()) if model.kappa_e else []))
    if compute_tendencies:
        compute_tendencies_(model)


def compute_amd_diffusivities(model):
    """compute_diffusivities!(…, ::AnisotropicMinimumDissipation, …) over :xyz (anisotropic_minimum_dissipation.jl:191-208)."""
    from .amd import amd_viscosity, amd_diffusivity
    grid = model.grid
    R = interior_region(grid)
    shape = model.nu_e.at(*R).shape
    model.nu_e.at(*R)[...] = np.broadcast_to(amd_viscosity(grid, model.velocities, model.closure.Cnu, R), shape)
    for n, c in model.tracers.items():
        model.kappa_e[n].at(*R)[...] = np.broadcast_to(
            amd_diffusivity(grid, model.velocities, c, model.closure.Ckappa_for(n), R), shape)


def compute_tendencies_(model):
    for a in range(3):
        R, G = momentum_tendency(model, a)
        if G is not None:
            model.Gn["uvw"[a]].at(*R)[...] = G
    for n in model.tracers:
        R, G = tracer_tendency(model, n)
        model.Gn[n].at(*R)[...] = G
    apply_flux_bcs(model)


def apply_flux_bcs(model):
    """compute_boundary_tendency_contributions! (compute_nonhydrostatic_tendencies.jl:202-207) with
    apply_{x,y,z}_bcs! of src/BoundaryConditions/apply_flux_bcs.jl:50-160: Gc[1] += q·A/V, Gc[N] -= q·A/V over the
    grid-interior transverse range; A at the boundary face (flip(loc)), V at the field's own location."""
    grid = model.grid
    for n, f in zip(model.prognostic_names(), model.prognostic_fields()):
        G = model.Gn[n]
        for d in range(3):
            for side in (0, 1):
                bc = f.side_bcs[d][side]
                if bc is None or bc[0] != "flux":
                    continue
                N = grid.N[d]
                r = [(1, grid.N[e]) for e in range(3)]
                rf = list(r)
                r[d] = (1, 1) if side == 0 else (N, N)
                rf[d] = (1, 1) if side == 0 else (N + 1, N + 1)
                floc = list(f.loc)
                floc[d] = "f" if f.loc[d] == "c" else "c"
                inc = bc[1] * area(grid, d, tuple(floc), rf) / volume(grid, f.loc, r)
                if side == 0:
                    G.at(*r)[...] += inc
                else:
                    G.at(*r)[...] -= inc


def compute_source_term(model, dt):
    """_compute_source_term! / _fourier_tridiagonal_source_term! into the solver's complex array."""
    grid = model.grid
    R = interior_region(grid)
    u, v, w = model.velocities
    div = divergence_ccc(grid, u, v, w, R)
    solver = model.pressure_solver
    dt = _as_dt(grid, dt)
    if isinstance(solver, FFTBasedPoissonSolver):
        solver.storage[...] = div / dt
    else:
        td = solver.td
        solver.source_term[...] = delta(grid, td, "c", R) * div / dt


def _as_dt(grid, dt):
    """Δt reaches the kernels as whatever the caller passed (Float64 literal, FT product…); keep numpy scalars,
    promote Python floats to Float64 like a Julia Float64 literal."""
    if isinstance(dt, (np.floating,)):
        return dt
    return np.float64(dt)


def calculate_pressure_correction(model, dt):
    fill_halo_regions(list(model.velocities))                       # with wall-normal zeroing
    compute_source_term(model, dt)
    model.pressure_solver.solve(model.pNHS)
    fill_halo_regions(model.pNHS)


def pressure_correct_velocities(model, dt):
    grid = model.grid
    dt = _as_dt(grid, dt)
    R = interior_region(grid)
    for a in range(3):
        U = model.velocities[a]
        g = deriv(grid, model.pNHS.at, a, VEL_LOC[a], R) * dt
        out = (U.at(*R) - g).astype(grid.FT)
        U.at(*R)[...] = out


def cache_previous_tendencies(model):
    R = interior_region(model.grid)
    for n in model.Gn:
        model.Gm[n].at(*R)[...] = model.Gn[n].at(*R)


def implicit_step(model, name, f, dt):
    """implicit_step!(field, solver, closure, …, Δt) — vertically_implicit_diffusion_solver.jl:57-121,231-273 with the Thomas
    sweep of batched_tridiagonal_solver.jl:209-235 on coefficient functions: (1 - Δt ∂z κ ∂z) ϕⁿ⁺¹ = ϕ★ per column, for the
    grid's Nz rows over the grid's Nx × Ny columns.  The coefficients are formed in the type of Δt·κ/Δz² — Float64 under RK3
    (stage_Δt = Δt·(γ + ζ) with the caller's Float64 Δt, runge_kutta_3.jl:157-158), eltype(grid) under AB2 (Δt is converted,
    quasi_adams_bashforth_2.jl:134) — and the scratch t has eltype(grid) (batched_tridiagonal_solver.jl:69-79)."""
    cl = model.closure
    if cl is None or not getattr(cl, "vitd", False):
        return
    grid = model.grid
    FT = grid.FT
    Nz = grid.N[2]
    cz = grid.coords[2]
    CT = np.float64 if isinstance(dt, np.float64) else FT        # coefficient type
    dt = CT(dt)

    def dzc(k):
        return CT(cz.dc if cz.regular else cz.dc[k])

    def dzf(k):
        return CT(cz.df if cz.regular else cz.df[k])
    w_field = f.loc[2] == "f"
    kap = CT(FT(cl.nu if name in "uvw" else cl.kappa_for(name)))
    if not w_field:
        def upper(k):
            return CT(0) if k > Nz - 1 else -dt * kap / (dzc(k) * dzf(k + 1))

        def lower(kp):
            k = kp + 1
            return CT(0) if kp < 1 else -dt * kap / (dzc(k) * dzf(k))
    else:
        def upper(k):
            return CT(0) if k < 1 else -dt * kap / (dzc(k) * dzf(k))

        def lower(k):
            kp = k + 2
            return CT(0) if k < 1 else -dt * kap / (dzc(kp) * dzf(kp - 1))

    def diag(k):
        return CT(1) - dt * CT(0) - upper(k) - lower(k - 1)
    phi = f.at((1, grid.N[0]), (1, grid.N[1]), (1, Nz))          # in place: ϕ and the right-hand side are the same array
    t = np.zeros(phi.shape[:2] + (Nz + 1,), dtype=FT)
    beta = diag(1)
    phi[:, :, 0] = (phi[:, :, 0] / beta).astype(FT)
    for k in range(2, Nz + 1):
        c_km1, b_k, a_km1 = upper(k - 1), diag(k), lower(k - 1)
        t[:, :, k] = (c_km1 / beta).astype(FT) if np.ndim(beta) else FT(c_km1 / beta)
        beta = b_k - a_km1 * t[:, :, k]
        fk = phi[:, :, k - 1]
        new = (fk - a_km1 * phi[:, :, k - 2]) / beta
        ok = ~(np.abs(beta) < 10 * np.finfo(beta.dtype).eps)
        phi[:, :, k - 1] = np.where(ok, new, fk).astype(FT)
    for k in range(Nz - 1, 0, -1):
        phi[:, :, k - 1] = (phi[:, :, k - 1] - t[:, :, k + 1] * phi[:, :, k]).astype(FT)


def rk3_substep(model, dt, gamma, zeta):
    FT = model.grid.FT
    dtj = _as_dt(model.grid, dt)
    stage_dt = dtj * gamma if zeta is None else dtj * (gamma + zeta)       # stage_Δt(Δt, γⁿ, ζⁿ), runge_kutta_3.jl:153-154
    dt = FT(dt)
    for n, f in zip(model.prognostic_names(), model.prognostic_fields()):
        R = tendency_region(model.grid, f.loc)
        if any(r[1] < r[0] for r in R):
            continue
        if zeta is None:
            f.at(*R)[...] += dt * gamma * model.Gn[n].at(*R)
        else:
            f.at(*R)[...] += dt * (gamma * model.Gn[n].at(*R) + zeta * model.Gm[n].at(*R))
        implicit_step(model, n, f, np.float64(stage_dt))


def ab2_step(model, dt, chi):
    FT = model.grid.FT
    dt, chi = FT(dt), FT(chi)
    not_euler = chi != FT(-0.5)
    for n, f in zip(model.prognostic_names(), model.prognostic_fields()):
        R = tendency_region(model.grid, f.loc)
        if any(r[1] < r[0] for r in R):
            continue
        if not_euler:
            G = (FT(1.5) + chi) * model.Gn[n].at(*R) - (FT(0.5) + chi) * model.Gm[n].at(*R)
        else:   # `x * false` is a strong zero in Julia: stale NaNs in G⁻ do not propagate
            G = (FT(1.5) + chi) * model.Gn[n].at(*R) - FT(0)
        f.at(*R)[...] += dt * G
        implicit_step(model, n, f, dt)


def time_step(model, dt, euler=False):
    """time_step!(model, Δt)."""
    clock = model.clock
    if clock.iteration == 0:
        update_state(model, compute_tendencies=True)
    if model.timestepper == "RungeKutta3":
        c = model.rk3
        FT = model.grid.FT
        dtj = _as_dt(model.grid, dt)
        first = c["g1"] * dtj
        second = (c["g2"] + c["z2"]) * dtj
        third = (c["g3"] + c["z3"]) * dtj
        t_next = clock.time + float(dt)
        # stage 1
        rk3_substep(model, dt, c["g1"], None)
        clock.time += float(first); clock.stage += 1; clock.last_stage_dt = float(first)
        calculate_pressure_correction(model, first)
        pressure_correct_velocities(model, first)
        cache_previous_tendencies(model)
        update_state(model)
        # stage 2
        rk3_substep(model, dt, c["g2"], c["z2"])
        clock.time += float(second); clock.stage += 1; clock.last_stage_dt = float(second)
        calculate_pressure_correction(model, second)
        pressure_correct_velocities(model, second)
        cache_previous_tendencies(model)
        update_state(model)
        # stage 3
        rk3_substep(model, dt, c["g3"], c["z3"])
        corrected = t_next - clock.time
        clock.time += float(third); clock.iteration += 1; clock.stage = 1
        clock.last_stage_dt = corrected
        clock.last_dt = float(dt)
        calculate_pressure_correction(model, third)
        pressure_correct_velocities(model, third)
        update_state(model)
    elif model.timestepper == "QuasiAdamsBashforth2":
        FT = model.grid.FT
        euler = euler or (float(dt) != clock.last_dt)
        chi = FT(-0.5) if euler else model.chi
        ab2_step(model, dt, chi)
        clock.time += float(dt); clock.iteration += 1; clock.stage = 1
        clock.last_dt = float(dt); clock.last_stage_dt = float(dt)
        calculate_pressure_correction(model, dt)
        pressure_correct_velocities(model, dt)
        cache_previous_tendencies(model)
        update_state(model)
    else:
        raise ValueError(f"unknown timestepper {model.timestepper}")


def set_model(model, enforce_incompressibility=True, **kwargs):
    """set!(model; u=…, v=…, …): arrays of the field's interior size (Nx,Ny,Nz)[+1 on Bounded faces]."""
    names = model.prognostic_names()
    for name, value in kwargs.items():
        if name not in names:
            raise ValueError(f"name {name} not found in model.velocities or model.tracers.")
        f = model.prognostic_fields()[names.index(name)]
        f.set(value)
        fill_halo_regions(f)
    update_state(model, compute_tendencies=False)
    if enforce_incompressibility:
        one = model.grid.FT(1)
        calculate_pressure_correction(model, one)
        pressure_correct_velocities(model, one)
        update_state(model, compute_tendencies=False)
