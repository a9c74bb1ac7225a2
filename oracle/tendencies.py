"""Tendency terms of the NonhydrostaticModel (oracle; test infrastructure only).

Follows the reference's
  src/Models/NonhydrostaticModels/nonhydrostatic_tendency_kernel_functions.jl:47-78,107-138,170-200,228-259
  src/Models/NonhydrostaticModels/compute_nonhydrostatic_tendencies.jl:57-139  (launch ranges; exclude_periphery)
  src/Utils/kernel_launching.jl:90-111                                          (periphery_offset)
  src/Advection/momentum_advection_operators.jl:46-83, tracer_advection_operators.jl:30-34
  src/Advection/upwind_biased_advective_fluxes.jl:23-121, centered_advective_fluxes.jl:15-33
  src/Coriolis/f_plane.jl:44-46 + src/Operators/interpolation_operators.jl:121-131 (active-weighted ℑxy)
  src/BuoyancyFormulations/g_dot_b.jl:2-4, buoyancy_tracer.jl:12, linear_equation_of_state.jl:72-80
  src/TurbulenceClosures/closure_kernel_operators.jl:22-48,
      turbulence_closure_implementations/abstract_scalar_diffusivity_closure.jl:158-211, velocity_tracer_gradients.jl:25-42
  src/Models/NonhydrostaticModels/update_hydrostatic_pressure.jl:12-20,41-49
"""
import numpy as np

from .grid import Bounded, Flat, cbrt_correctly_rounded
from .advection import (is_upwind, symmetric_interpolate, biased_interpolate, shift)
from .operators import area, volume, delta, diff, interp, deriv

VEL_LOC = (("f", "c", "c"), ("c", "f", "c"), ("c", "c", "f"))
CCC = ("c", "c", "c")


class Cached:
    """Evaluate ``fn(R)`` once over region R and serve sub-regions as views (each face flux is evaluated once;
    the reference evaluates it at both adjacent cells with identical results)."""

    def __init__(self, fn, R):
        self.R = [tuple(r) for r in R]
        v = fn(self.R)
        shape = tuple(r[1] - r[0] + 1 for r in self.R)
        self.v = np.broadcast_to(v, shape) if np.ndim(v) else v

    def __call__(self, ri, rj, rk):
        if not np.ndim(self.v):
            return self.v
        sl = tuple(slice(r[0] - R0[0], r[1] - R0[0] + 1) for r, R0 in zip((ri, rj, rk), self.R))
        return self.v[sl]


def flux_location(a, b):
    if a == b:
        return CCC
    loc = ["c", "c", "c"]
    loc[a] = loc[b] = "f"
    return tuple(loc)


def tendency_region(grid, loc):
    """`:xyz` with exclude_periphery=true: Face-located fields skip index 1 in Bounded dims (N > 1)."""
    R = []
    for d in range(3):
        off = 1 if (loc[d] == "f" and grid.topology[d] == Bounded and grid.N[d] > 1) else 0
        R.append((1 + off, grid.N[d]))
    return R


def _extended(R, d, target):
    """Region of the flux needed to difference onto `target` ('f': F[i]-F[i-1]; 'c': F[i+1]-F[i])."""
    R = [tuple(r) for r in R]
    R[d] = (R[d][0] - 1, R[d][1]) if target == "f" else (R[d][0], R[d][1] + 1)
    return R


# ---------------------------------------------------------------------------------------------
# advection
# ---------------------------------------------------------------------------------------------

def advective_momentum_flux(grid, adv, U, a, b, R):
    """advective_momentum_flux_{U_b}{u_a} over the flux-location region R."""
    scheme = adv[b]
    if grid.topology[b] == Flat or scheme is None:
        return grid.FT(0)
    if a == b:
        sym_dim, stag = b, "c"
    else:
        sym_dim, stag = a, "f"
    if is_upwind(scheme):
        def Aq(*r):
            return area(grid, b, VEL_LOC[b], list(r)) * U[b].at(*r)
        ut = symmetric_interpolate(grid, scheme, Aq, sym_dim, stag, R)
        qR = biased_interpolate(grid, scheme, U[a].at, b, stag, R, ut > 0)
        return ut * qR
    A = area(grid, b, flux_location(a, b), R)
    return (A * symmetric_interpolate(grid, scheme, U[b].at, sym_dim, stag, R)
            * symmetric_interpolate(grid, scheme, U[a].at, b, stag, R))


def div_Uu(grid, adv, U, a, R):
    """div_𝐯u / div_𝐯v / div_𝐯w for component a over region R (momentum_advection_operators.jl:46-83)."""
    if all(s is None for s in adv):
        return grid.FT(0)
    total = None
    for b in range(3):
        target = "f" if a == b else "c"
        if grid.topology[b] == Flat:
            t = grid.FT(0)
        else:
            F = Cached(lambda RR, b=b: advective_momentum_flux(grid, adv, U, a, b, RR), _extended(R, b, target))
            t = diff(grid, F, b, target, R)
        total = t if total is None else total + t
    return grid.FT(1) / volume(grid, VEL_LOC[a], R) * total


def advective_tracer_flux(grid, adv, U, c, b, R):
    scheme = adv[b]
    if grid.topology[b] == Flat or scheme is None:
        return grid.FT(0)
    A = area(grid, b, VEL_LOC[b], R)
    ub = U[b].at(*R)
    if is_upwind(scheme):
        cR = biased_interpolate(grid, scheme, c.at, b, "f", R, ub > 0)
        return A * ub * cR
    return (A * ub) * symmetric_interpolate(grid, scheme, c.at, b, "f", R)


def div_Uc(grid, adv, U, c, R):
    if all(s is None for s in adv):
        return grid.FT(0)
    total = None
    for b in range(3):
        if grid.topology[b] == Flat:
            t = grid.FT(0)
        else:
            F = Cached(lambda RR, b=b: advective_tracer_flux(grid, adv, U, c, b, RR), _extended(R, b, "c"))
            t = diff(grid, F, b, "c", R)
        total = t if total is None else total + t
    return grid.FT(1) / volume(grid, CCC, R) * total


# ---------------------------------------------------------------------------------------------
# Coriolis (FPlane)
# ---------------------------------------------------------------------------------------------

def _inactive_cell(grid, R):
    """inactive_cell(i,j,k) over region R (Grids/inactive_node.jl:43-108): outside 1:N in Bounded dims."""
    out = np.zeros((1, 1, 1), dtype=bool)
    for d in range(3):
        if grid.topology[d] == Bounded:
            idx = np.arange(R[d][0], R[d][1] + 1)
            shape = [1, 1, 1]
            shape[d] = len(idx)
            out = out | ((idx < 1) | (idx > grid.N[d])).reshape(shape)
    return np.broadcast_to(out, tuple(r[1] - r[0] + 1 for r in R))


def _not_peripheral(grid, loc):
    def get(*r):
        r = list(r)
        bad = _inactive_cell(grid, r)
        for d in range(3):
            if loc[d] == "f":
                bad = bad | _inactive_cell(grid, shift(r, d, -1))
        return (~bad).astype(np.int64)
    return get


def _interp_xy(grid, get, tx, ty, R):
    """ℑxy = ℑy(ℑx ·) onto (tx, ty)."""
    def ix(*r):
        return interp(grid, get, 0, tx, list(r))
    return interp(grid, ix, 1, ty, R)


def f_cross_U(grid, coriolis, U, a, R):
    """x_f_cross_U / y_f_cross_U / z_f_cross_U (f_plane.jl:44-46)."""
    FT = grid.FT
    if coriolis is None or a == 2:
        return FT(0)
    f = FT(coriolis.f)
    if a == 0:      # at (f,c,c): -f · active-weighted ℑxyᶠᶜ(v)
        q, nodeloc, tx, ty, sign = U[1], VEL_LOC[1], "f", "c", -1
    else:           # at (c,f,c): +f · active-weighted ℑxyᶜᶠ(u)
        q, nodeloc, tx, ty, sign = U[0], VEL_LOC[0], "c", "f", +1
    active = _interp_xy(grid, _not_peripheral(grid, nodeloc), tx, ty, R)
    active = np.asarray(active, dtype=FT) if np.ndim(active) else FT(active)
    val = _interp_xy(grid, q.at, tx, ty, R)
    with np.errstate(divide="ignore", invalid="ignore"):
        aw = np.where(active == 0, FT(0), val / active)
    return (-f if sign < 0 else f) * aw


# ---------------------------------------------------------------------------------------------
# buoyancy
# ---------------------------------------------------------------------------------------------

def buoyancy_perturbation(grid, buoyancy, tracers):
    """getter for buoyancy_perturbationᶜᶜᶜ."""
    FT = grid.FT
    if buoyancy.kind == "tracer":
        return tracers["b"].at
    g, al, be = FT(buoyancy.g), FT(buoyancy.alpha), FT(buoyancy.beta)

    def get(*r):
        return g * (al * tracers["T"].at(*r) - be * tracers["S"].at(*r))
    return get


def update_hydrostatic_pressure(grid, pHY, buoyancy, tracers):
    """_update_hydrostatic_pressure! over i ∈ 0:Nx+1, j ∈ 0:Ny+1 (1:N in Flat dims)."""
    if pHY is None or buoyancy is None or grid.topology[2] == Flat:
        return
    ri = (1, grid.N[0]) if grid.topology[0] == Flat else (0, grid.N[0] + 1)
    rj = (1, grid.N[1]) if grid.topology[1] == Flat else (0, grid.N[1] + 1)
    b = buoyancy_perturbation(grid, buoyancy, tracers)
    Nz = grid.N[2]

    def zdgb(k):     # ĝ_z · ℑzᵃᵃᶠ(b) with ĝ_z = 1
        R = [ri, rj, (k, k)]
        return interp(grid, b, 2, "f", R)

    def dzf(k):
        return delta(grid, 2, "f", [ri, rj, (k, k)])
    pHY.at(ri, rj, (Nz, Nz))[...] = -zdgb(Nz + 1) * dzf(Nz + 1)
    for k in range(Nz - 1, 0, -1):
        pHY.at(ri, rj, (k, k))[...] = pHY.at(ri, rj, (k + 1, k + 1)) - zdgb(k + 1) * dzf(k + 1)


# ---------------------------------------------------------------------------------------------
# ScalarDiffusivity (explicit, isotropic)
# ---------------------------------------------------------------------------------------------

def _strain(grid, U, a, b, R):
    """Σ_ab at its natural location over R (velocity_tracer_gradients.jl:25-42)."""
    if a == b:
        return deriv(grid, U[a].at, a, CCC, R)
    lo, hi = min(a, b), max(a, b)
    loc = flux_location(lo, hi)
    return grid.FT(0.5) * (deriv(grid, U[lo].at, hi, loc, R) + deriv(grid, U[hi].at, lo, loc, R))


def strain_dot_strain_ccc(grid, U, R):
    """ΣᵢⱼΣᵢⱼᶜᶜᶜ (Smagorinskys/scale_invariant_operators.jl:10-13; velocity_tracer_gradients.jl:44-46,78):
    tr_Σ² + 2·ℑxyᶜᶜᵃ(Σ₁₂²) + 2·ℑxzᶜᵃᶜ(Σ₁₃²) + 2·ℑyzᵃᶜᶜ(Σ₂₃²), ℑ_ab = ℑ_b(ℑ_a ·) (interpolation_operators.jl:45-53)."""
    FT = grid.FT
    s = [_strain(grid, U, a, a, R) for a in range(3)]
    total = s[0] * s[0] + s[1] * s[1] + s[2] * s[2]
    for lo, hi in ((0, 1), (0, 2), (1, 2)):
        def sq(*r, lo=lo, hi=hi):
            x = _strain(grid, U, lo, hi, list(r))
            return x * x

        def inner(*r, lo=lo, sq=sq):
            return interp(grid, sq, lo, "c", list(r))
        total = total + FT(2) * interp(grid, inner, hi, "c", R)
    return total


def dz_b(grid, buoyancy, tracers):
    """getter for ∂z_b at (c,c,f) — buoyancy_tracer.jl:16, seawater_buoyancy.jl:219-224, no_buoyancy.jl:9."""
    FT = grid.FT
    ccf = ("c", "c", "f")
    if buoyancy is None:
        return lambda *r: FT(0)
    if buoyancy.kind == "tracer":
        return lambda *r: deriv(grid, tracers["b"].at, 2, ccf, list(r))
    g, al, be = FT(buoyancy.g), FT(buoyancy.alpha), FT(buoyancy.beta)
    return lambda *r: g * (al * deriv(grid, tracers["T"].at, 2, ccf, list(r)) - be * deriv(grid, tracers["S"].at, 2, ccf, list(r)))


def compute_smagorinsky_viscosity(model):
    """_compute_smagorinsky_viscosity! over :xyz (Smagorinskys/smagorinsky.jl:88-103) with the constant coefficient
    (:105) or the Lilly stability correction (lilly_coefficient.jl:125-139): νₑ = cˢ²·Δᶠ²·√(2Σ²), Δᶠ = cbrt(ΔxΔyΔz)."""
    grid, FT, cl = model.grid, model.grid.FT, model.closure
    R = [(1, grid.N[0]), (1, grid.N[1]), (1, grid.N[2])]
    S2 = np.asarray(strain_dot_strain_ccc(grid, model.velocities, R), dtype=FT)
    D3 = delta(grid, 0, "c", R) * delta(grid, 1, "c", R) * delta(grid, 2, "c", R)
    D3 = np.asarray(D3, dtype=FT)
    Df = np.vectorize(lambda v: cbrt_correctly_rounded(v, FT), otypes=[FT])(D3)      # a few distinct values (one per k)
    if cl.kind == "smagorinsky_lilly":
        N2 = interp(grid, dz_b(grid, model.buoyancy, model.tracers), 2, "c", R)
        N2p = np.maximum(FT(0), N2)
        with np.errstate(divide="ignore", invalid="ignore"):
            ratio = FT(cl.Cb) * N2p / S2
            # Julia's min(1, NaN) is NaN; the ifelse below discards it
            sig2 = FT(1) - np.where(np.isnan(ratio), ratio, np.minimum(FT(1), ratio))
            sig = np.where(S2 == 0, FT(0), np.sqrt(sig2))
        cs2 = sig * (FT(cl.C) * FT(cl.C))
    else:
        cs2 = np.float64(cl.C) * np.float64(cl.C)      # the constant coefficient is NOT converted to eltype(grid)
    nu = cs2 * (Df * Df) * np.sqrt(FT(2) * S2)
    model.nu_e.at(*R)[...] = np.broadcast_to(nu, model.nu_e.at(*R).shape).astype(FT)


def _nu_getter(grid, nu, a, b):
    """ν at the location of Σ_ab: a number, or the field νₑ at ccc (a == b) / ℑ_hi(ℑ_lo νₑ) at the edge
    (abstract_scalar_diffusivity_closure.jl:60-75 with a Center-located viscosity field)."""
    if not hasattr(nu, "at"):
        c = grid.FT(nu)
        return lambda RR: c
    if a == b:
        return lambda RR: nu.at(*RR)
    lo, hi = min(a, b), max(a, b)

    def inner(*r):
        return interp(grid, nu.at, lo, "f", list(r))
    return lambda RR: interp(grid, inner, hi, "f", RR)


def _wall_mask_z(grid, RR):
    """(k == 1) | (k == Nz+1) over the flux region RR, broadcastable."""
    k = np.arange(RR[2][0], RR[2][1] + 1)
    return ((k == 1) | (k == grid.N[2] + 1)).reshape(1, 1, -1)


def div_viscous_flux(grid, nu, U, a, R, vitd=False):
    """∂ⱼ_τ_aⱼ (closure_kernel_operators.jl:22-41) with viscous_flux = -2·(ν·Σ); ν a number or the eddy-viscosity field.
    vitd: VerticallyImplicitTimeDiscretization on a vertically Bounded grid (abstract_scalar_diffusivity_closure.jl:219-252):
    the z-fluxes keep only what is not solved implicitly — -(ν·∂x w), -(ν·∂y w), 0 for u, v, w — except on the two walls
    k = 1, Nz+1, where they stay explicit."""
    FT = grid.FT
    total = None
    for b in range(3):
        target = "f" if a == b else "c"
        floc = flux_location(a, b)
        nu_at = _nu_getter(grid, nu, a, b)

        def Aq(RR, b=b, floc=floc, nu_at=nu_at):
            explicit = FT(-2) * (nu_at(RR) * _strain(grid, U, a, b, RR))
            if vitd and b == 2:
                if a == 2:
                    rest = FT(0)
                else:
                    rest = -(nu_at(RR) * deriv(grid, U[2].at, a, floc, RR))
                explicit = np.where(_wall_mask_z(grid, RR), explicit, rest)
            return area(grid, b, floc, RR) * explicit
        if grid.topology[b] == Flat:
            t = FT(0)
        else:
            t = diff(grid, Cached(Aq, _extended(R, b, target)), b, target, R)
        total = t if total is None else total + t
    return FT(1) / volume(grid, VEL_LOC[a], R) * total


def div_diffusive_flux(grid, kappa, c, R, Pr=None, vitd=False):
    """∇_dot_qᶜ (closure_kernel_operators.jl:43-48) with diffusive_flux = -κ·∂c; κ a number, or (Pr given) the field νₑ:
    κ at the face = ℑ_b(νₑ) / Pr (Smagorinskys/smagorinsky.jl:135-137)."""
    FT = grid.FT
    total = None
    for b in range(3):
        floc = VEL_LOC[b]
        if Pr is None:
            kc = FT(kappa)
            k_at = lambda RR, kc=kc: kc
        else:
            k_at = lambda RR, b=b: interp(grid, kappa.at, b, "f", RR) / FT(Pr)

        def Aq(RR, b=b, floc=floc, k_at=k_at):
            flux = -k_at(RR) * deriv(grid, c.at, b, floc, RR)
            if vitd and b == 2:          # diffusive_flux_z: explicit on the walls only
                flux = np.where(_wall_mask_z(grid, RR), flux, FT(0))
            return area(grid, b, floc, RR) * flux
        if grid.topology[b] == Flat:
            t = FT(0)
        else:
            t = diff(grid, Cached(Aq, _extended(R, b, "c")), b, "c", R)
        total = t if total is None else total + t
    return FT(1) / volume(grid, CCC, R) * total


# ---------------------------------------------------------------------------------------------
# tendencies
# ---------------------------------------------------------------------------------------------

def momentum_tendency(model, a):
    """u/v/w_velocity_tendency over the launch region; returns (R, values)."""
    grid, FT = model.grid, model.grid.FT
    U = model.velocities
    R = tendency_region(grid, VEL_LOC[a])
    if any(r[1] < r[0] for r in R):
        return R, None
    G = -div_Uu(grid, model.advection, U, a, R)
    # buoyancy: x,y components are zero for the default gravity vector; z only when pHY′ is nothing
    if model.buoyancy is not None and a == 2 and model.pHY is None:
        b = buoyancy_perturbation(grid, model.buoyancy, model.tracers)
        G = G + interp(grid, b, 2, "f", R)
    else:
        G = G + FT(0)
    G = G - f_cross_U(grid, model.coriolis, U, a, R)
    if a < 2:
        if model.pHY is not None:
            G = G - deriv(grid, model.pHY.at, a, VEL_LOC[a], R)
        else:
            G = G - FT(0)
    if model.closure is not None:
        nu = model.nu_e if model.closure.kind != "scalar" else model.closure.nu
        G = G - div_viscous_flux(grid, nu, U, a, R, vitd=getattr(model.closure, "vitd", False))
    else:
        G = G - FT(0)
    return R, G


def tracer_tendency(model, name):
    grid, FT = model.grid, model.grid.FT
    c = model.tracers[name]
    R = tendency_region(grid, CCC)
    G = -div_Uc(grid, model.advection, model.velocities, c, R)
    if model.closure is not None and model.closure.kind != "scalar":
        kfield = model.kappa_e[name] if model.closure.kind == "amd" else model.nu_e
        G = G - div_diffusive_flux(grid, kfield, c, R, Pr=model.closure.Pr_for(name))
    elif model.closure is not None:
        G = G - div_diffusive_flux(grid, model.closure.kappa_for(name), c, R, vitd=getattr(model.closure, "vitd", False))
    else:
        G = G - FT(0)
    return R, G
