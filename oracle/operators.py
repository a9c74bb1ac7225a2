"""Grid metrics and stencil operators on index regions (oracle; test infrastructure only).

Follows the reference's
  src/Operators/spacings_and_areas_and_volumes.jl:106-140,263-300,336-345  (Δ, A = Δ·Δ, V = Az·Δz)
  src/Operators/difference_operators.jl:7-49       (δ; zero in Flat dimensions)
  src/Operators/derivative_operators.jl:20-30      (∂ = δ / Δ)
  src/Operators/interpolation_operators.jl:8-28    (ℑ = FT(0.5)·(a + b); identity in Flat dimensions)
  src/Operators/divergence_operators.jl:16-19      (divᶜᶜᶜ = 1/V · (δx(Ax u) + δy(Ay v) + δz(Az w)))

A region R is a list of three inclusive 1-based index ranges [(i0,i1),(j0,j1),(k0,k1)].
"""
import numpy as np

from .grid import Flat
from .advection import shift


def delta(grid, d, loc, R):
    """Δ along dim d at location loc ('c'/'f') over region R, broadcastable against a region array."""
    c = grid.coords[d]
    if c.topo == Flat or c.regular:
        return c.dc if loc == "c" else c.df       # equal scalars when regular
    v = c.dc.r(*R[d]) if loc == "c" else c.df.r(*R[d])
    shape = [1, 1, 1]
    shape[d] = len(v)
    return v.reshape(shape)


def area(grid, d, loc, R):
    """A_d at location loc=(ℓx,ℓy,ℓz): product of the two transverse spacings, in the reference's order
    (Ax = Δy·Δz, Ay = Δx·Δz, Az = Δx·Δy)."""
    a, b = [e for e in range(3) if e != d]
    return delta(grid, a, loc[a], R) * delta(grid, b, loc[b], R)


def volume(grid, loc, R):
    """V = Az · Δz."""
    return area(grid, 2, loc, R) * delta(grid, 2, loc[2], R)


def face_loc(loc, d):
    """Location after moving from loc to the opposite stagger in dim d."""
    l = list(loc)
    l[d] = "f" if loc[d] == "c" else "c"
    return tuple(l)


def diff(grid, get, d, target_loc, R):
    """δ_d onto target_loc[d]: centre target → f[i+1]-f[i]; face target → f[i]-f[i-1]. Zero if Flat."""
    if grid.topology[d] == Flat:
        return grid.FT(0)
    if target_loc == "c":
        return get(*shift(R, d, 1)) - get(*R)
    return get(*R) - get(*shift(R, d, -1))


def interp(grid, get, d, target_loc, R):
    """ℑ_d onto target_loc: FT(0.5)·(a + b); identity if Flat."""
    if grid.topology[d] == Flat:
        return get(*R)
    half = grid.FT(0.5)
    if target_loc == "c":
        return half * (get(*R) + get(*shift(R, d, 1)))
    return half * (get(*shift(R, d, -1)) + get(*R))


def deriv(grid, field_get, d, target_loc3, R):
    """∂_d at full location target_loc3 (a 3-tuple): δ / Δ (Δ at the target location)."""
    if grid.topology[d] == Flat:
        return grid.FT(0)
    return diff(grid, field_get, d, target_loc3[d], R) / delta(grid, d, target_loc3[d], R)


def interior_region(grid):
    return [(1, grid.N[0]), (1, grid.N[1]), (1, grid.N[2])]


def divergence_ccc(grid, u, v, w, R):
    """divᶜᶜᶜ — divergence_operators.jl:16-19."""
    U = (u, v, w)
    locs = (("f", "c", "c"), ("c", "f", "c"), ("c", "c", "f"))
    total = None
    for d in range(3):
        def Aq(*r, d=d):
            return area(grid, d, locs[d], list(r)) * U[d].at(*r)
        t = diff(grid, Aq, d, "c", R)
        total = t if total is None else total + t
    return grid.FT(1) / volume(grid, ("c", "c", "c"), R) * total


def laplacian_ccc(grid, phi, R):
    """∇²ᶜᶜᶜ — src/Operators/laplacian_operators.jl:36-40: 1/V · (δx(Ax ∂x c) + δy(Ay ∂y c) + δz(Az ∂z c));
    used by the reference's Poisson tests (test/utils_for_runtests.jl:125-128)."""
    total = None
    ccc = ("c", "c", "c")
    for d in range(3):
        floc = face_loc(ccc, d)

        def A_dphi(*r, d=d, floc=floc):
            r = list(r)
            return area(grid, d, floc, r) * deriv(grid, phi.at, d, floc, r)
        t = diff(grid, A_dphi, d, "c", R)
        total = t if total is None else total + t
    return grid.FT(1) / volume(grid, ccc, R) * total
