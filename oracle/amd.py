"""AnisotropicMinimumDissipation eddy viscosity and diffusivities (oracle; test infrastructure only).

Follows the reference's
  src/TurbulenceClosures/turbulence_closure_implementations/anisotropic_minimum_dissipation.jl:143-208   (kernels)
  …/anisotropic_minimum_dissipation.jl:216-227                                                           (filter widths = 2Δ)
  …/anisotropic_minimum_dissipation.jl:233-281, 287-313, 333-365                                         ("the 30 terms")
  src/TurbulenceClosures/velocity_tracer_gradients.jl:8-23,129-262                                       (normalised gradients)
  src/Operators/interpolation_operators.jl:45-53                                                         (ℑ_ab = ℑ_b(ℑ_a ·))
Every sum and product keeps the reference's association (left to right inside each parenthesised group).  The
`ℑxzᶜᵃᶜ(norm_∂y_w)` of cy_uy (anisotropic_minimum_dissipation.jl:347) is reproduced as written.
Cb (buoyancy modification) is `nothing`, the reference's default (it warns that anything else is unvalidated).
"""
import numpy as np

from .operators import delta, interp, deriv

CCC = ("c", "c", "c")
FFC = ("f", "f", "c")
FCF = ("f", "c", "f")
CFF = ("c", "f", "f")
FCC = ("f", "c", "c")
CFC = ("c", "f", "c")
CCF = ("c", "c", "f")


def _filter_width(grid, d):
    """getter of Δᶠ_d = 2·Δ_dᶜᶜᶜ evaluated at the index it is asked for (Δᶠx_ffc etc. are all aliases of the ccc one)."""
    FT = grid.FT

    def get(*r):
        return FT(2) * delta(grid, d, "c", list(r))
    return get


def _to_center(grid, f, dims):
    """ℑ onto Centers along `dims` in the given order (first = innermost): ℑxyᶜᶜᵃ f = _to_center(f, (0, 1))."""
    g = f
    for d in dims:
        def g(*r, d=d, inner=g):
            return interp(grid, inner, d, "c", list(r))
    return g


class _Terms:
    def __init__(self, grid, U):
        self.grid, self.FT = grid, grid.FT
        self.u, self.v, self.w = U
        self.D = [_filter_width(grid, d) for d in range(3)]

    # --- normalised velocity gradients (velocity_tracer_gradients.jl:129-158)
    def dx_u(self, *r):
        return deriv(self.grid, self.u.at, 0, CCC, list(r))

    def dy_v(self, *r):
        return deriv(self.grid, self.v.at, 1, CCC, list(r))

    def dz_w(self, *r):
        return deriv(self.grid, self.w.at, 2, CCC, list(r))

    def _norm(self, num, den, field, d, loc):
        def get(*r):
            return self.D[num](*r) / self.D[den](*r) * deriv(self.grid, field.at, d, loc, list(r))
        return get

    def setup(self):
        self.dx_v = self._norm(0, 1, self.v, 0, FFC)
        self.dy_u = self._norm(1, 0, self.u, 1, FFC)
        self.dx_w = self._norm(0, 2, self.w, 0, FCF)
        self.dz_u = self._norm(2, 0, self.u, 2, FCF)
        self.dy_w = self._norm(1, 2, self.w, 1, CFF)
        self.dz_v = self._norm(2, 1, self.v, 2, CFF)
        half = self.FT(0.5)
        self.S12 = lambda *r: half * (self.dy_u(*r) + self.dx_v(*r))
        self.S13 = lambda *r: half * (self.dz_u(*r) + self.dx_w(*r))
        self.S23 = lambda *r: half * (self.dz_v(*r) + self.dy_w(*r))
        return self


def _sq(f):
    def get(*r):
        x = f(*r)
        return x * x
    return get


def _prod(f, g):
    return lambda *r: f(*r) * g(*r)


def delta_squared(grid, R):
    """δ² = 3 / (1/Δᶠx² + 1/Δᶠy² + 1/Δᶠz²)."""
    FT = grid.FT
    Dx, Dy, Dz = (_filter_width(grid, d)(*R) for d in range(3))
    return FT(3) / (FT(1) / (Dx * Dx) + FT(1) / (Dy * Dy) + FT(1) / (Dz * Dz))


def amd_viscosity(grid, U, C_nu, R):
    """_compute_AMD_viscosity! over region R."""
    FT = grid.FT
    t = _Terms(grid, U).setup()
    xy, xz, yz = (0, 1), (0, 2), (1, 2)
    I = lambda f, dims: _to_center(grid, f, dims)(*R)
    ux, vy, wz = t.dx_u(*R), t.dy_v(*R), t.dz_w(*R)
    # norm_tr_∇uᶜᶜᶜ
    q = (ux * ux + vy * vy + wz * wz
         + I(_sq(t.dx_v), xy) + I(_sq(t.dy_u), xy)
         + I(_sq(t.dx_w), xz) + I(_sq(t.dz_u), xz)
         + I(_sq(t.dy_w), yz) + I(_sq(t.dz_v), yz))
    two = FT(2)
    g1 = (ux * (ux * ux)
          + vy * I(_sq(t.dx_v), xy)
          + wz * I(_sq(t.dx_w), xz)
          + two * ux * I(_prod(t.dx_v, t.S12), xy)
          + two * ux * I(_prod(t.dx_w, t.S13), xz)
          + two * I(t.dx_v, xy) * I(t.dx_w, xz) * I(t.S23, yz))
    g2 = (ux * I(_sq(t.dy_u), xy)
          + vy * (vy * vy)
          + wz * I(_sq(t.dy_w), yz)
          + two * vy * I(_prod(t.dy_u, t.S12), xy)
          + two * I(t.dy_u, xy) * I(t.dy_w, yz) * I(t.S13, xz)
          + two * vy * I(_prod(t.dy_w, t.S23), yz))
    g3 = (ux * I(_sq(t.dz_u), xz)
          + vy * I(_sq(t.dz_v), yz)
          + wz * (wz * wz)
          + two * I(t.dz_u, xz) * I(t.dz_v, yz) * I(t.S12, xy)
          + two * wz * I(_prod(t.dz_u, t.S13), xz)
          + two * wz * I(_prod(t.dz_v, t.S23), yz))
    r = g1 + g2 + g3
    d2 = delta_squared(grid, R)
    with np.errstate(divide="ignore", invalid="ignore"):
        nu = -FT(C_nu) * d2 * (r - FT(0)) / q
    nu = np.where(q == 0, FT(0), nu)
    return np.maximum(FT(0), nu).astype(FT)


def amd_diffusivity(grid, U, c, C_kappa, R):
    """_compute_AMD_diffusivity! for one tracer over region R."""
    FT = grid.FT
    t = _Terms(grid, U).setup()
    Dx, Dy, Dz = t.D
    cx = lambda *r: Dx(*r) * deriv(grid, c.at, 0, FCC, list(r))
    cy = lambda *r: Dy(*r) * deriv(grid, c.at, 1, CFC, list(r))
    cz = lambda *r: Dz(*r) * deriv(grid, c.at, 2, CCF, list(r))
    I = lambda f, dims: _to_center(grid, f, dims)(*R)
    xy, xz, yz = (0, 1), (0, 2), (1, 2)
    sigma = I(_sq(cx), (0,)) + I(_sq(cy), (1,)) + I(_sq(cz), (2,))
    ux, vy, wz = t.dx_u(*R), t.dy_v(*R), t.dz_w(*R)
    Icx, Icy, Icz = I(cx, (0,)), I(cy, (1,)), I(cz, (2,))
    cx_ux = (ux * I(_sq(cx), (0,))
             + I(t.dx_v, xy) * Icx * Icy
             + I(t.dx_w, xz) * Icx * Icz)
    cy_uy = (I(t.dy_u, xy) * Icy * Icx
             + vy * I(_sq(cy), (1,))
             + I(t.dy_w, xz) * Icy * Icz)          # ℑxzᶜᵃᶜ of a cff quantity, as the reference writes it
    cz_uz = (I(t.dz_u, xz) * Icz * Icx
             + I(t.dz_v, yz) * Icz * Icy
             + wz * I(_sq(cz), (2,)))
    theta = cx_ux + cy_uy + cz_uz
    d2 = delta_squared(grid, R)
    with np.errstate(divide="ignore", invalid="ignore"):
        kappa = -FT(C_kappa) * d2 * theta / sigma
    kappa = np.where(sigma == 0, FT(0), kappa)
    return np.maximum(FT(0), kappa).astype(FT)
