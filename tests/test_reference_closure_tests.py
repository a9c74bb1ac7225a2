"""The reference's ScalarDiffusivity known-answer test (test/test_turbulence_closures.jl:36-66,340-348) restated over the
oracle's closure operators with the reference's exact equalities."""
import numpy as np
import pytest

import oracle as o
from oracle.fields import Field, fill_halo_regions
from oracle.tendencies import VEL_LOC, CCC, div_viscous_flux, div_diffusive_flux


@pytest.mark.parametrize("FT", [np.float32, np.float64])
def test_constant_isotropic_diffusivity_fluxdiv(FT):
    nu, kappa = FT(0.3), FT(0.7)
    closure = o.ScalarDiffusivity(nu=nu, kappa={"T": kappa, "S": kappa})
    assert closure.nu == FT(0.3) and closure.kappa_for("T") == FT(0.7)
    grid = o.RectilinearGrid(FT, size=(3, 1, 4), extent=(3, 1, 4))
    U = [Field(grid, VEL_LOC[a], name="uvw"[a]) for a in range(3)]
    T = Field(grid, CCC, name="T")
    for f, mid in zip(U + [T], (-0.5, -2, -3, -1)):
        f.interior[:, 0, :4] = np.array([0, mid, 0], dtype=FT)[:, None]
    fill_halo_regions(U + [T])
    R = [(2, 2), (1, 1), (3, 3)]
    assert div_diffusive_flux(grid, kappa, T, R)[0, 0, 0] == -2 * kappa          # ∇_dot_qᶜ(2, 1, 3, …) == −2κ
    assert div_viscous_flux(grid, nu, U, 0, R)[0, 0, 0] == -2 * nu               # ∂ⱼ_τ₁ⱼ == −2ν
    assert div_viscous_flux(grid, nu, U, 1, R)[0, 0, 0] == -4 * nu               # ∂ⱼ_τ₂ⱼ == −4ν
    assert div_viscous_flux(grid, nu, U, 2, R)[0, 0, 0] == -6 * nu               # ∂ⱼ_τ₃ⱼ == −6ν


@pytest.mark.gpu
@pytest.mark.parametrize("FT", [np.float32, np.float64])
def test_library_constant_isotropic_diffusivity_tendencies(FT):
    """The same known answers through the C ABI: with no advection, buoyancy or rotation the tendencies of the library's model are
    minus the closure's flux divergences, so Gu, Gv, Gw, GT at (2, 1, 3) are 2ν, 4ν, 6ν, 2κ — exactly."""
    import oceananigans_jl_b200 as oc
    nu, kappa = FT(0.3), FT(0.7)
    grid = oc.RectilinearGrid(FT, size=(3, 1, 4), extent=(3, 1, 4))
    m = oc.NonhydrostaticModel(grid, advection=None, closure=oc.ScalarDiffusivity(nu=float(nu), kappa={"T": float(kappa), "S": float(kappa)}),
                               tracers=("T", "S"))
    col = lambda mid, nz: np.broadcast_to(np.array([0, mid, 0], dtype=FT)[:, None, None], (3, 1, nz)).copy()
    w = col(-3, 5)
    w[:, :, 4] = 0                                              # the reference fills k = 1 … 4 only; k = 5 is the top wall
    oc.set_model(m, enforce_incompressibility=False, u=col(-0.5, 4), v=col(-2, 4), w=w, T=col(-1, 4))
    oc.update_state(m, compute_tendencies=True)
    G = m.timestepper.Gn
    assert G["T"].interior()[1, 0, 2] == 2 * kappa
    assert G["u"].interior()[1, 0, 2] == 2 * nu
    assert G["v"].interior()[1, 0, 2] == 4 * nu
    assert G["w"].interior()[1, 0, 2] == 6 * nu
