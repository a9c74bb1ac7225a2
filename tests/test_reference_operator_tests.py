"""The reference's operator tests (test/test_operators.jl:127-222 and the un-immersed part of :8-125) restated against the
oracle's operators with the reference's exact `==`: lengths, areas, volume of a (π, 2π, 3π) cell; differences, derivatives and
two-point interpolations of a function of the field at the centre of a 3³ grid."""
import itertools

import numpy as np
import pytest

import oracle as o
from oracle.operators import area, delta, deriv, diff, interp, volume

LOCS = list(itertools.product("cf", repeat=3))


def test_lengths_areas_and_volumes_of_the_reference_cell():
    FT = np.float64
    g = o.RectilinearGrid(FT, size=(1, 1, 1), extent=(np.pi, 2 * np.pi, 3 * np.pi))
    R = [(1, 1), (1, 1), (1, 1)]
    for loc in "cf":
        assert delta(g, 0, loc, R) == FT(np.pi) and delta(g, 1, loc, R) == FT(2 * np.pi) and delta(g, 2, loc, R) == FT(3 * np.pi)
    for loc in LOCS:
        assert area(g, 0, loc, R) == FT(6 * np.pi ** 2)
        assert area(g, 1, loc, R) == FT(3 * np.pi ** 2)
        assert area(g, 2, loc, R) == FT(2 * np.pi ** 2)
        assert volume(g, loc, R) == FT(6 * np.pi ** 3)


@pytest.mark.parametrize("FT", [np.float32, np.float64])
def test_differences_derivatives_and_interpolations_of_a_function(FT):
    rng = np.random.default_rng(3)
    g = o.RectilinearGrid(FT, size=(3, 3, 3), extent=(3, 3, 3), topology=("Periodic", "Periodic", "Bounded"))
    phi = rng.random((3, 3, 3)).astype(FT)
    p2 = phi ** 2

    def f(ri, rj, rk):                                       # f(i, j, k, grid, ϕ) = ϕ[i, j, k]^2 over a region of the interior
        return p2[ri[0] - 1: ri[1], rj[0] - 1: rj[1], rk[0] - 1: rk[1]]

    R = [(2, 2), (2, 2), (2, 2)]
    val = lambda a: np.asarray(a).reshape(-1)[0]
    # (0-based [1,1,1] is the reference's [2,2,2])
    d_c = [p2[2, 1, 1] - p2[1, 1, 1], p2[1, 2, 1] - p2[1, 1, 1], p2[1, 1, 2] - p2[1, 1, 1]]
    d_f = [p2[1, 1, 1] - p2[0, 1, 1], p2[1, 1, 1] - p2[1, 0, 1], p2[1, 1, 1] - p2[1, 1, 0]]
    i_c = [(p2[2, 1, 1] + p2[1, 1, 1]) / 2, (p2[1, 2, 1] + p2[1, 1, 1]) / 2, (p2[1, 1, 2] + p2[1, 1, 1]) / 2]
    i_f = [(p2[1, 1, 1] + p2[0, 1, 1]) / 2, (p2[1, 1, 1] + p2[1, 0, 1]) / 2, (p2[1, 1, 1] + p2[1, 1, 0]) / 2]
    for d in range(3):
        assert val(diff(g, f, d, "c", R)) == d_c[d] and val(diff(g, f, d, "f", R)) == d_f[d]
        assert val(interp(g, f, d, "c", R)) == i_c[d] and val(interp(g, f, d, "f", R)) == i_f[d]
        for loc in LOCS:                                     # Δ = 1 here: ∂ = δ / Δ at every location of the target stagger
            want = d_c[d] if loc[d] == "c" else d_f[d]
            assert val(deriv(g, f, d, loc, R)) == want / FT(1)
