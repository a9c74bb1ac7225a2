"""WENO(grid): the per-index reconstruction coefficients of a stretched direction, generated on the host.

The reference computes them on the CPU when the scheme is constructed (src/Advection/reconstruction_coefficients.jl:237-324:
``compute_reconstruction_coefficients`` → ``create_reconstruction_coefficients`` → ``stencil_coefficients``) and the kernels only
look them up (``retrieve_coeff``, weno_reconstruction.jl:161-167); here they are handed to the library with
``nh_model_set_weno_z_coefficients``.  The arithmetic follows ``stencil_coefficients`` (:100-115): products and sums of coordinate
differences in the grid's float type, the j-th coefficient accumulated without rounding (BigFloat there, exact rationals here),
one rounding to the float type, the last coefficient = 1 − Σ of the others.
"""
from fractions import Fraction

import numpy as np


def _round(fr, FT):
    """Correctly rounded FT(::BigFloat) of an exact rational."""
    x = FT(float(fr))
    if FT is np.float64:
        return x
    best = min((x, np.nextafter(x, FT(np.inf)), np.nextafter(x, FT(-np.inf))),
               key=lambda c: (abs(Fraction(float(c)) - fr), int(np.float32(c).view(np.uint32)) & 1))
    return FT(best)


def stencil_coefficients(FT, i, r, coord, first, order):
    """stencil_coefficients(FT, i, r, xr, xi; order) with xr = xi = coord; ``coord[n - first]`` is the coordinate of index n."""
    FT = np.dtype(FT).type
    x = lambda n: FT(coord[n - first])
    here = x(i)
    nodes = [x(i - (r - q + 1)) for q in range(order + 1)]
    exact = [Fraction(0)] * order
    for j in range(order):
        for m in range(j + 1, order + 1):
            numerator = None
            for l in range(order + 1):
                if l == m:
                    continue
                p = None
                for q in range(order + 1):
                    if q != m and q != l:
                        term = FT(here - nodes[q])
                        p = term if p is None else FT(p * term)
                numerator = p if numerator is None else FT(numerator + p)
            denominator = None
            for l in range(order + 1):
                if l != m:
                    term = FT(nodes[m] - nodes[l])
                    denominator = term if denominator is None else FT(denominator * term)
            spacing = FT(x(i - (r - j)) - x(i - (r - j + 1)))
            exact[j] += Fraction(float(FT(FT(numerator / denominator) * spacing)))
    c = [_round(v, FT) for v in exact[:-1]]
    total = c[0]
    for v in c[1:]:
        total = FT(total + v)
    return c + [FT(FT(1) - total)]


def z_nodes_with_halo(z_faces, FT, halo):
    """Face and Center coordinates of a Bounded stretched z with ``halo`` extra points on either side
    (src/Grids/grid_generation.jl:34-95: the halo continues with the first / last spacing) → (faces, first face index,
    centres, first centre index)."""
    FT = np.dtype(FT).type
    Fi = np.asarray(z_faces, dtype=FT)
    N = len(Fi) - 1

    def seqsum(v):
        s = FT(0)
        for t in v:
            s = FT(s + t)
        return s

    dlo, dhi = [FT(Fi[1] - Fi[0])] * halo, [FT(Fi[N] - Fi[N - 1])] * halo
    lo = [FT(Fi[0] - seqsum(dlo[q:])) for q in range(halo)]
    hi = [FT(Fi[N] + seqsum(dhi[q:])) for q in range(halo)][::-1]
    F = np.array(lo + list(Fi) + hi, dtype=FT)                               # indices 1 − halo … N + 1 + halo
    C = np.array([FT(FT(F[q + 1] + F[q]) / FT(2)) for q in range(len(F) - 1)], dtype=FT)   # indices 1 − halo … N + halo
    return F, 1 - halo, C, 1 - halo


def weno_z_tables(z_faces, FT, buffer, halo):
    """The two arrays nh_model_set_weno_z_coefficients takes for one buffer: {0 (Face targets), 1 (Center targets)} →
    (Nz + 2) × (buffer + 1) × buffer values, entry [i][r + 1][t] (compute_reconstruction_coefficients on the grid rebuilt with one
    more halo point, :251-272)."""
    FT = np.dtype(FT).type
    N = len(z_faces) - 1
    F, f0, C, c0 = z_nodes_with_halo(z_faces, FT, max(halo, buffer) + 1)
    out = {}
    for location, (coord, first) in enumerate(((F, f0), (C, c0))):
        a = np.empty((N + 2, buffer + 1, buffer), dtype=FT)
        for i in range(N + 2):
            for r in range(-1, buffer):
                a[i, r + 1, :] = stencil_coefficients(FT, i, r, coord, first, buffer)
        out[location] = a
    return out
