"""Advection schemes and reconstructions (oracle; test infrastructure only).

Follows the reference's
  src/Advection/reconstruction_coefficients.jl:74-115,138-140    (Lagrange FV coefficients in BigFloat → FT, last = 1 - Σ)
  src/Advection/reconstruction_coefficients.jl:163-199           (uniform stencil expression: coefficient order, left-assoc +)
  src/Advection/centered_reconstruction.jl:39-59,100-112         (Centered{B}; buffer_scheme = Centered(order-2))
  src/Advection/upwind_biased_reconstruction.jl:41-66,101-140    (UpwindBiased{B}; advecting_velocity_scheme; left/right)
  src/Advection/weno_reconstruction.jl:98-124                    (WENO{B}; buffer_scheme = WENO(order-2); Centered(order-1))
  src/Advection/weno_interpolants.jl:70,77-93,181-183,213-225,299-348,386-526  (Z-WENO weights, stencils, reconstruction)
  src/Advection/topologically_conditional_interpolation.jl:44-125 (order reduction near Bounded walls)
  src/Advection/flat_advective_fluxes.jl:26-44                   (Flat direction: interpolation is the identity)

A "getter" is a callable ``get(ri, rj, rk) -> ndarray`` over 1-based inclusive index ranges; a Field's
``.at`` is one, and so is ``lambda *r: area(*r) * U.at(*r)`` (the reference's ``Ax_qᶠᶜᶜ`` function argument).
"""
from fractions import Fraction

import numpy as np

from .grid import Bounded, Flat, round_fraction


# ---------------------------------------------------------------------------------------------
# coefficients — reconstruction_coefficients.jl:100-115 evaluated on the uniform grid xr = xi = 1:100, i = 50
# ---------------------------------------------------------------------------------------------

def stencil_coefficients(FT, r, order):
    FT = np.dtype(FT).type
    exact = []
    for j in range(order):
        acc = Fraction(0)
        for m in range(j + 1, order + 1):
            num = Fraction(0)
            for l in range(order + 1):
                if l == m:
                    continue
                p = Fraction(1)
                for q in range(order + 1):
                    if q != m and q != l:
                        p *= (r - q + 1)
                num += p
            den = Fraction(1)
            for l in range(order + 1):
                if l != m:
                    den *= (m - l)
            acc += num / den
        exact.append(acc)
    c = [round_fraction(x, FT) for x in exact[:-1]]
    s = c[0]
    for x in c[1:]:
        s = FT(s + x)
    return tuple(c) + (FT(FT(1) - s),)


def stretched_stencil_coefficients(FT, i, r, coord, order):
    """stencil_coefficients(FT, i, r, xr, xi; order) with xr = xi = ``coord`` (reconstruction_coefficients.jl:100-115,74-83), the form
    create_reconstruction_coefficients (:316-324) calls for a stretched direction: numerators, denominators and spacings in FT
    arithmetic (left folds, as Julia's sum / prod over a generator), the j-th coefficient accumulated exactly (BigFloat there),
    rounded to FT, and the last coefficient = 1 − Σ of the others in FT."""
    FT = np.dtype(FT).type
    x = lambda idx: FT(coord[idx])
    exact = [Fraction(0)] * order
    for j in range(order):
        for m in range(j + 1, order + 1):
            numerator = None
            for l in range(order + 1):
                if l == m:
                    continue
                p = None
                for q in range(order + 1):
                    if q == m or q == l:
                        continue
                    term = FT(x(i) - x(i - (r - q + 1)))
                    p = term if p is None else FT(p * term)
                numerator = p if numerator is None else FT(numerator + p)
            denominator = None
            for l in range(order + 1):
                if l == m:
                    continue
                term = FT(x(i - (r - m + 1)) - x(i - (r - l + 1)))
                denominator = term if denominator is None else FT(denominator * term)
            value = FT(FT(numerator / denominator) * FT(x(i - (r - j)) - x(i - (r - j + 1))))
            exact[j] += Fraction(float(value))
    c = [round_fraction(v, FT) for v in exact[:-1]]
    total = c[0]
    for v in c[1:]:
        total = FT(total + v)
    return tuple(c) + (FT(FT(1) - total),)


def stretched_weno_tables(grid, FT, B, d=2):
    """compute_reconstruction_coefficients(grid, FT, :WENO; order = B) for the stretched direction d (:251-272,306-314): for the Face
    and the Center coordinate of the grid rebuilt with one more halo point, one table per stencil r = −1 … B−1, entries i = 0 … N+1.
    → {"f" | "c": {r: [B-tuples for i = 0 … N+1]}}"""
    H = tuple(max(h, B) + 1 if grid.topology[e] != Flat else 0 for e, h in enumerate(grid.H))
    c = grid.with_halo(H).coords[d]
    N = grid.N[d]
    return {loc: {r: [stretched_stencil_coefficients(FT, i, r, coord, B) for i in range(N + 2)] for r in range(-1, B)}
            for loc, coord in (("f", c.F), ("c", c.C))}


class _Scheme:
    kind = None

    def __repr__(self):
        return f"{self.kind}(order={self.order})"

    def __eq__(self, other):
        return type(self) is type(other) and self.order == other.order and self.FT == other.FT

    def __hash__(self):
        return hash((self.kind, self.order, self.FT))


class Centered(_Scheme):
    kind = "Centered"

    def __init__(self, FT=np.float64, order=2):
        if order % 2 != 0:
            raise ValueError("Centered reconstruction scheme is defined only for even orders")
        self.FT = np.dtype(FT).type
        self.order = order
        self.B = order // 2
        self.buffer_scheme = Centered(FT, order - 2) if self.B > 1 else None
        self.advecting_velocity_scheme = self
        self.sym_coeff = stencil_coefficients(FT, self.B - 1, 2 * self.B)


class UpwindBiased(_Scheme):
    kind = "UpwindBiased"

    def __init__(self, FT=np.float64, order=3):
        if order % 2 == 0:
            raise ValueError("UpwindBiased reconstruction scheme is defined only for odd orders")
        self.FT = FT = np.dtype(FT).type
        self.order = order
        self.B = B = (order + 1) // 2
        if B > 1:
            self.advecting_velocity_scheme = Centered(FT, order - 1)
            self.buffer_scheme = UpwindBiased(FT, order - 2)
            self.left_coeff = stencil_coefficients(FT, B - 2, 2 * B - 1)
            self.right_coeff = stencil_coefficients(FT, B - 1, 2 * B - 1)
        else:
            self.advecting_velocity_scheme = Centered(FT, 2)
            self.buffer_scheme = None
            self.left_coeff = self.right_coeff = (FT(1),)


_CSTAR = {2: (Fraction(2, 3), Fraction(1, 3)),
          3: (Fraction(3, 10), Fraction(3, 5), Fraction(1, 10))}
_SMOOTH = {2: ((1, -2, 1), (1, -2, 1)),
           3: ((10, -31, 11, 25, -19, 4), (4, -13, 5, 13, -13, 4), (4, -19, 11, 25, -31, 10))}


def WENO(FT=np.float64, order=5, grid=None):
    """WENO(FT; order, grid): order < 3 returns UpwindBiased(order=1) (weno_reconstruction.jl:98-124).  With ``grid`` the scheme
    carries per-index reconstruction coefficients for the grid's stretched direction (z on this path); the Centered and
    UpwindBiased schemes it hands down keep uniform coefficients, as in the reference (centered_reconstruction.jl:47-58,
    upwind_biased_reconstruction.jl:51-65)."""
    if grid is not None:
        FT = grid.FT
    if order % 2 == 0:
        raise ValueError("WENO reconstruction scheme is defined only for odd orders")
    if order < 3:
        return UpwindBiased(FT, 1)
    return _WENO(FT, order, grid)


class _WENO(_Scheme):
    kind = "WENO"

    def __init__(self, FT, order, grid=None):
        if order > 5:
            raise NotImplementedError("oracle restates WENO orders 3 and 5 only")
        self.FT = FT = np.dtype(FT).type
        self.order = order
        self.B = B = (order + 1) // 2
        self.advecting_velocity_scheme = Centered(FT, order - 1)
        self.buffer_scheme = WENO(FT, order - 2, grid)
        self.grid = grid
        # ztab[loc][r][i]: stretched coefficients of the z direction (None ⇒ uniform, reconstruction_coefficients.jl:276-283)
        self.ztab = stretched_weno_tables(grid, FT, B) if (grid is not None and not grid.regular_dim(2)) else None
        self.coeff = tuple(stencil_coefficients(FT, r, B) for r in range(B))
        self.cstar = tuple(round_fraction(c, FT) for c in _CSTAR[B])
        self.smooth = tuple(tuple(FT(c) for c in row) for row in _SMOOTH[B])
        self.eps = FT(np.float32(1e-8))          # ε = 1f-8, a Float32 literal (weno_interpolants.jl:70)


def is_upwind(scheme):
    return isinstance(scheme, (UpwindBiased, _WENO))


# ---------------------------------------------------------------------------------------------
# region helpers
# ---------------------------------------------------------------------------------------------

def shift(R, d, s):
    R = list(R)
    R[d] = (R[d][0] + s, R[d][1] + s)
    return R


def _index_mask(R, d, cond):
    idx = np.arange(R[d][0], R[d][1] + 1)
    shape = [1, 1, 1]
    shape[d] = len(idx)
    return cond(idx).reshape(shape)


# ---------------------------------------------------------------------------------------------
# symmetric interpolation
# ---------------------------------------------------------------------------------------------

def _inner_symmetric(scheme, get, d, loc, R):
    """inner_symmetric_interpolate_*ᶠ at index i (+1 if the target is a Center) — uniform coefficients."""
    cs = scheme.advecting_velocity_scheme if is_upwind(scheme) else scheme
    B, coeff = cs.B, cs.sym_coeff
    s0 = 1 if loc == "c" else 0
    order = 2 * B
    out = None
    for idx in range(1, order + 1):          # n = idx, offset c = n - B - 1, coefficient coeff[order - idx + 1]
        c = idx - B - 1
        term = coeff[order - idx] * get(*shift(R, d, c + s0))
        out = term if out is None else out + term
    return out


def symmetric_interpolate(grid, scheme, get, d, loc, R):
    """_symmetric_interpolate_{x,y,z}{ᶠ,ᶜ} incl. order reduction near Bounded walls."""
    if grid.topology[d] == Flat:
        return get(*R)
    high = _inner_symmetric(scheme, get, d, loc, R)
    if grid.topology[d] != Bounded or scheme.B == 1:
        return high
    B, N = scheme.B, grid.N[d]
    if loc == "f":
        mask = _index_mask(R, d, lambda i: (i >= B + 1) & (i <= N + 1 - B))
    else:
        mask = _index_mask(R, d, lambda i: (i >= B) & (i <= N + 1 - B))
    if mask.all():
        return high
    low = symmetric_interpolate(grid, scheme.buffer_scheme, get, d, loc, R)
    return np.where(mask, high, low)


# ---------------------------------------------------------------------------------------------
# biased interpolation
# ---------------------------------------------------------------------------------------------

def _upwind_inner(scheme, get, d, s0, R, left):
    B = scheme.B
    if B == 1:
        return np.where(left, scheme.left_coeff[0] * get(*shift(R, d, -1 + s0)),
                        scheme.right_coeff[0] * get(*shift(R, d, 0 + s0)))
    N = 2 * B - 1
    outs = []
    for coeff, rng in ((scheme.left_coeff, range(1, N + 1)), (scheme.right_coeff, range(2, N + 2))):
        out = None
        for idx, n in enumerate(rng, start=1):
            c = n - B - 1
            term = coeff[N - idx] * get(*shift(R, d, c + s0))
            out = term if out is None else out + term
        outs.append(out)
    return np.where(left, outs[0], outs[1])


def weno_window_offsets(B):
    """load_weno_stencil(buffer, dir) — weno_interpolants.jl:376-404: the 2B values ψ[i + c], c = -B … B-1."""
    return range(-B, B)


def _weno_inner(scheme, get, d, s0, R, left):
    """inner_biased_interpolate for WENO{B}: stencil selection, Z-WENO weights, reconstruction."""
    B, FT = scheme.B, scheme.FT
    S = [get(*shift(R, d, c + s0)) for c in weno_window_offsets(B)]                  # S[0..2B-1] ↔ reference S[1..2B]
    # q[m], m = 0..2B-2: left bias takes S[1..2B-1], right bias takes S[2B..2] (reversed)
    q = [np.where(left, S[m], S[2 * B - 1 - m]) for m in range(2 * B - 1)]
    # ψ_r = (q[B-1-r], ..., q[2B-2-r])  (S₀ is the most upwind-centred stencil)
    psi = [tuple(q[B - 1 - r + t] for t in range(B)) for r in range(B)]
    beta = []
    for r in range(B):
        C, p = scheme.smooth[r], psi[r]
        if B == 3:
            b = (p[0] * (C[0] * p[0] + C[1] * p[1] + C[2] * p[2])
                 + p[1] * (C[3] * p[1] + C[4] * p[2])
                 + p[2] * p[2] * C[5])
        else:
            b = p[0] * (C[0] * p[0] + C[1] * p[1]) + p[1] * p[1] * C[2]
        beta.append(b)
    tau = np.abs(beta[0] - beta[B - 1])
    one = FT(1)
    alpha = []
    for r in range(B):
        x = tau / (beta[r] + scheme.eps)
        alpha.append(scheme.cstar[r] * (one + x * x))
    asum = alpha[0]
    for a in alpha[1:]:
        asum = asum + a
    out = None
    ztab = getattr(scheme, "ztab", None) if d == 2 else None
    if ztab is not None:
        # coeff_p for a stretched direction (weno_interpolants.jl:124-127): the table of stencil r at the target index for a left
        # bias, the reversed table of stencil B − 2 − r for a right bias (ψ is stored in reversed order there)
        tab = ztab["c" if s0 else "f"]
        ks = range(R[2][0], R[2][1] + 1)
        if ks[0] < 0 or ks[-1] > len(tab[0]) - 1:
            raise IndexError("stretched WENO coefficients exist for indices 0 … N+1 only")
    for r in range(B):
        w = alpha[r] / asum
        c, p = scheme.coeff[r], psi[r]
        if ztab is not None:
            cl = [np.array([tab[r][k][t] for k in ks], dtype=FT).reshape(1, 1, -1) for t in range(B)]
            cr = [np.array([tab[B - 2 - r][k][B - 1 - t] for k in ks], dtype=FT).reshape(1, 1, -1) for t in range(B)]
            c = [np.where(left, cl[t], cr[t]) for t in range(B)]
        pr = c[0] * p[0]
        for t in range(1, B):
            pr = pr + c[t] * p[t]
        term = w * pr
        out = term if out is None else out + term
    return out


def biased_interpolate(grid, scheme, get, d, loc, R, left):
    """_biased_interpolate_{x,y,z}{ᶠ,ᶜ}(…, bias, ψ): ``left`` is the boolean array ``ũ > 0`` (LeftBias)."""
    if grid.topology[d] == Flat:
        return get(*R)
    if not is_upwind(scheme):
        return symmetric_interpolate(grid, scheme, get, d, loc, R)
    s0 = 1 if loc == "c" else 0
    if isinstance(scheme, _WENO):
        high = _weno_inner(scheme, get, d, s0, R, left)
    else:
        high = _upwind_inner(scheme, get, d, s0, R, left)
    if grid.topology[d] != Bounded or scheme.B == 1:
        return high
    B, N = scheme.B, grid.N[d]
    if loc == "f":
        mask = _index_mask(R, d, lambda i: (i >= B + 1) & (i <= N + 1 - (B - 1)) & (i >= B) & (i <= N + 1 - B))
    else:
        mask = _index_mask(R, d, lambda i: (i >= B) & (i <= N + 1 - (B - 1)) & (i >= B - 1) & (i <= N + 1 - B))
    if mask.all():
        return high
    low = biased_interpolate(grid, scheme.buffer_scheme, get, d, loc, R, left)
    return np.where(mask, high, low)


# ---------------------------------------------------------------------------------------------
# order adaptation — adapt_advection_order.jl:18-51,73-110
# ---------------------------------------------------------------------------------------------

def adapt_advection_order(advection, grid):
    """Returns a per-direction list [sx, sy, sz] (the reference's FluxFormAdvection when anything changed)."""
    if advection is None:
        return [None, None, None]
    if isinstance(advection, (list, tuple)):
        per_dir = list(advection)
    else:
        per_dir = [advection] * 3
    out = []
    for d in range(3):
        s = per_dir[d]
        if grid.topology[d] == Flat or s is None:
            out.append(s)
            continue
        N = grid.N[d]
        if N >= s.B:
            out.append(s)
        elif isinstance(s, Centered):
            out.append(Centered(s.FT, 2 * N))
        elif isinstance(s, UpwindBiased):
            out.append(UpwindBiased(s.FT, 2 * N - 1))
        else:
            out.append(WENO(s.FT, 2 * N - 1, getattr(s, "grid", None)))
    return out
