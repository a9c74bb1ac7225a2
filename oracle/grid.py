"""RectilinearGrid restatement (oracle; test infrastructure only).

Follows the reference's
  src/Grids/rectilinear_grid.jl:248-275        (constructor, argument meaning)
  src/Grids/grid_generation.jl:34-95           (stretched coordinate: faces, centres, Δᶜ, Δᶠ incl. halos)
  src/Grids/grid_generation.jl:98-135          (regular coordinate: Δ = FT(BigFloat(L)/N))
  src/Grids/grid_generation.jl:138-153         (Flat: Δ = 1, L = 1)
  src/Grids/grid_utils.jl:66-72,52-58          (total_length / length of Face vs Center per topology)
  src/Grids/input_validation.jl:66-92          (default halo 3, Flat halo 0)

Index convention: all public indices are the reference's 1-based indices (first interior cell = 1,
halo cells ≤ 0 and > N); ``OffsetVec`` mirrors an OffsetArray over a dense vector.
"""
from fractions import Fraction

import numpy as np

Periodic, Bounded, Flat = "Periodic", "Bounded", "Flat"
_TOPOS = (Periodic, Bounded, Flat)


def round_fraction(fr, FT):
    """Correctly rounded conversion of an exact rational to FT (Julia: FT(::BigFloat))."""
    FT = np.dtype(FT).type
    x = FT(float(fr))
    if FT is np.float64:
        return x
    # float32: guard against double rounding by checking the neighbours exactly
    cands = [x, np.nextafter(x, FT(np.inf)), np.nextafter(x, FT(-np.inf))]
    best = min(cands, key=lambda c: (abs(Fraction(float(c)) - fr), int(np.float32(c).view(np.uint32)) & 1))
    return FT(best)


def cbrt_correctly_rounded(x, FT):
    """Correctly rounded cube root of a positive FT number by exact integer arithmetic (implementation-independent: the
    CUDA library's host code computes the same bits; the reference's Base.cbrt is accurate to < 1 ulp)."""
    import math
    FT = np.dtype(FT).type
    P = 53 if FT == np.float64 else 24
    if not (FT(x) > 0) or not np.isfinite(FT(x)):
        return FT(np.cbrt(FT(x)))
    f, E = math.frexp(float(FT(x)))
    m = int(math.ldexp(f, P))
    q, rem = divmod(E, 3)
    S = P if (m << rem) >= (1 << P) else P + 1
    X = m << (rem - P + 3 * S)
    r = 1 << ((X.bit_length() + 2) // 3)
    while True:                                  # integer Newton: ⌊cbrt(X)⌋
        nr = (2 * r + X // (r * r)) // 3
        if nr >= r:
            break
        r = nr
    return FT(math.ldexp((r + 1) >> 1, q - (S - 1)))


class OffsetVec:
    """Dense vector with an arbitrary first index (mirrors OffsetArrays.OffsetVector)."""

    def __init__(self, a, first):
        self.a = np.asarray(a)
        self.first = int(first)

    @property
    def last(self):
        return self.first + len(self.a) - 1

    def __getitem__(self, idx):
        return self.a[idx - self.first]

    def r(self, lo, hi):
        """Inclusive range lo:hi as a numpy view."""
        return self.a[lo - self.first: hi - self.first + 1]

    def __len__(self):
        return len(self.a)


class _Coord:
    """One dimension of the grid."""

    def __init__(self, FT, topo, N, H, spec, name):
        self.FT, self.topo, self.N, self.H, self.name = FT, topo, N, H, name
        self.regular = True
        if topo == Flat:
            # grid_generation.jl:138-153
            self.L = FT(1)
            self.dc = FT(1)   # Δ at centres  (Δxᶜᵃᵃ)
            self.df = FT(1)   # Δ at faces    (Δxᶠᵃᵃ)
            self.F = self.C = None
            return
        if isinstance(spec, tuple) and len(spec) == 2 and all(np.isscalar(s) for s in spec):
            self._regular(spec)
        else:
            self._stretched(spec)

    # grid_generation.jl:98-135
    def _regular(self, interval):
        FT, N, H = self.FT, self.N, self.H
        c1, c2 = (Fraction(float(FT(interval[0]))), Fraction(float(FT(interval[1]))))
        if not c1 < c2:
            raise ValueError(f"{self.name} must be an increasing interval!")
        L = c2 - c1
        d = L / N
        self.L = round_fraction(L, FT)
        self.dc = self.df = round_fraction(d, FT)
        TF = N + 2 * H + (1 if self.topo == Bounded else 0)
        TC = N + 2 * H
        Fm = c1 - H * d
        Cm = Fm + d / 2
        self.F = OffsetVec(np.array([round_fraction(Fm + i * d, FT) for i in range(TF)], dtype=FT), 1 - H)
        self.C = OffsetVec(np.array([round_fraction(Cm + i * d, FT) for i in range(TC)], dtype=FT), 1 - H)

    # grid_generation.jl:34-95
    def _stretched(self, spec):
        FT, N, H, topo = self.FT, self.N, self.H, self.topo
        self.regular = False
        if callable(spec):
            Fi = np.array([spec(i) for i in range(1, N + 2)], dtype=FT)
        else:
            Fi = np.asarray(spec, dtype=FT)
            if len(Fi) != N + 1:
                raise ValueError(f"length({self.name}) must be N+1 = {N + 1}")
        if not np.all(np.diff(Fi) >= 0):
            raise ValueError(f"The values of {self.name} must increase as the index increases!")
        self.L = FT(Fi[N] - Fi[0])

        def seqsum(v):
            s = FT(0)
            for x in v:
                s = FT(s + x)
            return s

        if topo == Bounded:
            dlo = [FT(Fi[1] - Fi[0])] * H
            dhi = [FT(Fi[N] - Fi[N - 1])] * H            # already "reversed" (constant)
        else:
            # lower halo spacings repeat the top interior spacings; upper repeat the bottom ones
            dlo = [FT(Fi[N - H + i] - Fi[N - H + i - 1]) for i in range(1, H + 1)]
            dhi = [FT(Fi[i] - Fi[i - 1]) for i in range(1, H + 1)][::-1]
        Flo = [FT(Fi[0] - seqsum(dlo[i:])) for i in range(H)]
        Fhi = [FT(Fi[N] + seqsum(dhi[i:])) for i in range(H)][::-1]
        F = np.array(Flo + list(Fi) + Fhi, dtype=FT)          # N + 1 + 2H entries
        TC = N + 2 * H
        C = np.array([FT(FT(F[i + 1] + F[i]) / FT(2)) for i in range(TC)], dtype=FT)
        d_raw = np.array([FT(C[i] - C[i - 1]) for i in range(1, TC)], dtype=FT)   # TC-1 entries
        TF = N + 2 * H + (1 if topo == Bounded else 0)
        F = F[:TF]
        dc = np.array([FT(F[i + 1] - F[i]) for i in range(TF - 1)], dtype=FT)
        # Δᶠ[idx] = C[idx] - C[idx-1]; the two lowest entries replicate the first difference
        df = np.concatenate([[d_raw[0], d_raw[0]], d_raw]).astype(FT)             # TC+1 entries
        self.F = OffsetVec(F, 1 - H)
        self.C = OffsetVec(C, 1 - H)
        self.dc = OffsetVec(dc, 1 - H)
        self.df = OffsetVec(df, -H)

    def dc_at(self, lo, hi):
        """Δ at centres for indices lo:hi (scalar broadcast if regular)."""
        if self.regular:
            return self.dc
        return self.dc.r(lo, hi)

    def df_at(self, lo, hi):
        if self.regular:
            return self.df
        return self.df.r(lo, hi)


class RectilinearGrid:
    """RectilinearGrid(FT; size, x, y, z, halo, extent, topology) — rectilinear_grid.jl:248-256."""

    def __init__(self, FT=np.float64, size=None, x=None, y=None, z=None, halo=None,
                 extent=None, topology=(Periodic, Periodic, Bounded)):
        FT = np.dtype(FT).type
        topo = tuple(topology)
        if len(topo) != 3 or any(t not in _TOPOS for t in topo):
            raise ValueError("topology must be a 3-tuple of Periodic/Bounded/Flat")
        nonflat = [d for d in range(3) if topo[d] != Flat]

        def inflate(val, default):
            if val is None:
                return None
            val = (val,) if np.isscalar(val) else tuple(val)
            if len(val) == 3:
                return val
            if len(val) != len(nonflat):
                raise ValueError("size/halo/extent must have one entry per non-Flat dimension")
            out = [default] * 3
            for n, d in enumerate(nonflat):
                out[d] = val[n]
            return tuple(out)

        size = inflate(size, 1)
        if halo is None:
            halo = tuple(min(3, size[d]) for d in range(3))   # input_validation.jl:66-72
        else:
            halo = inflate(halo, 0)
        halo = tuple(0 if topo[d] == Flat else int(halo[d]) for d in range(3))
        for d in range(2):
            if halo[d] > size[d]:
                raise ValueError(f"halo={halo[d]} must be ≤ size={size[d]} for coordinate {'xyz'[d]}")
        specs = [x, y, z]
        if extent is not None:
            if any(s is not None for s in specs):
                raise ValueError("Cannot specify both 'extent' and 'x, y, z' keyword arguments.")
            ext = inflate(extent, 0)
            # rectilinear_grid.jl / input_validation.jl:110-125: x,y from 0, z from -Lz to 0
            specs = [(0, ext[0]), (0, ext[1]), (-ext[2], 0)]
        for d in range(3):
            if topo[d] != Flat and specs[d] is None:
                raise ValueError(f"Must supply extent or {'xyz'[d]} keyword when {'xyz'[d]}-direction is {topo[d]}")
        self.FT = FT
        self._specs = specs
        self.topology = topo
        self.N = tuple(int(n) for n in size)
        self.H = halo
        self.coords = [_Coord(FT, topo[d], self.N[d], self.H[d], specs[d], "xyz"[d]) for d in range(3)]
        self.Nx, self.Ny, self.Nz = self.N
        self.Hx, self.Hy, self.Hz = self.H
        self.Lx, self.Ly, self.Lz = (c.L for c in self.coords)

    def with_halo(self, halo):
        """Grids.with_halo: rebuild the same grid with a different halo."""
        g = object.__new__(RectilinearGrid)
        g.__dict__.update(self.__dict__)
        g.H = tuple(0 if self.topology[d] == Flat else int(halo[d]) for d in range(3))
        g.Hx, g.Hy, g.Hz = g.H
        g.coords = [_Coord(self.FT, self.topology[d], self.N[d], g.H[d], self._specs[d], "xyz"[d])
                    for d in range(3)]
        return g

    # field sizes --------------------------------------------------------------------------
    def length(self, loc, d):
        """Interior length of a field at loc ('c'/'f') along dim d (grid_utils.jl:52-58)."""
        return self.N[d] + (1 if (loc == "f" and self.topology[d] == Bounded) else 0)

    def regular_dim(self, d):
        return self.coords[d].regular

    def __repr__(self):
        return (f"RectilinearGrid{{{self.FT.__name__}}} {self.N} halo {self.H} topology {self.topology}")
