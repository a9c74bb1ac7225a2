"""Field data layout and fill_halo_regions! restatement (oracle; test infrastructure only).

Follows the reference's
  src/Fields/field.jl:19-33, src/Grids/new_data.jl:15-70      (parent array = interior + halos; Face+Bounded → N+1)
  src/BoundaryConditions/field_boundary_conditions.jl:13-37   (default BCs by topology and location)
  src/BoundaryConditions/fill_halo_regions.jl:50-196          (open fill → tasks sorted by BC kind → per-task fill)
  src/BoundaryConditions/fill_halo_regions_periodic.jl:18-111 (periodic copies over the full transverse parent extent)
  src/BoundaryConditions/fill_halo_regions_flux.jl:14-33      (no-flux: mirror the first interior cell, one halo cell)
  src/BoundaryConditions/fill_halo_regions_open.jl:9-70       (impenetrable: wall-normal velocity := 0 at 1 and N+1)
  src/Fields/field_tuples.jl:56-101                           (tupled fill: one task list for all ordinary fields)
  src/BoundaryConditions/fill_halo_regions_value_gradient.jl:7-99  (Value / Gradient: linear extrapolation into one halo cell)
  src/BoundaryConditions/field_boundary_conditions.jl:199-201  (Number conditions are converted to eltype(grid))
"""
import numpy as np

from .grid import Periodic, Bounded, Flat

PBC, FBC, OBC = "periodic", "flux", "open"   # None = no boundary condition (Flat, or face field on a wall)


def default_bcs(grid, loc, prognostic=True):
    """field_boundary_conditions.jl:13-37 — one BC kind per dimension (west/east are always equal by default)."""
    bcs = []
    for d in range(3):
        t = grid.topology[d]
        if t == Periodic:
            bcs.append(PBC)
        elif t == Flat:
            bcs.append(None)
        elif loc[d] == "c":
            bcs.append(FBC)
        else:
            bcs.append(OBC if prognostic else None)
    return bcs


class Field:
    def __init__(self, grid, loc, prognostic=True, name=""):
        self.grid, self.loc, self.name = grid, tuple(loc), name
        self.size = tuple(grid.length(loc[d], d) for d in range(3))
        self.H = grid.H
        self.data = np.zeros(tuple(self.size[d] + 2 * self.H[d] for d in range(3)), dtype=grid.FT)
        self.bcs = default_bcs(grid, self.loc, prognostic)
        # non-default conditions: side_bcs[d][side] = None | ("value" | "gradient" | "flux", number or wall array)
        self.side_bcs = [[None, None], [None, None], [None, None]]

    # 1-based inclusive ranges → parent view
    def at(self, ri, rj, rk):
        H = self.H
        for d, r in enumerate((ri, rj, rk)):
            if r[0] - 1 + H[d] < 0 or r[1] + H[d] > self.data.shape[d]:
                raise IndexError(f"range {r} outside the parent array of '{self.name}' in dim {d}")
        return self.data[ri[0] - 1 + H[0]: ri[1] + H[0],
                         rj[0] - 1 + H[1]: rj[1] + H[1],
                         rk[0] - 1 + H[2]: rk[1] + H[2]]

    @property
    def interior(self):
        return self.at((1, self.size[0]), (1, self.size[1]), (1, self.size[2]))

    def set(self, values):
        self.interior[...] = np.asarray(values, dtype=self.grid.FT)

    def copy(self):
        f = object.__new__(Field)
        f.__dict__.update(self.__dict__)
        f.data = self.data.copy()
        f.bcs = list(self.bcs)
        f.side_bcs = [list(x) for x in self.side_bcs]
        return f

    def set_boundary_condition(self, d, side, kind, value):
        """FieldBoundaryConditions(west/east/south/north/bottom/top = Value/Gradient/FluxBoundaryCondition(number))."""
        if self.bcs[d] != FBC:
            raise ValueError(f"{kind} boundary condition needs a Bounded direction in which '{self.name}' sits at Centers")
        if kind not in ("value", "gradient", "flux"):
            raise ValueError(kind)
        if np.ndim(value) == 0:
            self.side_bcs[d][side] = (kind, self.grid.FT(value))
            return
        # array condition (boundary_condition.jl:118): condition[i, j] over the wall's transverse grid range
        want = tuple(self.grid.N[e] for e in range(3) if e != d)
        a = np.asarray(value, dtype=self.grid.FT)
        if a.shape != want:
            raise ValueError(f"the array condition must have shape {want}, got {a.shape}")
        self.side_bcs[d][side] = (kind, np.expand_dims(a, d).copy())


# ---------------------------------------------------------------------------------------------
# fill_halo_regions!
# ---------------------------------------------------------------------------------------------

def _task_order(kinds):
    """sortperm(bcs_array, lt=fill_first) as Julia's small-array insertion sort executes it
    (fill_halo_regions.jl:121-123,159-177): lt(a, b) is true for every pair except (periodic, non-periodic)."""
    def lt(a, b):
        return not (a == PBC and b != PBC)
    order = [0, 1, 2]
    for i in range(1, 3):
        x = order[i]
        j = i
        while j > 0 and lt(kinds[x], kinds[order[j - 1]]):
            order[j] = order[j - 1]
            j -= 1
        order[j] = x
    return order


def _sl(lo, hi):
    return slice(lo, hi)


def _fill_open(f):
    """fill_halo_regions_open.jl:65-70 with the impenetrable condition (value 0)."""
    g = f.grid
    for d in range(3):
        if f.bcs[d] != OBC or f.loc[d] != "f":
            continue
        # only (f,c,c) in x, (c,f,c) in y, (c,c,f) in z have a filling function
        if not all((f.loc[e] == "c") for e in range(3) if e != d):
            continue
        rng = [(1, g.N[e]) for e in range(3)]
        for idx in (1, g.N[d] + 1):
            r = list(rng)
            r[d] = (idx, idx)
            f.at(*r)[...] = 0


def _fill_flux_dir(f, d):
    """fill_halo_regions_flux.jl:14-33 over the grid-interior transverse range (`:yz` etc.)."""
    g = f.grid
    if f.H[d] == 0:
        return
    rng = [(1, g.N[e]) for e in range(3)]
    N = g.N[d]
    lo_dst, lo_src, hi_dst, hi_src = list(rng), list(rng), list(rng), list(rng)
    lo_dst[d], lo_src[d] = (0, 0), (1, 1)
    hi_dst[d], hi_src[d] = (N + 1, N + 1), (N, N)
    c = g.coords[d]
    for side, dst, src, bidx, sign in ((0, lo_dst, lo_src, 1, -1), (1, hi_dst, hi_src, N + 1, +1)):
        bc = f.side_bcs[d][side]
        if bc is None or bc[0] == "flux":
            f.at(*dst)[...] = f.at(*src)
            continue
        # fill_halo_regions_value_gradient.jl: Δ at the boundary face (flip(loc) = Face index 1 or N+1)
        FT = g.FT
        D = c.df if (c.topo == Flat or c.regular) else c.df[bidx]
        D = FT(D)
        cI = f.at(*src)
        if bc[0] == "gradient":
            grad = bc[1]
        elif side == 0:
            grad = (cI - bc[1]) / (D / FT(2))
        else:
            grad = (bc[1] - cI) / (D / FT(2))
        f.at(*dst)[...] = cI + grad * (D if sign > 0 else -D)


def _fill_periodic_dir(fields, d):
    """fill_halo_regions_periodic.jl:40-111: parent-indexed copies; the transverse extent is the minimum
    parent size over the tuple (parent_size_and_offset, :18-26)."""
    others = [e for e in range(3) if e != d]
    ext = [min(f.data.shape[e] for f in fields) for e in others]
    for f in fields:
        H, N = f.H[d], f.grid.N[d]
        if H == 0:
            continue
        sl = [slice(None)] * 3
        for e, n in zip(others, ext):
            sl[e] = slice(0, n)

        def S(a, b):
            s = list(sl)
            s[d] = slice(a, b)
            return tuple(s)
        p = f.data
        p[S(0, H)] = p[S(N, N + H)]                 # west  ← interior east end
        p[S(N + H, N + 2 * H)] = p[S(H, 2 * H)]     # east  ← interior west end


def fill_halo_regions(fields, fill_boundary_normal_velocities=True):
    """fill_halo_regions!(fields...) for one Field or a tuple/list of Fields sharing a grid."""
    if isinstance(fields, Field):
        fields = [fields]
    fields = [f for f in fields if f is not None]
    if not fields:
        return
    if fill_boundary_normal_velocities:
        for f in fields:
            _fill_open(f)
    # BC kind of each direction's task (a tuple is "periodic" only if every member is periodic)
    kinds = []
    for d in range(3):
        ks = [f.bcs[d] for f in fields]
        kinds.append(PBC if all(k == PBC for k in ks) else "other")
    for d in _task_order(kinds):
        if kinds[d] == PBC:
            _fill_periodic_dir(fields, d)
        else:
            for f in fields:
                if f.bcs[d] == FBC:
                    _fill_flux_dir(f, d)
