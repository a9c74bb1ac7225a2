"""oceananigans.jl_b200 — host-side mirror of the reference's API for the NonhydrostaticModel time step, driving the
hand-written sm_100a kernels in libnhstep.so through the C ABI of include/nhstep.h.

Names and argument meaning follow samlewin/Oceananigans.jl v0.96.19:
  RectilinearGrid               src/Grids/rectilinear_grid.jl:248-275
  NonhydrostaticModel           src/Models/NonhydrostaticModels/nonhydrostatic_model.jl:114-239
  set!  (here ``set_model``)    src/Models/NonhydrostaticModels/set_nonhydrostatic_model.jl:33-60
  time_step!, update_state!, compute_tendencies!, calculate_pressure_correction!, pressure_correct_velocities!,
  fill_halo_regions!, solve!    
Julia's ``f!(x)`` is spelled ``f(x)`` here.  There is no CPU path: everything below needs a CUDA device.

(The directory name contains a dot, so it is imported through the top-level shim ``oceananigans_jl_b200``.)
"""
import ctypes as C
from fractions import Fraction

import numpy as np

from . import _lib as L
from ._lib import NhError, lib, check, device_count  # noqa: F401

Periodic, Bounded, Flat = "Periodic", "Bounded", "Flat"
_TOPO = {Periodic: L.PERIODIC, Bounded: L.BOUNDED, Flat: L.FLAT}


def _ft_code(FT):
    FT = np.dtype(FT)
    if FT == np.float64:
        return L.F64
    if FT == np.float32:
        return L.F32
    raise ValueError("FT must be Float32 or Float64")


class RectilinearGrid:
    """RectilinearGrid(FT; size, x, y, z, halo, extent, topology) on one B200 (``device``)."""

    def __init__(self, FT=np.float64, size=None, x=None, y=None, z=None, halo=None, extent=None,
                 topology=(Periodic, Periodic, Bounded), device=0):
        topo = tuple(topology)
        if len(topo) != 3 or any(t not in _TOPO for t in topo):
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
        if size is None:
            raise ValueError("size is required")
        halo = tuple(min(3, size[d]) for d in range(3)) if halo is None else inflate(halo, 0)
        halo = tuple(0 if topo[d] == Flat else int(halo[d]) for d in range(3))
        specs = [x, y, z]
        if extent is not None:
            if any(s is not None for s in specs):
                raise ValueError("Cannot specify both 'extent' and 'x, y, z' keyword arguments.")
            ext = inflate(extent, 0)
            specs = [(0, ext[0]), (0, ext[1]), (-ext[2], 0)]
        for d in range(3):
            if topo[d] != Flat and specs[d] is None:
                raise ValueError(f"Must supply extent or {'xyz'[d]} keyword when {'xyz'[d]}-direction is {topo[d]}")
        self.FT = np.dtype(FT).type
        self.topology, self.N, self.H, self.device = topo, tuple(int(n) for n in size), halo, device
        self.Nx, self.Ny, self.Nz = self.N
        self._specs = specs
        desc = L.GridDesc()
        desc.float_type = _ft_code(FT)
        self._zfaces = None
        for d in range(3):
            desc.N[d], desc.H[d], desc.topology[d] = self.N[d], self.H[d], _TOPO[topo[d]]
        for d, name in enumerate("xyz"):
            s = specs[d]
            if topo[d] == Flat or s is None:
                continue
            regular = isinstance(s, tuple) and len(s) == 2 and all(np.isscalar(v) for v in s)
            if regular:
                getattr(desc, name)[0], getattr(desc, name)[1] = float(s[0]), float(s[1])
            else:
                if d != 2:
                    raise NotImplementedError("the B200 path supports stretched spacing in z only (SURVEY.md §8a a1)")
                faces = np.array([s(k) for k in range(1, self.N[2] + 2)] if callable(s) else s, dtype=np.float64)
                if faces.shape != (self.N[2] + 1,):
                    raise ValueError(f"length(z) must be N+1 = {self.N[2] + 1}")
                self._zfaces = np.ascontiguousarray(faces)
                desc.z_faces = self._zfaces.ctypes.data_as(C.POINTER(C.c_double))
        self._h = C.c_void_p()
        check(lib().nh_grid_create(C.byref(desc), device, C.byref(self._h)))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().nh_grid_destroy(self._h)
            self._h = None

    @property
    def z_stretched(self):
        return self._zfaces is not None

    def spacings(self):
        """(Δx, Δy, Δzᵃᵃᶜ[1-Hz … Nz+Hz], Δzᵃᵃᶠ[1-Hz … Nz+Hz]) as the library holds them (xspacings / yspacings / zspacings,
        src/Grids/nodes_and_spacings.jl); the z arrays start at index 1 - Hz."""
        n = self.N[2] + 2 * self.H[2]
        dxdy, dzc, dzf = (C.c_double * 2)(), (C.c_double * n)(), (C.c_double * n)()
        check(lib().nh_grid_spacings(self._h, dxdy, dzc, dzf, n))
        return (self.FT(dxdy[0]), self.FT(dxdy[1]), np.array(dzc, dtype=np.float64).astype(self.FT),
                np.array(dzf, dtype=np.float64).astype(self.FT))

    def extents(self):
        """(Lx, Ly, Lz) in the grid's float type: c₂ − c₁ of the FT-converted interval, 1 for a Flat direction."""
        FT, out = self.FT, []
        for d in range(3):
            if self.topology[d] == Flat:
                out.append(FT(1))
            elif d == 2 and self._zfaces is not None:
                F = self._zfaces.astype(FT)
                out.append(FT(F[-1] - F[0]))
            else:
                c1, c2 = (Fraction(float(FT(v))) for v in self._specs[d])
                out.append(FT(float(c2 - c1)))
        return tuple(out)

    def nodes(self, d, loc):
        """Interior node coordinates along dim d at 'c' or 'f' (src/Grids/grid_generation.jl:98-135, :34-95)."""
        FT, N = self.FT, self.N[d]
        if self.topology[d] == Flat:
            return np.ones(N, dtype=FT)
        n = N + (1 if (loc == "f" and self.topology[d] == Bounded) else 0)
        if d == 2 and self._zfaces is not None:
            F = self._zfaces.astype(FT)
            return F[:n].copy() if loc == "f" else ((F[1:] + F[:-1]) / FT(2)).astype(FT)
        c1, c2 = (Fraction(float(FT(v))) for v in self._specs[d])
        dl = (c2 - c1) / N
        start = c1 if loc == "f" else c1 + dl / 2
        return np.array([float(start + i * dl) for i in range(n)], dtype=np.float64).astype(FT)


def reconstruction_coefficients(FT, bias, buffer, r=0):
    """uniform_reconstruction_coefficients(FT, Val(bias), buffer) (bias = "symmetric" | "left" | "right"), the WENO sub-stencil
    coefficients (bias = "weno", stencil r) or optimal weights (bias = "weno_optimal") as the kernels' tables hold them —
    src/Advection/reconstruction_coefficients.jl:138-140, weno_interpolants.jl:77-93.  Host-only (no device needed)."""
    which = {"symmetric": 0, "left": 1, "right": 2, "weno": 3, "weno_optimal": 4}[bias]
    out, n = (C.c_double * 8)(), C.c_int32(0)
    check(lib().nh_reconstruction_coefficients(_ft_code(FT), which, int(buffer), int(r), out, C.byref(n)))
    return tuple(np.dtype(FT).type(out[q]) for q in range(n.value))


# ---- advection schemes, physics ----------------------------------------------------------------------------
class _Scheme:
    def __init__(self, kind, order):
        self.kind, self.order = kind, order


def Centered(FT=None, order=2, grid=None):
    """Centered(FT; grid, order): with or without ``grid`` the coefficients are the uniform ones, as in the reference
    (centered_reconstruction.jl:47-58 keeps constant coefficients for stretched directions)."""
    if order % 2:
        raise ValueError("Centered reconstruction scheme is defined only for even orders")
    return _Scheme(L.ADV_CENTERED, order)


def UpwindBiased(FT=None, order=3, grid=None):
    """UpwindBiased(FT; grid, order): uniform coefficients with or without ``grid`` (upwind_biased_reconstruction.jl:51-65)."""
    if order % 2 == 0:
        raise ValueError("UpwindBiased reconstruction scheme is defined only for odd orders")
    return _Scheme(L.ADV_UPWIND, order)


def WENO(FT=None, order=5, grid=None):
    """WENO(FT; order, grid) — weno_reconstruction.jl:98-124.  With ``grid`` (``WENO(grid)`` in the reference) the scheme reconstructs
    along the grid's stretched z with per-index coefficients (stretched_weno.py → nh_model_set_weno_z_coefficients); without it the
    uniform coefficients are used in every direction, also on a stretched grid, as in the reference."""
    if isinstance(FT, RectilinearGrid):                  # WENO(grid, …)
        FT, grid = None, FT
    if order % 2 == 0:
        raise ValueError("WENO reconstruction scheme is defined only for odd orders")
    s = _Scheme(L.ADV_WENO, order)
    s.grid = grid
    return s


class FluxFormAdvection:
    def __init__(self, x, y, z):
        self.x, self.y, self.z = x, y, z


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
    required_tracers = ("b",)


class SeawaterBuoyancy:
    """SeawaterBuoyancy with LinearEquationOfState (g = 9.80665, α = 1.67e-4, β = 7.80e-4 by default)."""
    required_tracers = ("T", "S")

    def __init__(self, gravitational_acceleration=9.80665, thermal_expansion=1.67e-4, haline_contraction=7.80e-4):
        self.g, self.alpha, self.beta = gravitational_acceleration, thermal_expansion, haline_contraction


class SmagorinskyLilly:
    """SmagorinskyLilly(FT; C = 0.16, Cb = 1, Pr = 1) — Smagorinskys/lilly_coefficient.jl:108-113; Pr a number or a
    dict by tracer name."""
    kind = L.CLOSURE_SMAGORINSKY_LILLY

    def __init__(self, C=0.16, Cb=1.0, Pr=1.0):
        self.C, self.Cb, self.Pr = C, Cb, Pr

    def Pr_for(self, name):
        return self.Pr[name] if isinstance(self.Pr, dict) else self.Pr


class Smagorinsky(SmagorinskyLilly):
    """Smagorinsky(FT; coefficient = 0.16, Pr = 1) with a number coefficient — Smagorinskys/smagorinsky.jl:78-84,105."""
    kind = L.CLOSURE_SMAGORINSKY

    def __init__(self, coefficient=0.16, Pr=1.0):
        self.C, self.Cb, self.Pr = coefficient, 0.0, Pr


class AnisotropicMinimumDissipation:
    """AnisotropicMinimumDissipation(FT; C = 1/12, Cν = C, Cκ = C (a number or a dict by tracer name)); Cb = nothing —
    anisotropic_minimum_dissipation.jl:110-121."""
    kind = L.CLOSURE_AMD

    def __init__(self, C=1 / 12, Cnu=None, Ckappa=None):
        self.Cnu = C if Cnu is None else Cnu
        self.Ckappa = C if Ckappa is None else Ckappa

    def Ckappa_for(self, name):
        return self.Ckappa[name] if isinstance(self.Ckappa, dict) else self.Ckappa


class ScalarDiffusivity:
    """ScalarDiffusivity([time_discretization,] ν, κ) (isotropic); vertically_implicit = True stands for
    VerticallyImplicitTimeDiscretization()."""
    kind = L.CLOSURE_SCALAR

    def __init__(self, nu=0.0, kappa=0.0, vertically_implicit=False):
        self.nu, self.kappa, self.vitd = nu, kappa, bool(vertically_implicit)

    def kappa_for(self, name):
        return self.kappa[name] if isinstance(self.kappa, dict) else self.kappa


from .model import (Field, NonhydrostaticModel, fill_halo_regions, update_state, compute_tendencies,  # noqa: E402,F401
                    calculate_pressure_correction, pressure_correct_velocities, rk3_substep, ab2_step,
                    cache_previous_tendencies, time_step, time_step_profiled, set_model, FFTBasedPoissonSolver,
                    FourierTridiagonalPoissonSolver, BatchedTridiagonalSolver, BoundaryCondition,
                    FieldBoundaryConditions, ValueBoundaryCondition, GradientBoundaryCondition, FluxBoundaryCondition)
from .checkpointer import (Checkpointer, write_output, checkpoint_path, latest_checkpoint, cleanup_checkpoints,  # noqa: E402,F401
                           set_from_checkpoint, pickup)
