"""NonhydrostaticModel, fields, time stepping and solvers over the C ABI (host mirror; see package docstring)."""
import ctypes as C
import os

import numpy as np

from . import _lib as L
from ._lib import lib, check
from . import (RectilinearGrid, Periodic, Bounded, Flat, Centered, FluxFormAdvection, BuoyancyTracer,
               SeawaterBuoyancy, _Scheme)


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class Field:
    """View of one device-resident field of a model (interior(field) / parent(field) copy through the ABI)."""

    def __init__(self, model, fid, name, loc):
        self.model, self.id, self.name, self.loc = model, fid, name, loc

    def _size(self, region):
        s = (C.c_int64 * 3)()
        check(lib().nh_field_size(self.model._h, self.id, region, s))
        return tuple(int(v) for v in s)

    @property
    def size(self):
        return self._size(L.INTERIOR)

    @property
    def parent_size(self):
        return self._size(L.PARENT)

    def _get(self, region, out=None):
        shape = self._size(region)
        if out is None:
            out = np.empty(shape, dtype=self.model.grid.FT, order="F")
        elif out.shape != shape or out.dtype != self.model.grid.FT or not out.flags.f_contiguous:
            raise ValueError(f"out must be a Fortran-ordered {shape} array of the grid's float type")
        check(lib().nh_get_field(self.model._h, self.id, region, _ptr(out), out.size, L.HOST))
        return out

    def interior(self, out=None):
        """interior(field) copied to the host; ``out`` (e.g. from ``pinned_empty``) avoids the allocation."""
        return self._get(L.INTERIOR, out)

    def parent(self, out=None):
        return self._get(L.PARENT, out)

    def _put(self, values, region):
        shape = self._size(region)
        a = np.asfortranarray(np.broadcast_to(np.asarray(values, dtype=self.model.grid.FT), shape))
        check(lib().nh_set_field(self.model._h, self.id, region, _ptr(a), a.size, L.HOST))

    def set(self, values):
        """set!(field, array or f(x,y,z)) — interior only."""
        if callable(values):
            g = self.model.grid
            xs = [g.nodes(d, self.loc[d]) for d in range(3)]
            n = self.size
            xs = [x[:n[d]] for d, x in enumerate(xs)]
            values = np.vectorize(values)(xs[0][:, None, None], xs[1][None, :, None], xs[2][None, None, :])
        self._put(values, L.INTERIOR)

    def set_parent(self, values):
        self._put(values, L.PARENT)


def pinned_empty(shape, FT=np.float64):
    """Fortran-ordered host array in page-locked memory (full-rate PCIe copies through nh_set_field / nh_get_field)."""
    import torch
    t = torch.empty(tuple(reversed(shape)), dtype=torch.float64 if np.dtype(FT) == np.float64 else torch.float32,
                    pin_memory=True)
    a = t.numpy().T
    assert a.flags.f_contiguous
    return a


class _Namespace(dict):
    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e


class BoundaryCondition:
    """BoundaryCondition(classification, condition) — src/BoundaryConditions/boundary_condition.jl:17-60,117-118 with a
    number (converted to eltype(grid), field_boundary_conditions.jl:199-201) or a 2-D array indexed by the wall's two
    transverse grid indices, shape (N_e1, N_e2)."""

    def __init__(self, kind, condition):
        self.kind = kind
        self.condition = float(condition) if np.ndim(condition) == 0 else np.array(condition)


def ValueBoundaryCondition(val):
    return BoundaryCondition(L.BC_VALUE, val)


def GradientBoundaryCondition(val):
    return BoundaryCondition(L.BC_GRADIENT, val)


def FluxBoundaryCondition(val):
    return BoundaryCondition(L.BC_FLUX, val)


_SIDES = {"west": (0, 0), "east": (0, 1), "south": (1, 0), "north": (1, 1), "bottom": (2, 0), "top": (2, 1)}


class FieldBoundaryConditions:
    """FieldBoundaryConditions(; west, east, south, north, bottom, top) — field_boundary_conditions.jl:60-110; sides that
    are not given keep the defaults of :13-37 (periodic / no-flux / impenetrable)."""

    def __init__(self, **sides):
        for k, bc in sides.items():
            if k not in _SIDES:
                raise ValueError(f"unknown side {k}")
            if not isinstance(bc, BoundaryCondition):
                raise ValueError("only Value/Gradient/FluxBoundaryCondition(number) cross the C ABI")
        self.sides = sides


class NonhydrostaticModel:
    """NonhydrostaticModel(; grid, advection, buoyancy, coriolis, closure, tracers, timestepper, hydrostatic_pressure_anomaly,
    boundary_conditions)."""

    def __init__(self, grid, advection="default", buoyancy=None, coriolis=None, closure=None, tracers=(),
                 timestepper="RungeKutta3", hydrostatic_pressure_anomaly="default", chi=0.1, exact_arithmetic=True,
                 boundary_conditions=None):
        if advection == "default":
            advection = Centered(order=2)
        tracers = (tracers,) if isinstance(tracers, str) else tuple(tracers)
        if len(tracers) > L.NH_MAX_TRACERS:
            raise ValueError(f"at most {L.NH_MAX_TRACERS} tracers")
        if buoyancy is not None:
            for req in buoyancy.required_tracers:
                if req not in tracers:
                    raise ValueError(f"buoyancy requires tracer {req}")
        d = L.ModelDesc()
        per_dir = ([advection.x, advection.y, advection.z] if isinstance(advection, FluxFormAdvection)
                   else [advection] * 3)
        for e, s in enumerate(per_dir):
            if s is None:
                d.advection_kind[e], d.advection_order[e] = L.ADV_NONE, 0
            elif isinstance(s, _Scheme):
                d.advection_kind[e], d.advection_order[e] = s.kind, s.order
            else:
                raise ValueError("advection must be Centered/UpwindBiased/WENO, FluxFormAdvection or None")
        if timestepper in ("RungeKutta3", "RK3"):
            d.timestepper = L.RK3
        elif timestepper in ("QuasiAdamsBashforth2", "AB2"):
            d.timestepper = L.AB2
        else:
            raise ValueError(f"unknown timestepper {timestepper}")
        d.ab2_chi = chi
        d.n_tracers = len(tracers)
        if buoyancy is None:
            d.buoyancy_kind = L.BUOY_NONE
        elif isinstance(buoyancy, BuoyancyTracer):
            d.buoyancy_kind = L.BUOY_TRACER
            d.buoyancy_tracer[0] = tracers.index("b")
        elif isinstance(buoyancy, SeawaterBuoyancy):
            d.buoyancy_kind = L.BUOY_SEAWATER
            d.buoyancy_tracer[0], d.buoyancy_tracer[1] = tracers.index("T"), tracers.index("S")
            d.gravitational_acceleration, d.thermal_expansion, d.haline_contraction = buoyancy.g, buoyancy.alpha, buoyancy.beta
        else:
            raise ValueError("unsupported buoyancy formulation")
        if coriolis is not None:
            d.has_coriolis, d.coriolis_f = 1, coriolis.f
        if closure is not None and closure.kind == L.CLOSURE_SCALAR:
            d.has_closure, d.nu = L.CLOSURE_SCALAR, closure.nu
            d.vertically_implicit = int(getattr(closure, "vitd", False))
            for n, name in enumerate(tracers):
                d.kappa[n] = closure.kappa_for(name)
        elif closure is not None and closure.kind == L.CLOSURE_AMD:
            d.has_closure, d.amd_Cnu = L.CLOSURE_AMD, closure.Cnu
            for n, name in enumerate(tracers):
                d.amd_Ckappa[n] = closure.Ckappa_for(name)
        elif closure is not None:
            d.has_closure, d.smagorinsky_C, d.smagorinsky_Cb = closure.kind, closure.C, closure.Cb
            for n, name in enumerate(tracers):
                d.smagorinsky_Pr[n] = closure.Pr_for(name)
        d.hydrostatic_pressure_anomaly = -1 if hydrostatic_pressure_anomaly == "default" else int(bool(hydrostatic_pressure_anomaly))
        d.exact_arithmetic = int(bool(exact_arithmetic))
        self.grid = grid
        self.tracer_names = tracers
        self.timestepper_name = timestepper
        self._h = C.c_void_p()
        check(lib().nh_model_create(grid._h, C.byref(d), C.byref(self._h)))
        # WENO(grid): the stretched z coefficients of the scheme and of its near-wall buffer scheme, generated on the host
        zs = per_dir[2]
        if isinstance(zs, _Scheme) and zs.kind == L.ADV_WENO and getattr(zs, "grid", None) is not None and grid.z_stretched:
            from .stretched_weno import weno_z_tables
            if grid.topology[2] != Bounded:
                raise NotImplementedError("WENO(grid): the stretched z direction must be Bounded (as the Fourier-tridiagonal solve needs)")
            if zs.grid is not grid and not (zs.grid.z_stretched and np.array_equal(zs.grid._zfaces, grid._zfaces)):
                raise ValueError("WENO(grid): the scheme was built for another vertical coordinate than the model's grid")
            top = min((zs.order + 1) // 2, grid.N[2])            # adapt_advection_order lowers the order on a short direction
            for buffer in range(2, top + 1):
                for location, a in weno_z_tables(grid._zfaces, grid.FT, buffer, grid.H[2]).items():
                    a = np.ascontiguousarray(a)
                    check(lib().nh_model_set_weno_z_coefficients(self._h, buffer, location, a.ctypes.data_as(C.c_void_p), a.size,
                                                                 L.HOST))
        for name, fbcs in (boundary_conditions or {}).items():
            names_ = ["u", "v", "w"] + list(tracers)
            if name not in names_:
                raise ValueError(f"boundary_conditions: {name} is not a prognostic field of this model")
            for side, bc in fbcs.sides.items():
                dim, s_ = _SIDES[side]
                if np.ndim(bc.condition) == 0:
                    check(lib().nh_model_set_boundary_condition(self._h, names_.index(name), dim, s_, bc.kind,
                                                                float(bc.condition)))
                    continue
                # array condition: condition[i, j] over the wall's transverse grid range, handed over column-major
                e1, e2 = [e for e in range(3) if e != dim]
                want = (grid.N[e1], grid.N[e2])
                a = np.asarray(bc.condition, dtype=grid.FT)
                if a.shape != want:
                    raise ValueError(f"boundary_conditions: the {side} condition of {name} must have shape {want}, got {a.shape}")
                a = np.asfortranarray(a)
                check(lib().nh_model_set_boundary_condition_array(self._h, names_.index(name), dim, s_, bc.kind,
                                                                  a.ctypes.data_as(C.c_void_p), a.size, L.HOST))
        gi = L.GridDesc()
        check(lib().nh_model_grid_info(self._h, C.byref(gi)))
        self.halo = tuple(gi.H)
        locs = [("f", "c", "c"), ("c", "f", "c"), ("c", "c", "f")] + [("c", "c", "c")] * len(tracers)
        names = ["u", "v", "w"] + list(tracers)
        self.velocities = _Namespace((n, Field(self, i, n, locs[i])) for i, n in enumerate("uvw"))
        self.tracers = _Namespace((n, Field(self, 3 + i, n, locs[3 + i])) for i, n in enumerate(tracers))
        self.pressures = _Namespace(pNHS=Field(self, L.FIELD_PNHS, "pNHS", ("c", "c", "c")))
        if (buoyancy is not None) if hydrostatic_pressure_anomaly == "default" else hydrostatic_pressure_anomaly:
            self.pressures["pHY"] = Field(self, L.FIELD_PHY, "pHY", ("c", "c", "c"))
        self.diffusivity_fields = _Namespace()
        if closure is not None and closure.kind != L.CLOSURE_SCALAR:
            self.diffusivity_fields["nu_e"] = Field(self, L.FIELD_NUE, "nu_e", ("c", "c", "c"))
            if closure.kind == L.CLOSURE_AMD:
                self.diffusivity_fields["kappa_e"] = _Namespace(
                    (n, Field(self, L.FIELD_KAPPAE0 + i, "kappa_e_" + n, ("c", "c", "c"))) for i, n in enumerate(tracers))
        self.timestepper = _Namespace(
            Gn=_Namespace((n, Field(self, L.FIELD_GN0 + i, "Gn_" + n, locs[i])) for i, n in enumerate(names)),
            Gm=_Namespace((n, Field(self, L.FIELD_GM0 + i, "Gm_" + n, locs[i])) for i, n in enumerate(names)))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().nh_model_destroy(self._h)
            self._h = None

    def prognostic_fields(self):
        return list(self.velocities.values()) + list(self.tracers.values())

    @property
    def clock(self):
        c = L.Clock()
        check(lib().nh_clock_get(self._h, C.byref(c)))
        return c

    def set_clock(self, **values):
        """Overwrite members of the clock (time, iteration, stage, last_dt, last_stage_dt): what set!(model, filepath) does with
        the checkpointed clock (checkpointer.jl:251-256)."""
        c = self.clock
        for k, v in values.items():
            if k not in ("time", "iteration", "stage", "last_dt", "last_stage_dt"):
                raise ValueError(f"the clock has no member {k}")
            setattr(c, k, v)
        check(lib().nh_clock_set(self._h, C.byref(c)))

    def update_state(self, compute_tendencies=True):
        check(lib().nh_update_state(self._h, int(compute_tendencies)))

    def launch_count(self):
        n = C.c_int64(0)
        check(lib().nh_launch_count(self._h, C.byref(n)))
        return n.value

    def synchronize(self):
        check(lib().nh_synchronize(self._h))

    def max_abs_divergence(self):
        v = C.c_double(0)
        check(lib().nh_max_abs_divergence(self._h, C.byref(v)))
        return v.value


def _mask(fields):
    if isinstance(fields, Field):
        fields = [fields]
    m = 0
    for f in fields:
        if f.id >= 32:
            raise ValueError("tendency fields have no halo fill")
        m |= 1 << f.id
    return m


def fill_halo_regions(fields, fill_boundary_normal_velocities=True):
    fields = [fields] if isinstance(fields, Field) else list(fields)
    check(lib().nh_fill_halo_regions(fields[0].model._h, _mask(fields), int(fill_boundary_normal_velocities)))


def update_state(model, compute_tendencies=True):
    check(lib().nh_update_state(model._h, int(compute_tendencies)))


def compute_tendencies(model):
    check(lib().nh_compute_tendencies(model._h))


def calculate_pressure_correction(model, dt):
    check(lib().nh_calculate_pressure_correction(model._h, float(dt)))


def pressure_correct_velocities(model, dt):
    check(lib().nh_pressure_correct_velocities(model._h, float(dt)))


def rk3_substep(model, dt, gamma, zeta=None):
    check(lib().nh_rk3_substep(model._h, float(dt), float(gamma), 0.0 if zeta is None else float(zeta), int(zeta is not None)))


def ab2_step(model, dt, chi):
    check(lib().nh_ab2_step(model._h, float(dt), float(chi)))


def cache_previous_tendencies(model):
    check(lib().nh_cache_previous_tendencies(model._h))


def time_step(model, dt, euler=False, n_steps=1):
    """time_step!(model, Δt; euler=false); ``n_steps`` repeats it without returning to Python."""
    check(lib().nh_time_step(model._h, float(dt), int(euler), int(n_steps)))


def time_step_profiled(model, dt, euler=False):
    """One time_step! with per-phase CUDA-event times (ms): tendencies+substep, pressure solve, pressure correction, rest."""
    ms = (C.c_double * 4)()
    check(lib().nh_time_step_profiled(model._h, float(dt), int(euler), ms))
    return dict(zip(("tendencies_substep", "pressure_solve", "pressure_correct", "other"), (float(v) for v in ms)))


def set_model(model, enforce_incompressibility=True, **kwargs):
    """set!(model; u=…, v=…, w=…, T=…, …, enforce_incompressibility=true); set!(model, filepath) with a checkpoint file's path
    in second place restores the checkpointed state (checkpointer.jl:223-260)."""
    if isinstance(enforce_incompressibility, (str, os.PathLike)):
        from .checkpointer import set_from_checkpoint
        if kwargs:
            raise ValueError("set!(model, filepath) takes no field values")
        return set_from_checkpoint(model, enforce_incompressibility)
    names = ["u", "v", "w"] + list(model.tracer_names)
    for name, value in kwargs.items():
        if name not in names:
            raise ValueError(f"name {name} not found in model.velocities or model.tracers.")
        f = model.prognostic_fields()[names.index(name)]
        f.set(value)
        fill_halo_regions(f)
    check(lib().nh_set_state_finalize(model._h, int(enforce_incompressibility)))


# ---- standalone solvers ------------------------------------------------------------------------------------
class _PoissonSolver:
    kind = None

    def __init__(self, grid):
        self.grid = grid
        self._h = C.c_void_p()
        check(lib().nh_poisson_create(grid._h, self.kind, C.byref(self._h)))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().nh_poisson_destroy(self._h)
            self._h = None

    def solve(self, b, m=0):
        """solve!(ϕ, solver, b, m): real (Nx,Ny,Nz) right-hand side in, real solution out (host arrays)."""
        b = np.asfortranarray(np.asarray(b, dtype=self.grid.FT))
        if b.shape != self.grid.N:
            raise ValueError(f"rhs must have shape {self.grid.N}")
        phi = np.empty(self.grid.N, dtype=self.grid.FT, order="F")
        check(lib().nh_poisson_solve(self._h, _ptr(b), _ptr(phi), float(m), L.HOST))
        return phi

    def launch_count(self):
        n = C.c_int64(0)
        check(lib().nh_poisson_launch_count(self._h, C.byref(n)))
        return n.value


class FFTBasedPoissonSolver(_PoissonSolver):
    kind = L.SOLVER_FFT


class FourierTridiagonalPoissonSolver(_PoissonSolver):
    kind = L.SOLVER_FOURIER_TRIDIAGONAL


class BatchedTridiagonalSolver:
    def __init__(self, grid, lower_diagonal, diagonal, upper_diagonal, direction=2):
        self.grid = grid
        a = np.ascontiguousarray(lower_diagonal, dtype=np.float64)
        c = np.ascontiguousarray(upper_diagonal, dtype=np.float64)
        b = np.asarray(diagonal, dtype=np.float64)
        b3d = b.ndim == 3
        b = np.asfortranarray(b)
        dp = C.POINTER(C.c_double)
        self._h = C.c_void_p()
        check(lib().nh_tridiag_create(grid._h, a.ctypes.data_as(dp), b.ctypes.data_as(dp), c.ctypes.data_as(dp),
                                      int(b3d), int(direction), C.byref(self._h)))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().nh_tridiag_destroy(self._h)
            self._h = None

    def solve(self, f):
        f = np.asfortranarray(np.asarray(f, dtype=self.grid.FT))
        phi = np.empty(self.grid.N, dtype=self.grid.FT, order="F")
        check(lib().nh_tridiag_solve(self._h, _ptr(f), _ptr(phi)))
        return phi
