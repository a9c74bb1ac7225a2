"""ctypes binding of libnhstep.so (include/nhstep.h).  No CPU fallback: a missing library or device is an error."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("NHSTEP_LIB", os.path.join(_HERE, "libnhstep.so"))   # override: development A/B builds only

NH_MAX_TRACERS = 8
TRIDIAG_STATIC_MAX_ELEMENTS = 1 << 20
F32, F64 = 0, 1
PERIODIC, BOUNDED, FLAT = 0, 1, 2
ADV_NONE, ADV_CENTERED, ADV_UPWIND, ADV_WENO = 0, 1, 2, 3
RK3, AB2 = 0, 1
BUOY_NONE, BUOY_TRACER, BUOY_SEAWATER = 0, 1, 2
HOST, DEVICE = 0, 1
INTERIOR, PARENT = 0, 1
SOLVER_FFT, SOLVER_FOURIER_TRIDIAGONAL = 0, 1
FIELD_U, FIELD_V, FIELD_W, FIELD_TRACER0, FIELD_PNHS, FIELD_PHY, FIELD_GN0, FIELD_GM0 = 0, 1, 2, 3, 16, 17, 32, 48
FIELD_NUE, FIELD_KAPPAE0 = 18, 19
CLOSURE_NONE, CLOSURE_SCALAR, CLOSURE_SMAGORINSKY, CLOSURE_SMAGORINSKY_LILLY, CLOSURE_AMD = range(5)


class GridDesc(C.Structure):
    _fields_ = [("float_type", C.c_int32), ("N", C.c_int32 * 3), ("H", C.c_int32 * 3), ("topology", C.c_int32 * 3),
                ("x", C.c_double * 2), ("y", C.c_double * 2), ("z", C.c_double * 2), ("z_faces", C.POINTER(C.c_double))]


class ModelDesc(C.Structure):
    _fields_ = [("advection_kind", C.c_int32 * 3), ("advection_order", C.c_int32 * 3), ("timestepper", C.c_int32),
                ("ab2_chi", C.c_double), ("n_tracers", C.c_int32), ("buoyancy_kind", C.c_int32),
                ("buoyancy_tracer", C.c_int32 * 2), ("gravitational_acceleration", C.c_double),
                ("thermal_expansion", C.c_double), ("haline_contraction", C.c_double), ("has_coriolis", C.c_int32),
                ("coriolis_f", C.c_double), ("has_closure", C.c_int32), ("nu", C.c_double),
                ("kappa", C.c_double * NH_MAX_TRACERS), ("vertically_implicit", C.c_int32), ("smagorinsky_C", C.c_double), ("smagorinsky_Cb", C.c_double),
                ("smagorinsky_Pr", C.c_double * NH_MAX_TRACERS), ("amd_Cnu", C.c_double),
                ("amd_Ckappa", C.c_double * NH_MAX_TRACERS), ("hydrostatic_pressure_anomaly", C.c_int32),
                ("exact_arithmetic", C.c_int32)]


class Clock(C.Structure):
    _fields_ = [("time", C.c_double), ("iteration", C.c_int64), ("stage", C.c_int32), ("last_dt", C.c_double),
                ("last_stage_dt", C.c_double)]


# every symbol include/nhstep.h declares (tests check that the library exports all of them)
SYMBOLS = [
    "nh_last_error", "nh_version", "nh_device_count", "nh_grid_create", "nh_grid_destroy", "nh_grid_info",
    "nh_grid_spacings", "nh_model_create", "nh_model_set_weno_z_coefficients", "nh_model_destroy", "nh_model_grid_info", "nh_field_size", "nh_set_field",
    "nh_get_field", "nh_set_state_finalize", "nh_fill_halo_regions", "nh_update_state", "nh_compute_tendencies",
    "nh_rk3_substep", "nh_ab2_step", "nh_cache_previous_tendencies", "nh_calculate_pressure_correction",
    "nh_pressure_correct_velocities", "nh_time_step", "nh_time_step_profiled", "nh_clock_get", "nh_clock_set", "nh_max_abs_divergence",
    "nh_synchronize", "nh_field_device_ptr", "nh_model_set_stream", "nh_launch_count", "nh_poisson_create",
    "nh_poisson_destroy", "nh_poisson_solve", "nh_poisson_synchronize", "nh_poisson_launch_count",
    "nh_poisson_time_solves", "nh_tridiag_create", "nh_tridiag_solve", "nh_tridiag_destroy",
    "nh_dist_configure", "nh_dist_buffer", "nh_dist_halo", "nh_dist_tendency_substep", "nh_dist_tendency_substep_part", "nh_dist_poisson",
    "nh_dist_pressure_correct", "nh_dist_ipc_export", "nh_dist_ipc_import", "nh_model_set_boundary_condition",
    "nh_model_set_boundary_condition_array", "nh_dist_mode",
    "nh_dist_barrier", "nh_dist_transpose", "nh_tridiag_solve_static", "nh_reconstruction_coefficients",
]
BC_DEFAULT, BC_FLUX, BC_VALUE, BC_GRADIENT = range(4)
DIST_SEND_LO, DIST_SEND_HI, DIST_RECV_LO, DIST_RECV_HI, DIST_SPECTRAL_A, DIST_SPECTRAL_B = range(6)

_lib = None


class NhError(RuntimeError):
    pass


def lib():
    """Load libnhstep.so (built in-tree by __graft_entry__.build() / csrc/Makefile)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise NhError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(there is no CPU fallback)")
        _lib = C.CDLL(LIB_PATH)
        _lib.nh_last_error.restype = C.c_char_p
        p = C.c_void_p
        sig = {
            "nh_grid_create": [C.POINTER(GridDesc), C.c_int32, C.POINTER(p)],
            "nh_grid_destroy": [p],
            "nh_grid_info": [p, C.POINTER(GridDesc)],
            "nh_grid_spacings": [p, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double), C.c_int64],
            "nh_model_create": [p, C.POINTER(ModelDesc), C.POINTER(p)],
            "nh_model_destroy": [p],
            "nh_model_grid_info": [p, C.POINTER(GridDesc)],
            "nh_field_size": [p, C.c_int32, C.c_int32, C.POINTER(C.c_int64)],
            "nh_set_field": [p, C.c_int32, C.c_int32, p, C.c_int64, C.c_int32],
            "nh_get_field": [p, C.c_int32, C.c_int32, p, C.c_int64, C.c_int32],
            "nh_set_state_finalize": [p, C.c_int32],
            "nh_fill_halo_regions": [p, C.c_uint32, C.c_int32],
            "nh_model_set_boundary_condition": [p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_double],
            "nh_dist_mode": [p, C.POINTER(C.c_int32)],
            "nh_model_set_boundary_condition_array": [p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_int64,
                                                      C.c_int32],
            "nh_model_set_weno_z_coefficients": [p, C.c_int32, C.c_int32, C.c_void_p, C.c_int64, C.c_int32],
            "nh_update_state": [p, C.c_int32],
            "nh_compute_tendencies": [p],
            "nh_rk3_substep": [p, C.c_double, C.c_double, C.c_double, C.c_int32],
            "nh_ab2_step": [p, C.c_double, C.c_double],
            "nh_cache_previous_tendencies": [p],
            "nh_calculate_pressure_correction": [p, C.c_double],
            "nh_pressure_correct_velocities": [p, C.c_double],
            "nh_time_step": [p, C.c_double, C.c_int32, C.c_int32],
            "nh_time_step_profiled": [p, C.c_double, C.c_int32, C.POINTER(C.c_double)],
            "nh_clock_get": [p, C.POINTER(Clock)],
            "nh_clock_set": [p, C.POINTER(Clock)],
            "nh_max_abs_divergence": [p, C.POINTER(C.c_double)],
            "nh_synchronize": [p],
            "nh_field_device_ptr": [p, C.c_int32, C.POINTER(p), C.POINTER(C.c_int64), C.POINTER(C.c_int64)],
            "nh_model_set_stream": [p, p],
            "nh_launch_count": [p, C.POINTER(C.c_int64)],
            "nh_dist_configure": [p, C.c_int32, C.c_int32, C.c_double, C.c_double, C.c_int32],
            "nh_dist_buffer": [p, C.c_int32, C.POINTER(p), C.POINTER(C.c_int64)],
            "nh_dist_halo": [p, C.c_uint32, C.c_int32, C.c_int32, C.POINTER(C.c_int64)],
            "nh_dist_tendency_substep": [p, C.c_double, C.c_int32, C.c_int32, C.POINTER(C.c_double)],
            "nh_dist_tendency_substep_part": [p, C.c_double, C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_double)],
            "nh_dist_poisson": [p, C.c_int32, C.c_double],
            "nh_dist_pressure_correct": [p, C.c_double],
            "nh_dist_ipc_export": [p, p, C.c_int64, C.POINTER(C.c_int64)],
            "nh_dist_ipc_import": [p, p, C.c_int64],
            "nh_dist_barrier": [p],
            "nh_dist_transpose": [p, C.c_int32],
            "nh_poisson_create": [p, C.c_int32, C.POINTER(p)],
            "nh_poisson_destroy": [p],
            "nh_poisson_solve": [p, p, p, C.c_double, C.c_int32],
            "nh_poisson_synchronize": [p],
            "nh_poisson_launch_count": [p, C.POINTER(C.c_int64)],
            "nh_poisson_time_solves": [p, p, p, C.c_int32, C.POINTER(C.c_float)],
            "nh_tridiag_create": [p, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double), C.c_int32,
                                  C.c_int32, C.POINTER(p)],
            "nh_tridiag_solve": [p, p, p],
            "nh_tridiag_destroy": [p],
            "nh_reconstruction_coefficients": [C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_double), C.POINTER(C.c_int32)],
            "nh_tridiag_solve_static": [C.c_int32, C.c_int32, C.POINTER(C.c_int32), C.c_int32, p, p, C.c_int32, p, p, p, p],
            "nh_device_count": [C.POINTER(C.c_int32)],
        }
        for name, args in sig.items():
            fn = getattr(_lib, name)
            fn.argtypes = args
            fn.restype = C.c_int32
    return _lib


def check(status):
    if status != 0:
        raise NhError(lib().nh_last_error().decode("utf-8", "replace"))


def device_count():
    n = C.c_int32(0)
    check(lib().nh_device_count(C.byref(n)))
    return n.value
